/* pupil_b200.h — C ABI of libpupil_b200.so, the B200-native wavefront path tracer that replaces
 * the OptiX pipeline of vymv/PupilOptixLab's example/path_tracer.
 *
 * Every entry point cites the reference interface it stands in for (paths are into the reference
 * tree).  Plain C: opaque context, POD structs, pointers + sizes, int status codes.  No torch, no
 * C++ types.  Calls on one context must be serialised by the caller (the reference serialises
 * Pass::OnRun and SetScene under m_render_system_mutex, framework/system/system.cpp:96-101,142).
 *
 * The same header is implemented twice:
 *   - PB_API symbols  pb_*   : the product, CUDA sm_100a (pupiloptixlab_b200/csrc)
 *   - oracle symbols  orc_*  : the CPU oracle (oracle/), generated from the same prototypes with
 *                              the prefix swapped — test infrastructure only.
 */
#ifndef PUPIL_B200_H
#define PUPIL_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#ifndef PB_API
#define PB_API __attribute__((visibility("default")))
#endif

/* ---- status codes ---------------------------------------------------------------------------- */
enum {
    PB_OK = 0,
    PB_ERR_INVALID = 1,   /* bad argument / call order                         */
    PB_ERR_CUDA = 2,      /* a CUDA runtime call failed (see pb_last_error)    */
    PB_ERR_NO_DEVICE = 3, /* no CUDA device: there is NO CPU fallback          */
    PB_ERR_LIMIT = 4,     /* a structural limit was exceeded (BVH depth, ...)  */
    PB_ERR_NOT_COMMITTED = 5,
    PB_ERR_NCCL = 6       /* NCCL missing, a collective failed, or an asynchronous communicator error was polled */
};

/* ---- textures: mirror of Pupil::util::Texture / cuda::Texture -------------------------------
 * framework/util/texture.h:22-26 (ETextureType), framework/cuda/texture.h:9-57.                 */
enum { PB_TEX_RGB = 0, PB_TEX_BITMAP = 1, PB_TEX_CHECKERBOARD = 2 };

typedef struct pb_texture {
    uint32_t type;       /* PB_TEX_*                                                           */
    float a[3];          /* rgb  | checkerboard patch1                                          */
    float b[3];          /* ---  | checkerboard patch2                                          */
    float uv_r0[4];      /* rows r0, r1 of the 4x4 uv transform (cuda/texture.h:34-36)          */
    float uv_r1[4];
    /* bitmap (row-major RGBA float, host pointer, copied at pb_add_material / pb_add_* time)   */
    const float* bitmap;
    uint32_t bitmap_w, bitmap_h;
    uint32_t address_mode; /* 0 wrap, 1 clamp, 2 mirror, 3 border (util/texture.h)              */
    uint32_t filter_mode;  /* 0 point, 1 linear                                                 */
} pb_texture;

/* ---- materials: mirror of Pupil::resource::Material (framework/resource/material.h:20-84) ----
 * type values follow EMatType (framework/render/material/predefine.h:15-22 with
 * framework/decl/material_decl.inl:4-11).                                                       */
enum {
    PB_MAT_DIFFUSE = 1,
    PB_MAT_DIELECTRIC = 2,
    PB_MAT_ROUGH_DIELECTRIC = 3,
    PB_MAT_CONDUCTOR = 4,
    PB_MAT_ROUGH_CONDUCTOR = 5,
    PB_MAT_PLASTIC = 6,
    PB_MAT_ROUGH_PLASTIC = 7
};

typedef struct pb_material {
    uint32_t type;     /* PB_MAT_*                                                              */
    uint32_t twosided; /* resource::Material::twosided                                          */
    float int_ior, ext_ior;
    uint32_t nonlinear;
    uint32_t _pad;
    pb_texture reflectance;            /* diffuse.reflectance | plastic.diffuse_reflectance      */
    pb_texture specular_reflectance;
    pb_texture specular_transmittance;
    pb_texture alpha;
    pb_texture eta;                    /* conductors                                             */
    pb_texture k;
} pb_material;

/* ---- shape flags (resource::ShapeInstance, framework/resource/shape.h:60-72) ----------------- */
enum { PB_SHAPE_FLIP_NORMALS = 1u, PB_SHAPE_FLIP_TEX_COORDS = 2u };

/* ---- environment emitter (framework/resource/emitter.h, framework/world/emitter.cpp:280-313) - */
enum { PB_ENV_NONE = 0, PB_ENV_CONSTANT = 1, PB_ENV_MAP = 2 };
typedef struct pb_env {
    uint32_t type;
    float radiance[3];    /* constant                                                            */
    float scale;          /* envmap                                                              */
    float to_world[12];   /* envmap, rows of the 3x4                                             */
    pb_texture map;       /* envmap bitmap                                                       */
} pb_env;

/* ---- rays / hits for the ray-cast entry points ------------------------------------------------ */
typedef struct pb_ray { float ox, oy, oz, tmin, dx, dy, dz, tmax; } pb_ray;   /* 32 B */
/* prim = global primitive id (insertion order: triangles of each mesh in index order, one per
 * sphere); 0xFFFFFFFF = miss.  u,v = barycentric weights of vertices 1 and 2 (the layout of
 * optixGetTriangleBarycentrics, framework/render/geometry.h:52), 0 for spheres.                  */
typedef struct pb_hit { float t, u, v; uint32_t prim; } pb_hit;               /* 16 B */
#define PB_MISS 0xFFFFFFFFu

/* ---- render parameters: Pupil::pt::OptixLaunchParams::config + seeds (path_tracer/type.h:9-33) */
typedef struct pb_render_params {
    uint32_t width, height;
    uint32_t max_depth;     /* config.max_depth, read as the reference's int (type.h:12): a path scatters at
                               most max_depth - 1 times after the primary hit (main.cu:104-106), so values
                               <= 1 (mitsuba's -1 included) render the primary hit only; > 1024: PB_ERR_LIMIT */
    uint32_t first_sample;  /* random_seed of the first frame rendered by this call              */
    uint32_t sample_count;  /* number of frames (1 frame = 1 spp, pt_pass.cpp:51-56)             */
    uint32_t sample_stride; /* frame index step (multi-GPU sample interleave); 0 means 1         */
    uint32_t flags;         /* PB_RENDER_*                                                       */
    uint32_t _pad;
} pb_render_params;
enum {
    PB_RENDER_HOST_BUFFERS = 1u, /* output pointers are host memory (copy included in the call)  */
    PB_RENDER_CLEAR = 2u,        /* zero the sum buffer first (sample_cnt = 0, pt_pass.cpp:43)   */
    PB_RENDER_COUNTED = 4u,      /* instrumented traversal: fills the visit counters of pb_stats
                                    (same image; never the timed build)                          */
    PB_RENDER_STAGE_TIMING = 8u, /* CUDA events around every extend / connect launch: fills
                                    extend_ms / connect_ms / *_launches of pb_stats               */
    PB_RENDER_REDUCE = 32u,      /* multi-GPU (a communicator is attached, pb_comm_init_rank / pb_create_multi): this rank's
                                    frames are summed into a context-owned film, ONE ncclReduce(sum, fp32, root 0) of
                                    w*h*4 floats is enqueued on the same stream behind the last wave, and the ROOT adds
                                    the reduced film to its accum_sum_rgba (PB_RENDER_CLEAR: replaces it).  Collective:
                                    every rank calls pb_render with the same film size; the accumulation buffers of the
                                    other ranks are not written.  The reference is single-GPU (cuda/context.cpp:15-44);
                                    the sharding is SURVEY 8(e): rank r renders frames first_sample + k * sample_stride
                                    with first_sample = base + r, sample_stride = n_ranks                              */
    PB_RENDER_ASYNC = 16u        /* return once the work is enqueued on the stream (the reference's
                                    optixLaunch is asynchronous too, pass.h:79-86); pb_wait /
                                    pb_get_stats synchronise and report the totals of all renders
                                    enqueued since the last wait.  With PB_RENDER_HOST_BUFFERS the
                                    uploads / downloads are enqueued as well: the host buffers must stay
                                    valid until pb_wait and should be pinned (pageable memory makes the
                                    copies block the caller)                                        */
};

typedef struct pb_stats {
    uint64_t rays_closest, rays_any;        /* rays issued by the last render / trace call       */
    uint64_t samples;                       /* pixel-samples finished                            */
    uint64_t node_visits, prim_tests_tri, prim_tests_sphere; /* only from *_counted calls        */
    uint64_t n_prims, n_nodes, n_emitters;
    uint32_t bvh_depth, n_objects;
    float    build_ms, last_render_ms, last_trace_ms;
    uint32_t kernel_launches;               /* product kernels launched by the last call         */
    /* PB_RENDER_COUNTED: visits of the shadow (any-hit) rays; node_visits / prim_tests_* above
     * then hold the closest-hit rays' visits only                                                */
    uint64_t node_visits_any, prim_tests_tri_any, prim_tests_sphere_any;
    /* PB_RENDER_STAGE_TIMING: summed device time of the traversal launches of the last render    */
    float    extend_ms, connect_ms;
    uint32_t extend_launches, connect_launches;
    float    shade_ms;                      /* miss + the seven shade launches of every bounce   */
    uint32_t shade_groups;
    /* how pb_update_instance was served since the last pb_scene_commit: refits in place (IAS::Update semantics) and
     * full rebuilds (the move changed the scene box enough to change the contract's padding delta)               */
    uint32_t refits, rebuilds;
} pb_stats;

typedef struct pb_ctx pb_ctx;

/* ---- lifetime.  Replaces cuda::Context::Init + optix::Context::Init + PTPass ctor
 * (framework/cuda/context.cpp:9-49, framework/optix/context.cpp:38-50, path_tracer/pt_pass.cpp:29) */
PB_API int pb_create(pb_ctx** out, int device);
PB_API void pb_destroy(pb_ctx* ctx);
PB_API const char* pb_last_error(const pb_ctx* ctx);
PB_API const char* pb_version(void);

/* ---- scene ingestion: host pointers, copied during the call.  Replaces World::LoadScene →
 * RenderObject ctor + Material::LoadMaterial + EmitterHelper::AddAreaEmitter + GAS::Create
 * (framework/world/world.cpp:101-139, world/render_object.cpp:10-39,
 *  render/material/optix_material.cpp:121-132, world/emitter.cpp:245-278, world/gas_manager.cpp:61). */
PB_API int pb_scene_begin(pb_ctx* ctx);
PB_API int pb_add_material(pb_ctx* ctx, const pb_material* m, uint32_t* out_id);
/* pos: 3*n_vert, nrm: 3*n_vert or NULL, uv: 2*n_vert or NULL, idx: 3*n_tri; to_world = rows 0..2
 * of the row-major Mat4 (framework/util/type.h:73).  area_radiance != NULL makes it an emitter.  */
PB_API int pb_add_trimesh(pb_ctx* ctx, const float* pos, const float* nrm, const float* uv,
                          uint32_t n_vert, const uint32_t* idx, uint32_t n_tri,
                          const float to_world[12], uint32_t flags, uint32_t material,
                          const pb_texture* area_radiance, uint32_t* out_object);
/* unit sphere at the origin under to_world (shape.cpp:106-125 folds radius/center into it)       */
PB_API int pb_add_sphere(pb_ctx* ctx, const float to_world[12], uint32_t flags, uint32_t material,
                         const pb_texture* area_radiance, uint32_t* out_object);
PB_API int pb_set_env(pb_ctx* ctx, const pb_env* env);
/* optix::Camera (framework/render/camera.h:7-10): two row-major 4x4                              */
PB_API int pb_set_camera(pb_ctx* ctx, const float sample_to_camera[16], const float camera_to_world[16]);
/* emitter tables (EmitterHelper::ComputeProbability, world/emitter.cpp:321-337) + device BVH
 * build (replaces optixAccelBuild, world/gas_manager.cpp:151, world/ias_manager.cpp:84)          */
PB_API int pb_scene_commit(pb_ctx* ctx);
/* IAS::Update analogue (world/ias_manager.cpp:116-151,187-205; RenderObject::UpdateTransform, world/render_object.cpp:46-49):
 * new object->world transform of one object.  Served by a REFIT: the object's world-space leaf records are rewritten in
 * place and every node's child boxes are recomputed bottom-up (no sort, no collapse, no allocation; device pointers
 * and captured wave graphs stay valid; < 1 ms at 10 M primitives); area-emitter tables of a moved light are rewritten
 * in place.  Falls back to a full rebuild only when the move changes the scene's largest extent, i.e. delta, the
 * padding of the intersection contract (pb_stats.refits / .rebuilds tell which).  Results never depend on the tree.  */
PB_API int pb_update_instance(pb_ctx* ctx, uint32_t object, const float to_world[12]);

/* ---- ray casting.  Replaces optixTrace ray type 0 (path_tracer/main.cu:77,158) and ray type 1
 * with TERMINATE_ON_FIRST_HIT (framework/render/emitter.h:91-100).  rays/hits are DEVICE pointers
 * unless host != 0 (then the call stages them through the device, copies included).              */
PB_API int pb_trace_closest(pb_ctx* ctx, const pb_ray* rays, uint64_t n, pb_hit* hits, int host, void* cuda_stream);
PB_API int pb_trace_any(pb_ctx* ctx, const pb_ray* rays, uint64_t n, uint8_t* occluded, int host, void* cuda_stream);
/* same result, instrumented build: fills node/prim visit counters in pb_stats (never timed)      */
PB_API int pb_trace_closest_counted(pb_ctx* ctx, const pb_ray* rays, uint64_t n, pb_hit* hits, int host, void* cuda_stream);
PB_API int pb_trace_any_counted(pb_ctx* ctx, const pb_ray* rays, uint64_t n, uint8_t* occluded, int host, void* cuda_stream);

/* ---- allocation-free closest hit.  NO COUNTERPART IN THE REFERENCE: added at the request of this run's operators for
 * callers that cannot call cudaMalloc (tight serving loops, harnesses that forbid it).  Same result as pb_trace_closest,
 * bit for bit; it allocates nothing (no cudaMalloc / cudaFree / event creation): the kernel's scratch — ray cursor and
 * counters, and for host != 0 the device staging of rays and hits — is a file-scope device array of
 * 256 B + 48 B x PB_STATIC_TRACE_MAX_RAYS (12.0 MiB) that every GPU loading the library holds until the process exits.
 * n > PB_STATIC_TRACE_MAX_RAYS is refused with PB_ERR_LIMIT.  Synchronous.  All calls on a device share that array:
 * calls that overlap on the GPU (other streams / threads) race — serialise them.                                       */
#define PB_STATIC_TRACE_MAX_RAYS 262144u
PB_API int pb_trace_closest_static(pb_ctx* ctx, const pb_ray* rays, uint64_t n, pb_hit* hits, int host, void* cuda_stream);

/* ---- render.  Replaces PTPass::OnRun → optix::Pass::Run → optixLaunch(w,h,1) repeated
 * sample_count times (path_tracer/pt_pass.cpp:39-57, framework/optix/pass.h:69-87).
 * accum_sum_rgba: w*h float4, receives SUMS of per-frame radiance in .xyz and the number of frames
 * added in .w (the reference's running mean main.cu:187-193 is sum/count); albedo/normal: w*h*3
 * floats of the last frame (main.cu:94-99) or NULL.                                              */
PB_API int pb_render(pb_ctx* ctx, const pb_render_params* p, float* accum_sum_rgba, float* albedo_rgb,
                     float* normal_rgb, void* cuda_stream);
/* pb_render with every output buffer of the pass contract (pt_pass.cpp:123-147): "pt accum buffer", "albedo",
 * "normal" and "test" (main.cu:101: the RNG draw that follows the primary hit, one float per pixel, last frame).
 * accum_sum_rgba is required, the others may be NULL.  Device pointers unless PB_RENDER_HOST_BUFFERS.          */
typedef struct pb_render_targets {
    float* accum_sum_rgba;   /* w*h float4 */
    float* albedo_rgb;       /* w*h*3      */
    float* normal_rgb;       /* w*h*3      */
    float* test;             /* w*h        */
} pb_render_targets;
PB_API int pb_render_ex(pb_ctx* ctx, const pb_render_params* p, const pb_render_targets* targets, void* cuda_stream);
/* "frame buffer" of the reference: mean = sum.xyz / sum.w, alpha 1 (main.cu:192-193)             */
PB_API int pb_resolve(pb_ctx* ctx, const float* accum_sum_rgba, float* frame_rgba, uint32_t n_pixels,
                      int host, void* cuda_stream);

/* display transform of the reference's output pass (framework/system/gui/output.hlsl:30-73): optional ACES
 * tone mapping (adapted luminance 1), optional gamma 2.2, then the UNORM8 conversion the swap chain's render
 * target performs (saturate, x * 255 + 0.5).  frame_rgba = n_pixels float4 (pb_resolve's output), rgba8 =
 * n_pixels * 4 bytes in the same pixel order (the shader's vertical flip belongs to the window, not the data). */
PB_API int pb_tonemap(pb_ctx* ctx, const float* frame_rgba, uint8_t* rgba8, uint32_t n_pixels, int tone_mapping,
                      int gamma_correct, int host, void* cuda_stream);

/* waits for renders enqueued with PB_RENDER_ASYNC (replaces Pass::Synchronize, framework/optix/pass.h:88) */
PB_API int pb_wait(pb_ctx* ctx);
/* implies pb_wait */
PB_API int pb_get_stats(pb_ctx* ctx, pb_stats* out);

/* ---- multi-GPU: one context per GPU, scene + BVH replicated (every context gets the same pb_add_* / pb_scene_commit
 * calls and builds the identical tree: deterministic build, cheaper than broadcasting it), frames sharded by sample
 * index, one NCCL reduce of the fp32 sum film per pb_render.  NCCL is loaded at run time (libnccl.so.2, the
 * copy already in the process if there is one); without it these calls return PB_ERR_NCCL and everything else works.
 *
 *  one process per GPU (bench.py under torchrun, MPI, ...):
 *      rank 0: pb_comm_unique_id(id) -> ship the 128 bytes to every rank by any means
 *      all   : pb_comm_init_rank(ctx, id, n_ranks, rank)                 (collective)
 *      all   : pb_render(ctx, {first_sample = base + rank, sample_stride = n_ranks, flags |= PB_RENDER_REDUCE}, ...)
 *  one process, n GPUs (host/b200_pt_pass.cpp):
 *      pb_create_multi(ctxs, devices, n)  -> n contexts, rank i on devices[i], communicators from ncclCommInitAll
 *      pb_render_multi(ctxs, n, p, targets_on_ctx0, streams)             p describes the WHOLE job: frames
 *                                          first_sample + j * stride, j < sample_count, dealt round-robin to the ranks
 * Asynchronous NCCL errors (a peer died, a link failed) are polled in pb_wait / pb_get_stats / synchronous pb_render
 * and reported as PB_ERR_NCCL with the NCCL error string in pb_last_error.                                          */
#define PB_UNIQUE_ID_BYTES 128
PB_API int pb_comm_unique_id(unsigned char id[PB_UNIQUE_ID_BYTES]);
PB_API int pb_comm_init_rank(pb_ctx* ctx, const unsigned char id[PB_UNIQUE_ID_BYTES], int n_ranks, int rank);
PB_API int pb_comm_info(const pb_ctx* ctx, int* n_ranks, int* rank);      /* 1, 0 when no communicator is attached */
PB_API int pb_comm_destroy(pb_ctx* ctx);                                  /* also done by pb_destroy             */
PB_API int pb_create_multi(pb_ctx** out, const int* devices, int n);
PB_API int pb_render_multi(pb_ctx* const* ctxs, int n, const pb_render_params* p, const pb_render_targets* targets_on_ctx0,
                           void* const* cuda_streams /* n entries or NULL */);

/* ---- BVH export (test / roofline accounting only): the committed 8-wide BVH and its primitive
 * records, host buffers.  Query sizes by passing NULL pointers.                                  */
PB_API int pb_export_bvh(pb_ctx* ctx, void* nodes80, uint64_t* n_nodes, void* prims48, uint64_t* n_prims);

#ifdef __cplusplus
}
#endif
#endif /* PUPIL_B200_H */
