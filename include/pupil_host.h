/* pupil_host.h — C entry points of libpupil_host.so, the C++ HOST layer above libpupil_b200.so:
 * mitsuba3-style scene XML loading (Pupil::resource::Scene::LoadFromXML,
 * framework/resource/scene.cpp:27), the World::LoadScene hand-off (framework/world/world.cpp:101) and
 * a headless run of the System / Pass loop with B200PTPass (framework/system/system.cpp:81-101,
 * example/path_tracer/pt_pass.cpp:39-57).  The C++ classes themselves are in
 * pupiloptixlab_b200/host/{system,b200_pt_pass}.h; this header is what tests / tools bind.        */
#ifndef PUPIL_HOST_H
#define PUPIL_HOST_H

#include <stdint.h>

#include "pupil_b200.h"

#ifdef __cplusplus
extern "C" {
#endif

#ifndef PH_API
#define PH_API __attribute__((visibility("default")))
#endif

/* A scene consumer: the pb_* scene-ingestion entry points of libpupil_b200.so (ph_default_sink), or
 * any other implementation of the same signatures.                                                 */
typedef struct ph_sink {
    int (*scene_begin)(void* ctx);
    int (*add_material)(void* ctx, const pb_material* m, uint32_t* out_id);
    int (*add_trimesh)(void* ctx, const float* pos, const float* nrm, const float* uv, uint32_t n_vert,
                       const uint32_t* idx, uint32_t n_tri, const float to_world[12], uint32_t flags,
                       uint32_t material, const pb_texture* area_radiance, uint32_t* out_object);
    int (*add_sphere)(void* ctx, const float to_world[12], uint32_t flags, uint32_t material,
                      const pb_texture* area_radiance, uint32_t* out_object);
    int (*set_env)(void* ctx, const pb_env* env);
    int (*set_camera)(void* ctx, const float sample_to_camera[16], const float camera_to_world[16]);
    int (*scene_commit)(void* ctx);
} ph_sink;

typedef struct ph_scene ph_scene;

typedef struct ph_scene_info {
    int32_t  max_depth;            /* integrator.max_depth (default 1, scene.cpp:79-82)            */
    int32_t  film_w, film_h;       /* hdrfilm width / height (768 x 576, scene.cpp:86-95)          */
    float    fov_y;                /* vertical fov in degrees after the fov_axis conversion         */
    float    near_clip, far_clip;
    uint32_t n_shapes, n_emissive_shapes, n_triangles, n_spheres;
    uint32_t has_env;              /* PB_ENV_*                                                      */
    uint32_t n_warnings;
} ph_scene_info;

PH_API int  ph_scene_load_xml(const char* path, ph_scene** out);                        /* Scene::LoadFromXML(file) */
PH_API int  ph_scene_load_xml_string(const char* xml, const char* root_dir, ph_scene** out);
PH_API void ph_scene_free(ph_scene* s);
PH_API const char* ph_last_error(void);
PH_API int  ph_scene_info_get(const ph_scene* s, ph_scene_info* out);
PH_API const char* ph_scene_warning(const ph_scene* s, uint32_t i);
/* optix::Camera matrices (framework/render/camera.h:7-10) as CameraHelper::GetCudaMemory uploads them */
PH_API int  ph_scene_camera(const ph_scene* s, float sample_to_camera[16], float camera_to_world[16]);
/* shape i: type (0 obj, 1 sphere, 2 cube, 3 rectangle), rows 0..2 of its to_world, material, emitter */
PH_API int  ph_scene_shape(const ph_scene* s, uint32_t i, uint32_t* type, float to_world[12], pb_material* mat,
                           uint32_t* flags, uint32_t* is_emitter, pb_texture* radiance);
/* World::LoadScene: feed the scene to a sink (NULL = the linked libpupil_b200.so), ending in commit */
PH_API int  ph_scene_upload(const ph_scene* s, const ph_sink* sink, void* ctx);
PH_API const ph_sink* ph_default_sink(void);

/* Headless System::Run with one B200PTPass: load `xml_path`, render `frames` frames (1 frame = 1 spp,
 * pt_pass.cpp:51-56) on CUDA device `device`, read back the "final result" buffer (float4 w*h,
 * running mean) and optionally albedo / normal (float3 w*h).  Buffers are host pointers or NULL.   */
PH_API int  ph_run_path_tracer(const char* xml_path, int device, uint32_t frames, float* final_rgba,
                               float* albedo_rgb, float* normal_rgb, uint32_t* out_w, uint32_t* out_h);

/* The same loop as an object, for callers that interleave rendering with scene edits (the reference's GUI thread does
 * this through the event system): System + one B200PTPass kept alive between calls.
 *   ph_session_run          n more frames (System::Run's loop body, system.cpp:96-101)
 *   ph_session_move_object  RenderObject::UpdateTransform (render_object.cpp:46-49) of shape instance `index`: raises
 *                           EWorldEvent::RenderInstanceTransform -> World updates the instance -> RenderInstanceUpdate ->
 *                           the pass restarts accumulation against the moved geometry (pt_pass.cpp:40-49,216-218)
 *   ph_session_read         copy a named BufferManager buffer ("final result", "pt accum buffer", "albedo", "normal",
 *                           "test") to host memory; bytes must equal the buffer size
 *   ph_session_sample_count frames accumulated since the last restart (sample_cnt, pt_pass.cpp:43,55)           */
typedef struct ph_session ph_session;
PH_API int  ph_session_create(const char* xml_path, int device, ph_session** out);
/* n GPUs of this box behind one pass: scene + BVH replicated, frames of a run dealt round-robin, one NCCL reduce into
 * devices[0]'s buffers (pb_create_multi / pb_render_multi).  ph_session_set_frames_per_run: frames per System frame
 * (B200PTPass::frames_per_run; >= n keeps every GPU busy).                                                         */
PH_API int  ph_session_create_multi(const char* xml_path, const int* devices, int n, ph_session** out);
PH_API int  ph_session_set_frames_per_run(ph_session* s, uint32_t frames);
PH_API void ph_session_destroy(ph_session* s);
PH_API int  ph_session_run(ph_session* s, uint32_t frames);
PH_API int  ph_session_move_object(ph_session* s, uint32_t index, const float to_world[16]);
PH_API int  ph_session_read(ph_session* s, const char* buffer_name, void* dst, uint64_t bytes);
PH_API int  ph_session_film(const ph_session* s, uint32_t* w, uint32_t* h);
PH_API uint32_t ph_session_sample_count(const ph_session* s);

/* util::BitmapTexture::Load / Save (framework/util/texture.cpp:87-180).  Load: PNG / BMP / TGA / PPM through the
 * reference's pow(x/255, 2.2) decode, Radiance .hdr / OpenEXR (scanline; none / zips / zip) / PFM as float data;
 * returns RGBA float texels (row 0 = top) in a buffer released by ph_image_free.  Save: by extension .hdr / .exr
 * (rows flipped and channel order as the reference writes them) / .pfm.  ph_image_save_png8 writes pb_tonemap's output. */
PH_API int  ph_image_load(const char* path, float** rgba, uint32_t* w, uint32_t* h);
PH_API void ph_image_free(float* rgba);
PH_API int  ph_image_save(const char* path, const float* rgba, uint32_t w, uint32_t h);
PH_API int  ph_image_save_png8(const char* path, const uint8_t* rgba8, uint32_t w, uint32_t h);

#ifdef __cplusplus
}
#endif
#endif /* PUPIL_HOST_H */
