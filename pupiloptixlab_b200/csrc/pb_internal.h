// pb_internal.h — the context behind pb_ctx and declarations shared by the translation units.
#pragma once
#include <cuda_runtime.h>

#include <string>
#include <vector>

#include "../../include/pupil_b200.h"
#include "pb_types.h"
#include "pb_traverse.cuh"

namespace pb {

struct HostObject {
    bool sphere = false;
    float M[12]{};
    uint32_t flags = 0, material = 0;
    bool is_emitter = false;
    pb_texture area_radiance{};
    uint32_t vert_offset = 0, n_vert = 0, index_offset = 0, n_tri = 0;
    bool has_nrm = false, has_uv = false;
};

struct Wavefront;  // pb_wavefront.cu

}  // namespace pb

struct pb_ctx {
    int device = 0;
    std::string err;
    cudaStream_t stream = nullptr;   // internal stream used when the caller passes none
    bool in_scene = false, committed = false;

    // ---- host copy of the scene (kept for pb_update_instance)
    std::vector<pb_material> materials;
    std::vector<pb::HostObject> objects;
    std::vector<float> h_pos, h_nrm, h_uv;
    std::vector<uint32_t> h_idx;
    std::vector<std::vector<float>> bitmaps;   // owned copies of bitmap texels
    pb_env env{};
    pb::Camera cam{};

    // ---- device scene
    float* d_pos = nullptr; float* d_nrm = nullptr; float* d_uv = nullptr; uint32_t* d_idx = nullptr;
    uint64_t scene_epoch = 1;            // bumped by commit / set_camera: captured wave graphs of older epochs are dropped
    bool use_graphs = true;              // PB_NO_GRAPH=1 launches kernel by kernel
    bool wave_target_env = false;        // wave_target came from PB_WAVE_PATHS
    uint32_t wave_target = 0;            // paths per wave, decided once per context (cudaMemGetInfo is too slow for every pb_render)
    float* d_stage = nullptr;            // film staging of the host-buffer pb_render path (sum 16 B + albedo 12 B + normal 12 B per pixel), grow-only
    size_t stage_px = 0;
    float* d_trishade = nullptr;         // 24 floats per triangle (pb_scene.cuh)
    pb::ObjRec* d_objs = nullptr;
    pb::MatRec* d_mats = nullptr;
    pb::EmitRec* d_emit = nullptr;       // area emitters, then (optionally) the env emitter
    float* d_emit_prefix = nullptr;      // running sums of select_probability (emitter.h:110-121)
    float* d_env_tables = nullptr;
    uint32_t n_area_emitters = 0;
    pb::EmitRec h_env_rec{};             // the environment record of the last commit (table pointers included), reused by refits
    bool has_env = false;
    std::vector<pb::ObjRec> h_objs;
    std::vector<float*> d_bitmaps;       // device copies of `bitmaps` (same order), RGBA float, sampled in software (pb_shade.cuh)

    // ---- BVH
    pb::Node8* d_nodes = nullptr;
    pb::PrimRec* d_prims = nullptr;      // leaf order
    uint32_t n_nodes = 0, n_prims = 0, bvh_levels = 0;
    float delta = 0.f;
    pb::V3 scene_lo{}, scene_hi{};
    // ---- refit state (pb_update_instance without a rebuild): BFS level boundaries of d_nodes, one exact (lo, hi) box
    // per node, the exact world box of every object (the scene box = their union decides delta)
    std::vector<uint32_t> level_first;   // level l = nodes [level_first[l], level_first[l + 1])
    pb::V3* d_node_lo = nullptr; pb::V3* d_node_hi = nullptr;
    std::vector<pb::V3> h_obj_lo, h_obj_hi;
    uint32_t* d_obj_box = nullptr;       // 6 ordered uints per object (scratch of the box reductions)
    bool allow_refit = true;             // PB_NO_REFIT=1: every pb_update_instance rebuilds (A/B, tests)
    uint32_t refits = 0, rebuilds = 0;   // how pb_update_instance was served since the last commit

    // ---- scratch
    unsigned long long* d_counters = nullptr;   // [0] ray fetch cursor, [1..3] visit counters
    pb::Wavefront* wf = nullptr;
    int sm_count = 148;
    uint32_t leaf_max = 2;               // primitives per leaf slot (PB_LEAF_MAX, 1..3)
    uint32_t bottom_max = 8;             // BVH collapse: subtrees up to this size become one wide node (PB_BOTTOM_MAX)
    uint32_t prim_iters = 2u;            // primitives a lane tests per primitive phase (PB_PRIM_ITERS).  Round 2, C4, same box:
                                         // unlimited 520.6, 1: 521.4, 2: 529.2 Msamples/s (round 1's optimum was unlimited: the
                                         // sphere records were dependent fetches then)
    uint32_t prim_batch = 8;             // lanes holding primitives before a warp runs its primitive tests (PB_PRIM_BATCH)
    uint32_t refill = 20;                // dynamic-fetch threshold (PB_REFILL overrides, tuning only)
    pb_stats stats{};

    // ---- multi-GPU (pb_comm.cu): NCCL communicator of this context, film the reduce works on
    void* nccl_comm = nullptr;           // ncclComm_t
    int rank = 0, n_ranks = 1;
    float* d_reduce = nullptr;           // w*h float4: this rank's frames of one PB_RENDER_REDUCE call, reduced in place
    size_t reduce_px = 0;

    // the any-hit kernel (k_connect) ends a ray at its first accepted hit, so its batching knobs are tuned separately
    // (PB_PRIM_BATCH_ANY / PB_PRIM_ITERS_ANY / PB_REFILL_ANY; 0 = the closest-hit value)
    uint32_t prim_iters_any = 0, prim_batch_any = 0, refill_any = 0;
    pb::TraceScene trace_scene_any() const {
        pb::TraceScene t = trace_scene();
        if (refill_any) t.refill = refill_any;
        if (prim_batch_any) t.prim_batch = prim_batch_any;
        if (prim_iters_any) t.prim_iters = prim_iters_any;
        return t;
    }
    pb::TraceScene trace_scene() const {
        return pb::TraceScene{reinterpret_cast<const pb::U4*>(d_nodes), reinterpret_cast<const pb::V4*>(d_prims), d_objs, delta, n_nodes, refill, prim_batch, 0x47000000u, prim_iters};
    }
};

#define PB_CUDA_TRY(ctx, call)                                                                       \
    do {                                                                                             \
        cudaError_t _e = (call);                                                                     \
        if (_e != cudaSuccess) {                                                                     \
            (ctx)->err = std::string(#call) + ": " + cudaGetErrorString(_e);                         \
            return PB_ERR_CUDA;                                                                      \
        }                                                                                            \
    } while (0)

namespace pb {
// pb_build.cu
int build_bvh(pb_ctx* c, const PrimRec* d_prims_in, uint32_t n, cudaStream_t s);
int make_world_prims(pb_ctx* c, PrimRec* d_out, cudaStream_t s);   // also computes scene box + delta + per-object boxes
// refit after object `object`'s transform changed (d_objs already holds the new matrices).  *done = false: the scene box
// (and with it delta, part of the intersection contract) changed, the caller must rebuild
int refit_object(pb_ctx* c, uint32_t object, bool* done, cudaStream_t s);
// pb_trace.cu
int trace_rays(pb_ctx* c, const pb_ray* d_rays, uint64_t n, pb_hit* d_hits, uint8_t* d_occ, bool any, bool counted, cudaStream_t s);
// pb_wavefront.cu
int wavefront_render(pb_ctx* c, const pb_render_params* p, float* d_sum, float* d_albedo, float* d_normal, float* d_test, cudaStream_t s);
void wavefront_free(pb_ctx* c);
int tonemap_frame(pb_ctx* c, const float* d_frame, uint8_t* d_rgba8, uint32_t n, int tone, int gamma, cudaStream_t s);
int wavefront_wait(pb_ctx* c);   // synchronises pending (PB_RENDER_ASYNC) renders and folds them into pb_stats
int resolve_frame(pb_ctx* c, const float* d_sum, float* d_frame, uint32_t n, cudaStream_t s);
int film_add(pb_ctx* c, float* d_dst, const float* d_src, uint32_t n_px, cudaStream_t s);   // dst += src (float4 per pixel)
// pb_comm.cu
int comm_reserve_film(pb_ctx* c, size_t n_px);
int comm_reduce_enqueue(pb_ctx* c, float* d_film, size_t n_px, cudaStream_t s);   // ncclReduce(sum, fp32, root 0), in place
int comm_poll(pb_ctx* c);                                                          // asynchronous NCCL errors -> PB_ERR_NCCL
void comm_free(pb_ctx* c);
int comm_group_begin(pb_ctx* c);
int comm_group_end(pb_ctx* c);
}  // namespace pb
