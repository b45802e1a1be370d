// pb_build.cu — kernels + orchestration of the device BVH build (bodies in pb_build.cuh /
// pb_collapse.cuh).  Replaces optixAccelBuild/Compact for GAS and IAS
// (framework/world/gas_manager.cpp:151,173; framework/world/ias_manager.cpp:84,106).
// Grids are sized to cover the element count in 256-thread blocks; all kernels are HBM-streaming.
#include <cub/cub.cuh>

#include "pb_collapse.cuh"
#include "pb_internal.h"
#include "pb_scene.cuh"

namespace pb {

// exact box of one primitive into its object's box: one reduction per warp when the warp's primitives belong to one object
// (meshes), per-lane atomics otherwise (object boundaries, spheres)
__device__ __forceinline__ void object_box_add(uint32_t* obj_box, uint32_t obj, bool valid, const uint32_t l[3], const uint32_t h[3]) {
    const uint32_t first = __shfl_sync(0xFFFFFFFFu, obj, 0);
    const bool uniform = __all_sync(0xFFFFFFFFu, valid && obj == first);
    if (uniform) {
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            const uint32_t a = __reduce_min_sync(0xFFFFFFFFu, l[k]), b = __reduce_max_sync(0xFFFFFFFFu, h[k]);
            if ((threadIdx.x & 31u) == 0u) { atomicMin(&obj_box[6 * first + k], a); atomicMax(&obj_box[6 * first + 3 + k], b); }
        }
    } else if (valid) {
#pragma unroll
        for (int k = 0; k < 3; ++k) { atomicMin(&obj_box[6 * obj + k], l[k]); atomicMax(&obj_box[6 * obj + 3 + k], h[k]); }
    }
}

__global__ void k_make_prims(SceneGeom g, PrimRec* out, uint32_t* box /* lo xyz, hi xyz as ordered uints */, uint32_t* obj_box) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t l[3] = {0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu}, h[3] = {0u, 0u, 0u};
    uint32_t obj = 0;
    if (i < g.n_prims) {
        PrimRec r = make_prim(g, i);
        out[i] = r;
        obj = f2u(r.b.w);
        V3 lo, hi;
        prim_bbox(r, lo, hi);
        l[0] = float_to_ordered(lo.x); l[1] = float_to_ordered(lo.y); l[2] = float_to_ordered(lo.z);
        h[0] = float_to_ordered(hi.x); h[1] = float_to_ordered(hi.y); h[2] = float_to_ordered(hi.z);
    }
    object_box_add(obj_box, obj, i < g.n_prims, l, h);
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        uint32_t a = __reduce_min_sync(0xFFFFFFFFu, l[k]);
        uint32_t b = __reduce_max_sync(0xFFFFFFFFu, h[k]);
        if ((threadIdx.x & 31u) == 0u) { atomicMin(&box[k], a); atomicMax(&box[3 + k], b); }
    }
}

// ---- refit (pb_update_instance): rewrite the moved object's leaf records, then the boxes bottom-up ----------------------
__global__ void k_update_prims(SceneGeom g, PrimRec* prims /* leaf order */, uint32_t n, uint32_t object, uint32_t* obj_box) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t l[3] = {0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu}, h[3] = {0u, 0u, 0u};
    bool mine = false;
    if (i < n) {
        const V4 b = ld_v4(reinterpret_cast<const V4*>(prims) + 3ull * i + 1);
        if (f2u(b.w) == object) {
            const V4 a = ld_v4(reinterpret_cast<const V4*>(prims) + 3ull * i);
            const PrimRec r = make_prim(g, f2u(a.w));     // the record of this global id under the new transform
            prims[i] = r;
            V3 lo, hi;
            prim_bbox(r, lo, hi);
            l[0] = float_to_ordered(lo.x); l[1] = float_to_ordered(lo.y); l[2] = float_to_ordered(lo.z);
            h[0] = float_to_ordered(hi.x); h[1] = float_to_ordered(hi.y); h[2] = float_to_ordered(hi.z);
            mine = true;
        }
    }
    if (__any_sync(0xFFFFFFFFu, mine)) object_box_add(obj_box, object, mine, l, h);
}
__global__ void k_refit_wide(Node8* nodes, const PrimRec* prims, V3* box_lo, V3* box_hi, float delta, uint32_t first, uint32_t count) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < count) refit_wide_node(nodes, prims, box_lo, box_hi, delta, first + i);
}

__global__ void k_make_trishade(SceneGeom g, float* out) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < g.n_prims) make_trishade(g, i, out);
}

__global__ void k_build_prim(BuildArrays b) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < b.n) build_prim(b, i);
}
__global__ void k_radix(BuildArrays b) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i + 1 < b.n) build_radix_node(b, i);
}
__global__ void k_refit(BuildArrays b) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= b.n) return;
    build_refit_from_leaf(b, i, [](uint32_t* f) {
        uint32_t old = atomicAdd(f, 1u);
        __threadfence();   // the sibling's box writes must be visible to the second arrival
        return old;
    });
}
__global__ void k_plan(BuildArrays b, CollapseArrays c) {
    const uint32_t w = blockIdx.x * blockDim.x + threadIdx.x;
    if (w < c.n_work) collapse_plan(b, c, w);
}
__global__ void k_emit(BuildArrays b, CollapseArrays c, const uint32_t* off_inner, const uint32_t* off_prims) {
    const uint32_t w = blockIdx.x * blockDim.x + threadIdx.x;
    if (w < c.n_work) collapse_emit(b, c, w, off_inner[w], off_prims[w]);
}
__global__ void k_level_totals(const uint32_t* cnt_inner, const uint32_t* off_inner, const uint32_t* cnt_prims,
                               const uint32_t* off_prims, uint32_t n, uint32_t* out) {
    out[0] = off_inner[n - 1] + cnt_inner[n - 1];
    out[1] = off_prims[n - 1] + cnt_prims[n - 1];
}
__global__ void k_root_item(WorkItem* w, uint32_t n) { w[0] = {n == 1u ? (0u | kLeafFlag) : 0u, 0u}; }
__global__ void k_set_u32(uint32_t* p, uint32_t v, uint32_t n) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) p[i] = v;
}

static inline unsigned blocks_for(uint64_t n, unsigned t = 256) { return (unsigned)((n + t - 1) / t); }

// world-space primitives in insertion order + exact scene box + delta
int make_world_prims(pb_ctx* c, PrimRec* d_out, cudaStream_t s) {
    uint32_t n = 0;
    for (auto& o : c->h_objs) n += o.n_prims;
    c->scene_lo = v3(0.f); c->scene_hi = v3(0.f); c->delta = 0.f;
    if (n == 0) return PB_OK;
    uint32_t* d_box = nullptr;
    PB_CUDA_TRY(c, cudaMalloc(&d_box, 6 * sizeof(uint32_t)));
    const uint32_t init[6] = {0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0u, 0u, 0u};
    PB_CUDA_TRY(c, cudaMemcpyAsync(d_box, init, sizeof(init), cudaMemcpyHostToDevice, s));
    const size_t n_objs = c->h_objs.size();
    std::vector<uint32_t> obj_init(6 * n_objs);
    for (size_t o = 0; o < n_objs; ++o) for (int k = 0; k < 6; ++k) obj_init[6 * o + k] = k < 3 ? 0xFFFFFFFFu : 0u;
    cudaFree(c->d_obj_box); c->d_obj_box = nullptr;
    PB_CUDA_TRY(c, cudaMalloc(&c->d_obj_box, std::max<size_t>(1, obj_init.size()) * sizeof(uint32_t)));
    PB_CUDA_TRY(c, cudaMemcpyAsync(c->d_obj_box, obj_init.data(), obj_init.size() * sizeof(uint32_t), cudaMemcpyHostToDevice, s));
    SceneGeom g{c->d_pos, c->d_nrm, c->d_uv, c->d_idx, c->d_objs, (uint32_t)c->h_objs.size(), n, nullptr};
    k_make_prims<<<blocks_for(n), 256, 0, s>>>(g, d_out, d_box, c->d_obj_box);
    // per-triangle shading records (pb_scene.cuh)
    cudaFree(c->d_trishade); c->d_trishade = nullptr;
    const size_t n_tri_total = c->h_idx.size() / 3;
    PB_CUDA_TRY(c, cudaMalloc(&c->d_trishade, std::max<size_t>(n_tri_total, 1) * kTriShadeFloats * sizeof(float)));
    k_make_trishade<<<blocks_for(n), 256, 0, s>>>(g, c->d_trishade);
    c->stats.kernel_launches += 1;
    uint32_t box[6];
    PB_CUDA_TRY(c, cudaMemcpyAsync(box, d_box, sizeof(box), cudaMemcpyDeviceToHost, s));
    PB_CUDA_TRY(c, cudaMemcpyAsync(obj_init.data(), c->d_obj_box, obj_init.size() * sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
    PB_CUDA_TRY(c, cudaStreamSynchronize(s));
    cudaFree(d_box);
    c->h_obj_lo.resize(n_objs); c->h_obj_hi.resize(n_objs);
    for (size_t o = 0; o < n_objs; ++o) {
        c->h_obj_lo[o] = {ordered_to_float(obj_init[6 * o]), ordered_to_float(obj_init[6 * o + 1]), ordered_to_float(obj_init[6 * o + 2])};
        c->h_obj_hi[o] = {ordered_to_float(obj_init[6 * o + 3]), ordered_to_float(obj_init[6 * o + 4]), ordered_to_float(obj_init[6 * o + 5])};
    }
    c->scene_lo = {ordered_to_float(box[0]), ordered_to_float(box[1]), ordered_to_float(box[2])};
    c->scene_hi = {ordered_to_float(box[3]), ordered_to_float(box[4]), ordered_to_float(box[5])};
    const float ext = fmaxf(fmaxf(c->scene_hi.x - c->scene_lo.x, c->scene_hi.y - c->scene_lo.y), c->scene_hi.z - c->scene_lo.z);
    c->delta = ext * (1.0f / 262144.0f);
    c->stats.kernel_launches += 1;
    return PB_OK;
}

}  // namespace pb

namespace pb {

namespace {
struct TempPool {  // frees every scratch buffer on scope exit
    std::vector<void*> ptrs;
    ~TempPool() { for (void* p : ptrs) cudaFree(p); }
    template<class T> cudaError_t get(T** out, size_t count) {
        void* p = nullptr;
        cudaError_t e = cudaMalloc(&p, (count ? count : 1) * sizeof(T));
        if (e == cudaSuccess) { ptrs.push_back(p); *out = static_cast<T*>(p); }
        return e;
    }
};
}  // namespace

int build_bvh(pb_ctx* c, const PrimRec* d_prims_in, uint32_t n, cudaStream_t s) {
    if (c->d_nodes) { cudaFree(c->d_nodes); c->d_nodes = nullptr; }
    if (c->d_prims) { cudaFree(c->d_prims); c->d_prims = nullptr; }
    c->n_nodes = 0; c->n_prims = n; c->bvh_levels = 0;
    if (n == 0) return PB_OK;
    TempPool tp;
    BuildArrays b{};
    b.prims_in = d_prims_in; b.n = n; b.delta = c->delta; b.scene_lo = c->scene_lo; b.scene_hi = c->scene_hi;
    b.leaf_max = c->leaf_max;
    b.bottom_max = std::min(c->bottom_max, 8u * c->leaf_max);
    if (n > 0x1fffffffu) { c->err = "more than 2^29 primitives"; return PB_ERR_LIMIT; }
    unsigned long long* keys_alt = nullptr;
    uint32_t* order_alt = nullptr;
    PB_CUDA_TRY(c, tp.get(&b.blo, n)); PB_CUDA_TRY(c, tp.get(&b.bhi, n));
    PB_CUDA_TRY(c, tp.get(&b.keys, n)); PB_CUDA_TRY(c, tp.get(&keys_alt, n));
    PB_CUDA_TRY(c, tp.get(&b.order, n)); PB_CUDA_TRY(c, tp.get(&order_alt, n));
    PB_CUDA_TRY(c, tp.get(&b.left, n)); PB_CUDA_TRY(c, tp.get(&b.right, n));
    PB_CUDA_TRY(c, tp.get(&b.parent_int, n)); PB_CUDA_TRY(c, tp.get(&b.parent_leaf, n));
    PB_CUDA_TRY(c, tp.get(&b.first, n)); PB_CUDA_TRY(c, tp.get(&b.last, n));
    PB_CUDA_TRY(c, tp.get(&b.ilo, n)); PB_CUDA_TRY(c, tp.get(&b.ihi, n));
    PB_CUDA_TRY(c, tp.get(&b.flag, n));
    PB_CUDA_TRY(c, cudaMemsetAsync(b.flag, 0, n * sizeof(uint32_t), s));

    // 1. padded boxes + Morton keys
    k_build_prim<<<blocks_for(n), 256, 0, s>>>(b);
    // 2. radix sort of (key, prim id)
    {
        cub::DoubleBuffer<unsigned long long> dk(b.keys, keys_alt);
        cub::DoubleBuffer<uint32_t> dv(b.order, order_alt);
        size_t bytes = 0;
        PB_CUDA_TRY(c, cub::DeviceRadixSort::SortPairs(nullptr, bytes, dk, dv, (int)n, 0, 63, s));
        char* tmp = nullptr;
        PB_CUDA_TRY(c, tp.get(&tmp, bytes));
        PB_CUDA_TRY(c, cub::DeviceRadixSort::SortPairs(tmp, bytes, dk, dv, (int)n, 0, 63, s));
        b.keys = dk.Current(); b.order = dv.Current();
    }
    // 3. binary radix tree + bottom-up boxes
    if (n > 1) k_radix<<<blocks_for(n - 1), 256, 0, s>>>(b);
    else k_set_u32<<<1, 32, 0, s>>>(b.parent_leaf, kNoRef, 1);
    k_refit<<<blocks_for(n), 256, 0, s>>>(b);
    c->stats.kernel_launches += 4;

    // 4. collapse level by level
    Node8* d_nodes_tmp = nullptr;
    PrimRec* d_prims_out = nullptr;
    const size_t work_cap = (size_t)n / 2 + 8;     // inner wide children have >= 4 primitives each
    PB_CUDA_TRY(c, tp.get(&d_nodes_tmp, (size_t)n / 2 + 8));
    PB_CUDA_TRY(c, cudaMalloc(&d_prims_out, (size_t)n * sizeof(PrimRec)));
    c->d_prims = d_prims_out;
    WorkItem *work = nullptr, *next = nullptr;
    uint32_t *plan = nullptr, *ci = nullptr, *cp = nullptr, *oi = nullptr, *op = nullptr, *d_tot = nullptr;
    PB_CUDA_TRY(c, tp.get(&work, work_cap)); PB_CUDA_TRY(c, tp.get(&next, work_cap));
    PB_CUDA_TRY(c, tp.get(&plan, 8 * work_cap));
    PB_CUDA_TRY(c, tp.get(&ci, work_cap)); PB_CUDA_TRY(c, tp.get(&cp, work_cap));
    PB_CUDA_TRY(c, tp.get(&oi, work_cap)); PB_CUDA_TRY(c, tp.get(&op, work_cap));
    PB_CUDA_TRY(c, tp.get(&d_tot, 2));
    size_t scan_bytes = 0;
    PB_CUDA_TRY(c, cub::DeviceScan::ExclusiveSum(nullptr, scan_bytes, ci, oi, (int)work_cap, s));
    char* scan_tmp = nullptr;
    PB_CUDA_TRY(c, tp.get(&scan_tmp, scan_bytes));
    k_root_item<<<1, 1, 0, s>>>(work, n);
    uint32_t n_work = 1, wide_count = 1, prim_cursor = 0, levels = 0;
    c->level_first.assign(1, 0u);            // level 0 = the root
    while (n_work) {
        ++levels;
        if (wide_count > (size_t)n / 2 + 8) { c->err = "BVH collapse produced more nodes than reserved"; return PB_ERR_LIMIT; }
        CollapseArrays ca{work, n_work, plan, ci, cp, next, wide_count, prim_cursor, d_nodes_tmp, d_prims_out};
        k_plan<<<blocks_for(n_work), 256, 0, s>>>(b, ca);
        PB_CUDA_TRY(c, cub::DeviceScan::ExclusiveSum(scan_tmp, scan_bytes, ci, oi, (int)n_work, s));
        PB_CUDA_TRY(c, cub::DeviceScan::ExclusiveSum(scan_tmp, scan_bytes, cp, op, (int)n_work, s));
        k_level_totals<<<1, 1, 0, s>>>(ci, oi, cp, op, n_work, d_tot);
        k_emit<<<blocks_for(n_work), 256, 0, s>>>(b, ca, oi, op);
        uint32_t tot[2];
        PB_CUDA_TRY(c, cudaMemcpyAsync(tot, d_tot, sizeof(tot), cudaMemcpyDeviceToHost, s));
        PB_CUDA_TRY(c, cudaStreamSynchronize(s));
        c->stats.kernel_launches += 5;
        c->level_first.push_back(wide_count);  // first node of the next level (== the end of this one)
        wide_count += tot[0]; prim_cursor += tot[1];
        if (tot[0] > work_cap) { c->err = "BVH collapse level overflow"; return PB_ERR_LIMIT; }
        n_work = tot[0];
        WorkItem* t = work; work = next; next = t;
    }
    if (prim_cursor != n) { c->err = "BVH collapse lost primitives"; return PB_ERR_LIMIT; }
    if (levels > PB_SMEM_STACK + PB_LOCAL_STACK) { c->err = "BVH deeper than the traversal stack"; return PB_ERR_LIMIT; }
    c->level_first.push_back(wide_count);
    PB_CUDA_TRY(c, cudaMalloc(&c->d_nodes, (size_t)wide_count * sizeof(Node8)));
    PB_CUDA_TRY(c, cudaMemcpyAsync(c->d_nodes, d_nodes_tmp, (size_t)wide_count * sizeof(Node8), cudaMemcpyDeviceToDevice, s));
    cudaFree(c->d_node_lo); cudaFree(c->d_node_hi); c->d_node_lo = c->d_node_hi = nullptr;   // node boxes: filled by the first refit
    PB_CUDA_TRY(c, cudaMalloc(&c->d_node_lo, (size_t)wide_count * sizeof(V3)));
    PB_CUDA_TRY(c, cudaMalloc(&c->d_node_hi, (size_t)wide_count * sizeof(V3)));
    PB_CUDA_TRY(c, cudaStreamSynchronize(s));
    PB_CUDA_TRY(c, cudaGetLastError());
    c->n_nodes = wide_count; c->bvh_levels = levels;
    return PB_OK;
}

// IAS::Update analogue (framework/world/ias_manager.cpp:116-151,187-205): the moved object's world-space leaf records are
// rewritten in place, its exact box is reduced again, and if the scene box still yields the same delta (the padding of
// the intersection contract) every node's child boxes are recomputed level by level from the leaves up.  No sort, no
// collapse, no allocation: device pointers, node / primitive counts and the tree topology stay, so captured wave graphs
// stay valid.  Results do not depend on the tree (pb_intersect.cuh), only speed does — as with OptiX's update builds, a
// large move degrades the hierarchy until the next pb_scene_commit.
int refit_object(pb_ctx* c, uint32_t object, bool* done, cudaStream_t s) {
    *done = false;
    if (!c->allow_refit || !c->d_nodes || !c->d_prims || !c->d_node_lo || c->n_prims == 0 || c->level_first.size() < 2 ||
        c->h_obj_lo.size() != c->h_objs.size() || object >= c->h_objs.size()) return PB_OK;
    const uint32_t n = c->n_prims;
    const uint32_t init[6] = {0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0u, 0u, 0u};
    PB_CUDA_TRY(c, cudaMemcpyAsync(c->d_obj_box + 6ull * object, init, sizeof(init), cudaMemcpyHostToDevice, s));
    SceneGeom g{c->d_pos, c->d_nrm, c->d_uv, c->d_idx, c->d_objs, (uint32_t)c->h_objs.size(), n, nullptr};
    k_update_prims<<<blocks_for(n), 256, 0, s>>>(g, c->d_prims, n, object, c->d_obj_box);
    uint32_t box[6];
    PB_CUDA_TRY(c, cudaMemcpyAsync(box, c->d_obj_box + 6ull * object, sizeof(box), cudaMemcpyDeviceToHost, s));
    PB_CUDA_TRY(c, cudaStreamSynchronize(s));
    c->h_obj_lo[object] = {ordered_to_float(box[0]), ordered_to_float(box[1]), ordered_to_float(box[2])};
    c->h_obj_hi[object] = {ordered_to_float(box[3]), ordered_to_float(box[4]), ordered_to_float(box[5])};
    // scene box = union of the objects' exact boxes; delta exactly as make_world_prims derives it
    const float fm = 3.402823466e38f;
    V3 lo = v3(fm), hi = v3(-fm);
    for (size_t o = 0; o < c->h_objs.size(); ++o) {
        if (c->h_objs[o].n_prims == 0) continue;
        lo = {fminf(lo.x, c->h_obj_lo[o].x), fminf(lo.y, c->h_obj_lo[o].y), fminf(lo.z, c->h_obj_lo[o].z)};
        hi = {fmaxf(hi.x, c->h_obj_hi[o].x), fmaxf(hi.y, c->h_obj_hi[o].y), fmaxf(hi.z, c->h_obj_hi[o].z)};
    }
    const float ext = fmaxf(fmaxf(hi.x - lo.x, hi.y - lo.y), hi.z - lo.z);
    const float delta = ext * (1.0f / 262144.0f);
    if (f2u(delta) != f2u(c->delta)) return PB_OK;      // the contract's padding changes: rebuild (leaf records are rewritten again there)
    c->scene_lo = lo; c->scene_hi = hi;
    for (size_t l = c->level_first.size() - 1; l-- > 0;) {
        const uint32_t first = c->level_first[l], count = c->level_first[l + 1] - first;
        if (count) k_refit_wide<<<blocks_for(count), 256, 0, s>>>(c->d_nodes, c->d_prims, c->d_node_lo, c->d_node_hi, c->delta, first, count);
    }
    PB_CUDA_TRY(c, cudaGetLastError());
    PB_CUDA_TRY(c, cudaStreamSynchronize(s));
    c->stats.kernel_launches += (uint32_t)c->level_first.size();
    *done = true;
    return PB_OK;
}

}  // namespace pb
