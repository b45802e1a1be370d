// pb_trace.cu — ray-cast kernels: closest hit and any hit over the 8-wide BVH.
// Replaces optixLaunch + optixTrace for the ray-cast entry points (pb_trace_closest / pb_trace_any).
//
// Persistent kernel: grid = SM count x resident CTAs, 128 threads per CTA.  A warp claims 32
// consecutive rays at a time from a global cursor (one atomicAdd per warp), so lanes of a warp
// read rays / write hits as contiguous 128-bit vectors.  Traversal stack head in shared memory.
#include "pb_internal.h"
#include "pb_trace_loop.cuh"

namespace pb {

constexpr int kTraceThreads = 128;

template<bool ANY>
struct RayIO {
    const float4* __restrict__ rays;
    float4* __restrict__ hits;
    uint8_t* __restrict__ occ;
    __device__ __forceinline__ bool load(uint32_t i, V3& o, V3& d, float& tmin, float& tmax) const {
        const float4 a = __ldg(rays + 2ull * i), b = __ldg(rays + 2ull * i + 1);
        o = {a.x, a.y, a.z}; d = {b.x, b.y, b.z}; tmin = a.w; tmax = b.w;
        return true;
    }
    __device__ __forceinline__ void store(uint32_t i, const Hit& h) const {
        if (ANY) occ[i] = h.prim != 0xFFFFFFFFu ? 1 : 0;
        else hits[i] = make_float4(h.t, h.u, h.v, __uint_as_float(h.prim));
    }
    __device__ __forceinline__ void flush() const {}
};

template<bool ANY, bool COUNT>
__global__ void __launch_bounds__(kTraceThreads)
k_trace(TraceScene sc, const float4* __restrict__ rays, uint32_t n, float4* __restrict__ hits,
        uint8_t* __restrict__ occ, unsigned long long* __restrict__ counters) {
    __shared__ U2 s_stack[PB_SMEM_STACK * kTraceThreads];
    U2 tail[PB_LOCAL_STACK];
    LaneStack st;
    st.init(s_stack + threadIdx.x, kTraceThreads, tail);
    TraceCounters cnt = {0ull, 0ull, 0ull};
    RayIO<ANY> io{rays, hits, occ};
    trace_loop<ANY, COUNT>(sc, io, st, reinterpret_cast<uint32_t*>(&counters[0]), n, sc.refill, cnt);
    if (COUNT) {
        atomicAdd(&counters[1], cnt.nodes);
        atomicAdd(&counters[2], cnt.tris);
        atomicAdd(&counters[3], cnt.spheres);
    }
}

int trace_rays(pb_ctx* c, const pb_ray* d_rays, uint64_t n_total, pb_hit* d_hits, uint8_t* d_occ, bool any, bool counted, cudaStream_t s) {
    if (n_total == 0) return PB_OK;
    const TraceScene sc = c->trace_scene();
    unsigned long long visits[3] = {0ull, 0ull, 0ull};
    // the ray cursor is 32-bit: batches of at most 2^30 rays per launch
    for (uint64_t first = 0; first < n_total; first += (1ull << 30)) {
    const uint32_t n = (uint32_t)std::min<uint64_t>(n_total - first, 1ull << 30);
    PB_CUDA_TRY(c, cudaMemsetAsync(c->d_counters, 0, 4 * sizeof(unsigned long long), s));
    const float4* r = reinterpret_cast<const float4*>(d_rays + first);
    float4* h = d_hits ? reinterpret_cast<float4*>(d_hits + first) : nullptr;
    uint8_t* d_occ_b = d_occ ? d_occ + first : nullptr;
    int per_sm = 8;
    unsigned long long want = ((unsigned long long)n + kTraceThreads - 1) / kTraceThreads;
    unsigned grid = (unsigned)std::min<unsigned long long>(want, (unsigned long long)c->sm_count * per_sm);
    if (any) {
        if (counted) k_trace<true, true><<<grid, kTraceThreads, 0, s>>>(sc, r, n, h, d_occ_b, c->d_counters);
        else k_trace<true, false><<<grid, kTraceThreads, 0, s>>>(sc, r, n, h, d_occ_b, c->d_counters);
    } else {
        if (counted) k_trace<false, true><<<grid, kTraceThreads, 0, s>>>(sc, r, n, h, d_occ_b, c->d_counters);
        else k_trace<false, false><<<grid, kTraceThreads, 0, s>>>(sc, r, n, h, d_occ_b, c->d_counters);
    }
    PB_CUDA_TRY(c, cudaGetLastError());
    c->stats.kernel_launches += 1;
    if (counted) {
        unsigned long long v[4];
        PB_CUDA_TRY(c, cudaMemcpyAsync(v, c->d_counters, sizeof(v), cudaMemcpyDeviceToHost, s));
        PB_CUDA_TRY(c, cudaStreamSynchronize(s));
        visits[0] += v[1]; visits[1] += v[2]; visits[2] += v[3];
    }
    }
    if (counted) { c->stats.node_visits = visits[0]; c->stats.prim_tests_tri = visits[1]; c->stats.prim_tests_sphere = visits[2]; }
    return PB_OK;
}

}  // namespace pb
