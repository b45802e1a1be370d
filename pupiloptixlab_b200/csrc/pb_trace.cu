// pb_trace.cu — ray-cast kernels: closest hit and any hit over the 8-wide BVH.
// Replaces optixLaunch + optixTrace for the ray-cast entry points (pb_trace_closest / pb_trace_any).
//
// Persistent kernel: grid = SM count x resident CTAs, 128 threads per CTA.  A warp claims 32
// consecutive rays at a time from a global cursor (one atomicAdd per warp), so lanes of a warp
// read rays / write hits as contiguous 128-bit vectors.  Traversal stack head in shared memory.
#include "pb_internal.h"

namespace pb {

constexpr int kTraceThreads = 128;

template<bool ANY, bool COUNT>
__global__ void __launch_bounds__(kTraceThreads)
k_trace(TraceScene sc, const float4* __restrict__ rays, unsigned long long n, float4* __restrict__ hits,
        uint8_t* __restrict__ occ, unsigned long long* __restrict__ counters) {
    __shared__ U2 s_stack[PB_SMEM_STACK * kTraceThreads];
    LaneStack st;
    st.init(s_stack + threadIdx.x, kTraceThreads);
    TraceCounters cnt = {0ull, 0ull, 0ull};
    const unsigned lane = threadIdx.x & 31u;
    while (true) {
        unsigned long long base = 0;
        if (lane == 0) base = atomicAdd(&counters[0], 32ull);
        base = __shfl_sync(0xFFFFFFFFu, base, 0);
        if (base >= n) break;
        const unsigned long long i = base + lane;
        if (i < n) {
            const float4 a = __ldg(rays + 2 * i), b = __ldg(rays + 2 * i + 1);
            Hit h;
            traverse<ANY, COUNT>(sc, {a.x, a.y, a.z}, {b.x, b.y, b.z}, a.w, b.w, h, st, &cnt);
            if (ANY) occ[i] = h.prim != 0xFFFFFFFFu ? 1 : 0;
            else hits[i] = make_float4(h.t, h.u, h.v, __uint_as_float(h.prim));
        }
    }
    if (COUNT) {
        atomicAdd(&counters[1], cnt.nodes);
        atomicAdd(&counters[2], cnt.tris);
        atomicAdd(&counters[3], cnt.spheres);
    }
}

int trace_rays(pb_ctx* c, const pb_ray* d_rays, uint64_t n, pb_hit* d_hits, uint8_t* d_occ, bool any, bool counted, cudaStream_t s) {
    if (n == 0) return PB_OK;
    PB_CUDA_TRY(c, cudaMemsetAsync(c->d_counters, 0, 4 * sizeof(unsigned long long), s));
    const TraceScene sc = c->trace_scene();
    const float4* r = reinterpret_cast<const float4*>(d_rays);
    float4* h = reinterpret_cast<float4*>(d_hits);
    int per_sm = 8;
    unsigned long long want = (n + kTraceThreads - 1) / kTraceThreads;
    unsigned grid = (unsigned)std::min<unsigned long long>(want, (unsigned long long)c->sm_count * per_sm);
    if (any) {
        if (counted) k_trace<true, true><<<grid, kTraceThreads, 0, s>>>(sc, r, n, h, d_occ, c->d_counters);
        else k_trace<true, false><<<grid, kTraceThreads, 0, s>>>(sc, r, n, h, d_occ, c->d_counters);
    } else {
        if (counted) k_trace<false, true><<<grid, kTraceThreads, 0, s>>>(sc, r, n, h, d_occ, c->d_counters);
        else k_trace<false, false><<<grid, kTraceThreads, 0, s>>>(sc, r, n, h, d_occ, c->d_counters);
    }
    PB_CUDA_TRY(c, cudaGetLastError());
    c->stats.kernel_launches += 1;
    if (counted) {
        unsigned long long v[4];
        PB_CUDA_TRY(c, cudaMemcpyAsync(v, c->d_counters, sizeof(v), cudaMemcpyDeviceToHost, s));
        PB_CUDA_TRY(c, cudaStreamSynchronize(s));
        c->stats.node_visits = v[1]; c->stats.prim_tests_tri = v[2]; c->stats.prim_tests_sphere = v[3];
    }
    return PB_OK;
}

}  // namespace pb
