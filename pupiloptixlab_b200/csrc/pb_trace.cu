// pb_trace.cu — ray-cast kernels: closest hit and any hit over the 8-wide BVH.
// Replaces optixLaunch + optixTrace for the ray-cast entry points (pb_trace_closest / pb_trace_any).
//
// Persistent kernel: grid = SM count x resident CTAs, 128 threads per CTA.  A warp claims 32
// consecutive rays at a time from a global cursor (one atomicAdd per warp), so lanes of a warp
// read rays / write hits as contiguous 128-bit vectors.  Traversal stack head in shared memory.
#include <mutex>

#include "pb_internal.h"
#include "pb_trace_loop.cuh"

namespace pb {

constexpr int kTraceThreads = 128;

template<bool ANY>
struct RayIO {
    static constexpr bool kFixedRange = false;
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
    __shared__ U2 s_stack[(PB_SMEM_STACK + PB_COLD_ENTRIES) * kTraceThreads];
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
    const TraceScene sc = any ? c->trace_scene_any() : c->trace_scene();
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

// ---- allocation-free closest-hit entry point (pb_trace_closest_static, include/pupil_b200.h) -------------------------------
// The scratch of k_trace - the ray cursor + visit counters, and for host buffers the device staging of rays and hits that
// pb_trace_closest cudaMallocs per call - lives in ONE file-scope device array of fixed size.  Every GPU that loads the
// library holds it (256 B + 48 B x PB_STATIC_TRACE_MAX_RAYS = 12.0 MiB) until the process exits.
//   [0, 256)                         cursor + counters (cleared at the head of every call)
//   [256, 256 + 32 MAX)              pb_ray staging (host-buffer calls)
//   [256 + 32 MAX, 256 + 48 MAX)     pb_hit staging (host-buffer calls)
// All calls on a device share the array and it keeps the last call's contents: two calls that overlap on the GPU
// (different streams, or asynchronous use) race on the cursor and on the staging area.  The entry point is synchronous.
static __device__ __align__(256) unsigned char g_static_scratch[256 + 48ull * PB_STATIC_TRACE_MAX_RAYS];

namespace {
// In host code `g_static_scratch` is only the host-side shadow of the symbol, not a device address: each device's real
// address is taken once, with that device current, by cudaGetSymbolAddress.
std::mutex g_static_mutex;
unsigned char* g_static_ptr[64] = {};
bool g_static_known[64] = {};

cudaError_t static_scratch_for(int device, unsigned char** out) {
    std::lock_guard<std::mutex> lock(g_static_mutex);
    if (device < 0 || device >= 64) return cudaErrorInvalidDevice;
    if (!g_static_known[device]) {
        void* p = nullptr;
        cudaError_t e = cudaSetDevice(device);
        if (e == cudaSuccess) e = cudaGetSymbolAddress(&p, g_static_scratch);
        if (e != cudaSuccess) return e;
        g_static_ptr[device] = static_cast<unsigned char*>(p);
        g_static_known[device] = true;
    }
    *out = g_static_ptr[device];
    return cudaSuccess;
}
}  // namespace

}  // namespace pb

extern "C" PB_API int pb_trace_closest_static(pb_ctx* c, const pb_ray* rays, uint64_t n, pb_hit* hits, int host, void* stream) {
    using namespace pb;
    if (!c) return PB_ERR_INVALID;
    if (!c->committed) { c->err = "scene not committed"; return PB_ERR_NOT_COMMITTED; }
    if (n > (uint64_t)PB_STATIC_TRACE_MAX_RAYS) {
        c->err = "pb_trace_closest_static: " + std::to_string(n) + " rays exceed the static scratch (PB_STATIC_TRACE_MAX_RAYS = " +
                 std::to_string((unsigned long long)PB_STATIC_TRACE_MAX_RAYS) + "); split the batch or use pb_trace_closest";
        return PB_ERR_LIMIT;
    }
    if (n == 0) return PB_OK;
    if (!rays || !hits) { c->err = "null ray / result pointer"; return PB_ERR_INVALID; }
    unsigned char* base = nullptr;
    PB_CUDA_TRY(c, static_scratch_for(c->device, &base));
    PB_CUDA_TRY(c, cudaSetDevice(c->device));
    cudaStream_t s = stream ? (cudaStream_t)stream : c->stream;
    unsigned long long* counters = reinterpret_cast<unsigned long long*>(base);
    pb_ray* d_rays = reinterpret_cast<pb_ray*>(base + 256);
    pb_hit* d_hits = reinterpret_cast<pb_hit*>(base + 256 + 32ull * PB_STATIC_TRACE_MAX_RAYS);
    PB_CUDA_TRY(c, cudaMemsetAsync(counters, 0, 256, s));                       // the cursor must start at zero
    if (host) PB_CUDA_TRY(c, cudaMemcpyAsync(d_rays, rays, n * sizeof(pb_ray), cudaMemcpyHostToDevice, s));
    const float4* r = reinterpret_cast<const float4*>(host ? d_rays : rays);
    float4* h = reinterpret_cast<float4*>(host ? d_hits : hits);
    const unsigned long long want = (n + kTraceThreads - 1) / kTraceThreads;
    const unsigned grid = (unsigned)std::min<unsigned long long>(want, (unsigned long long)c->sm_count * 8ull);
    k_trace<false, false><<<grid, kTraceThreads, 0, s>>>(c->trace_scene(), r, (uint32_t)n, h, nullptr, counters);
    PB_CUDA_TRY(c, cudaGetLastError());
    if (host) PB_CUDA_TRY(c, cudaMemcpyAsync(hits, d_hits, n * sizeof(pb_hit), cudaMemcpyDeviceToHost, s));
    PB_CUDA_TRY(c, cudaStreamSynchronize(s));
    c->stats.rays_closest = n;
    c->stats.kernel_launches += 1;
    return PB_OK;
}
