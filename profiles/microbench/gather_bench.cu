// microbench: scattered 96 B record gather, LDG.256 x3 per lane vs one cp.async.bulk (TMA 1D) per lane into shared memory
#include <cstdio>
#include <cstdint>
#include <vector>
#include <algorithm>
#include <random>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s:%d %s\n", __FILE__, __LINE__, cudaGetErrorString(e)); return 1; } } while (0)

__device__ __forceinline__ void ld256(const float4* p, float4& a, float4& b) {
    asm volatile("ld.global.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=f"(a.x), "=f"(a.y), "=f"(a.z), "=f"(a.w), "=f"(b.x), "=f"(b.y), "=f"(b.z), "=f"(b.w) : "l"(p));
}

__global__ void __launch_bounds__(128, 8) k_ldg(const float4* rec, const uint32_t* q, uint32_t n, float* out) {
    const unsigned lane = threadIdx.x & 31u;
    const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, n_warps = (gridDim.x * blockDim.x) >> 5;
    float acc = 0.f;
    for (uint32_t base = warp * 32u; base < n; base += n_warps * 32u) {
        const uint32_t i = base + lane;
        if (i < n) {
            const float4* r = rec + 6ull * q[i];
            float4 a, b, c, d, e, f;
            ld256(r, a, b); ld256(r + 2, c, d); ld256(r + 4, e, f);
            acc += a.x + b.y + c.z + d.w + e.x + f.y;
        }
    }
    if (acc == 123.456f) out[0] = acc;
}

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// each lane: one 96 B bulk copy into its shared-memory slot; one mbarrier per warp
template<int DEPTH>
__global__ void __launch_bounds__(128, 8) k_bulk(const float4* rec, const uint32_t* q, uint32_t n, float* out) {
    __shared__ alignas(128) float4 buf[DEPTH][128][6];
    __shared__ alignas(8) unsigned long long bar[DEPTH][4];
    const unsigned lane = threadIdx.x & 31u, w = threadIdx.x >> 5;
    const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, n_warps = (gridDim.x * blockDim.x) >> 5;
    if (lane == 0)
        for (int d = 0; d < DEPTH; ++d) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" :: "r"(smem_u32(&bar[d][w])));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    __syncwarp();
    float acc = 0.f;
    uint32_t phase[DEPTH];
    for (int d = 0; d < DEPTH; ++d) phase[d] = 0;
    auto issue = [&](uint32_t base, int d) {
        const uint32_t i = base + lane;
        const unsigned act = __ballot_sync(0xFFFFFFFFu, i < n);
        if (lane == 0) asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(smem_u32(&bar[d][w])), "r"(96u * __popc(act)) : "memory");
        __syncwarp();
        if (i < n) {
            const float4* r = rec + 6ull * q[i];
            asm volatile("cp.async.bulk.shared::cta.global.mbarrier::complete_tx::bytes [%0], [%1], 96, [%2];"
                         :: "r"(smem_u32(&buf[d][threadIdx.x][0])), "l"(r), "r"(smem_u32(&bar[d][w])) : "memory");
        }
    };
    uint32_t base = warp * 32u;
    const uint32_t step = n_warps * 32u;
    for (int d = 0; d < DEPTH - 1; ++d) if (base + d * step < n) issue(base + d * step, d);
    int slot = 0;
    for (; base < n; base += step) {
        const uint32_t ahead = base + (DEPTH - 1) * step;
        if (ahead < n) issue(ahead, (slot + DEPTH - 1) % DEPTH);
        // wait for this slot
        uint32_t done = 0;
        while (!done)
            asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                         : "=r"(done) : "r"(smem_u32(&bar[slot][w])), "r"(phase[slot]) : "memory");
        phase[slot] ^= 1u;
        const uint32_t i = base + lane;
        if (i < n) {
            const float4* b = buf[slot][threadIdx.x];
            acc += b[0].x + b[1].y + b[2].z + b[3].w + b[4].x + b[5].y;
        }
        __syncwarp();
        slot = (slot + 1) % DEPTH;
    }
    if (acc == 123.456f) out[0] = acc;
}

int main() {
    const uint32_t N = 48u << 20;   // records in the array (4.8 GB)
    const uint32_t n = 32u << 20;   // gathered per launch
    float4* rec; uint32_t* q; float* out;
    CK(cudaMalloc(&rec, (size_t)N * 96)); CK(cudaMalloc(&q, (size_t)n * 4)); CK(cudaMalloc(&out, 4));
    CK(cudaMemset(rec, 0, (size_t)N * 96));
    std::vector<uint32_t> h(n);
    std::mt19937 g(1);
    for (int mode = 0; mode < 4; ++mode) {
        if (mode == 0) for (uint32_t i = 0; i < n; ++i) h[i] = g() % N;                       // fully random
        else { const uint32_t W = mode == 1 ? 64u : mode == 2 ? 131072u : 1048576u; for (uint32_t i = 0; i < n; ++i) h[i] = (uint32_t)(((uint64_t)i * 3 / 2) % N); for (uint32_t i = 0; i + W <= n; i += W) std::shuffle(h.begin() + i, h.begin() + i + W, g); }   // locally shuffled
        CK(cudaMemcpy(q, h.data(), (size_t)n * 4, cudaMemcpyHostToDevice));
        cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
        auto time = [&](auto f, const char* name) {
            f(); f(); cudaDeviceSynchronize();
            cudaEventRecord(e0); for (int r = 0; r < 5; ++r) f(); cudaEventRecord(e1); cudaEventSynchronize(e1);
            float ms; cudaEventElapsedTime(&ms, e0, e1); ms /= 5;
            printf("mode %d %-12s %.3f ms  %.1f Grec/s  %.0f GB/s  (%s)\n", mode, name, ms, n / ms / 1e6, n * 96.0 / ms / 1e6, cudaGetErrorString(cudaGetLastError()));
        };
        time([&] { k_ldg<<<148 * 8, 128>>>(rec, q, n, out); }, "ldg256x3");
        time([&] { k_bulk<1><<<148 * 8, 128>>>(rec, q, n, out); }, "bulk d1");
        time([&] { k_bulk<2><<<148 * 8, 128>>>(rec, q, n, out); }, "bulk d2");
        time([&] { k_bulk<1><<<148 * 16, 128>>>(rec, q, n, out); }, "bulk d1 x16");
    }
    return 0;
}
