// pb_wavefront.cu — wavefront pt+MIS integrator (placeholder until the kernels land in this file).
#include "pb_internal.h"

namespace pb {

__global__ void k_resolve(const float4* __restrict__ sum, float4* __restrict__ frame, uint32_t n) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float4 s = sum[i];
    const float inv = s.w > 0.f ? 1.0f / s.w : 0.f;   // mean of the frames added so far (main.cu:187-193)
    frame[i] = make_float4(s.x * inv, s.y * inv, s.z * inv, 1.f);
}

int resolve_frame(pb_ctx* c, const float* d_sum, float* d_frame, uint32_t n, cudaStream_t s) {
    if (n == 0) return PB_OK;
    k_resolve<<<(n + 255) / 256, 256, 0, s>>>(reinterpret_cast<const float4*>(d_sum), reinterpret_cast<float4*>(d_frame), n);
    PB_CUDA_TRY(c, cudaGetLastError());
    return PB_OK;
}

void wavefront_free(pb_ctx*) {}

int wavefront_render(pb_ctx* c, const pb_render_params*, float*, float*, float*, cudaStream_t) {
    c->err = "wavefront integrator not built into this library yet";
    return PB_ERR_INVALID;
}

}  // namespace pb
