// tests/devsim/devsim.cu — TEST HARNESS ONLY (nvcc, sm_100a, the product's flags: -fmad=false).  Runs the product's
// BSDF headers on the GPU outside the integrator so that pb_bsdf.cuh itself — not only the oracle's port — is checked
// against the reference-header tables of tests/golden/reference_headers.npz.  Not a product export: built into
// tests/devsim/libdevsim.so, loaded by tests/test_gpu_bsdf_tables.py only.
#include <cuda_runtime.h>

#include "../../pupiloptixlab_b200/csrc/pb_host_prep.h"
#include "table_body.h"

using namespace pb;

__global__ void k_bsdf_table(MatRec m, const float* __restrict__ wo, const uint32_t* __restrict__ seeds, uint32_t n, float* __restrict__ out) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    bsdf_table_entry_any(m, wo + 3 * i, seeds[i], out + 16 * i);
}

extern "C" __attribute__((visibility("default")))
int ds_bsdf_table(const pb_material* mat, const float* wo, const uint32_t* seeds, uint32_t n, float* out) {
    if (!mat || mat->reflectance.type == PB_TEX_BITMAP) return 1;   // rgb / checkerboard parameters only
    const MatRec m = prep::mat_rec(*mat);
    float *d_wo = nullptr, *d_out = nullptr;
    uint32_t* d_seeds = nullptr;
    cudaError_t e = cudaMalloc(&d_wo, 12ull * n);
    if (e == cudaSuccess) e = cudaMalloc(&d_seeds, 4ull * n);
    if (e == cudaSuccess) e = cudaMalloc(&d_out, 64ull * n);
    if (e == cudaSuccess) e = cudaMemcpy(d_wo, wo, 12ull * n, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(d_seeds, seeds, 4ull * n, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) {
        k_bsdf_table<<<(n + 127) / 128, 128>>>(m, d_wo, d_seeds, n, d_out);
        e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaMemcpy(out, d_out, 64ull * n, cudaMemcpyDeviceToHost);
    cudaFree(d_wo); cudaFree(d_seeds); cudaFree(d_out);
    return e == cudaSuccess ? 0 : 2;
}
