// pb_entry.cu — ray-cast, render, resolve, stats and BVH-export entry points of the C ABI.
#include <chrono>
#include <cstring>

#include "pb_internal.h"

using namespace pb;

#define PB_REQUIRE(cond, msg)                       \
    do {                                            \
        if (!(cond)) { c->err = (msg); return PB_ERR_INVALID; } \
    } while (0)

// temporaries of the host-buffer entry points: released on every return path
struct DevBuf {
    void* p = nullptr;
    ~DevBuf() { if (p) cudaFree(p); }
    cudaError_t alloc(size_t bytes) { return cudaMalloc(&p, bytes ? bytes : 1); }
    template<class T> T* as() const { return static_cast<T*>(p); }
};
struct EventPair {
    cudaEvent_t a = nullptr, b = nullptr;
    ~EventPair() { if (a) cudaEventDestroy(a); if (b) cudaEventDestroy(b); }
};

static int trace_entry(pb_ctx* c, const pb_ray* rays, uint64_t n, void* out, int host, void* stream, bool any, bool counted) {
    if (!c) return PB_ERR_INVALID;
    if (!c->committed) { c->err = "scene not committed"; return PB_ERR_NOT_COMMITTED; }
    PB_REQUIRE((rays && out) || n == 0, "null ray / result pointer");
    cudaSetDevice(c->device);
    cudaStream_t s = stream ? (cudaStream_t)stream : c->stream;
    const size_t out_bytes = any ? n : n * sizeof(pb_hit);
    EventPair ev;
    PB_CUDA_TRY(c, cudaEventCreate(&ev.a)); PB_CUDA_TRY(c, cudaEventCreate(&ev.b));
    const cudaEvent_t e0 = ev.a, e1 = ev.b;
    int rc = PB_OK;
    if (host) {
        DevBuf b_rays, b_out;
        PB_CUDA_TRY(c, b_rays.alloc(n * sizeof(pb_ray)));
        PB_CUDA_TRY(c, b_out.alloc(out_bytes));
        pb_ray* d_rays = b_rays.as<pb_ray>();
        void* d_out = b_out.p;
        PB_CUDA_TRY(c, cudaMemcpyAsync(d_rays, rays, n * sizeof(pb_ray), cudaMemcpyHostToDevice, s));
        cudaEventRecord(e0, s);
        rc = trace_rays(c, d_rays, n, any ? nullptr : (pb_hit*)d_out, any ? (uint8_t*)d_out : nullptr, any, counted, s);
        cudaEventRecord(e1, s);
        if (rc == PB_OK) {
            cudaError_t e = cudaMemcpyAsync(out, d_out, out_bytes, cudaMemcpyDeviceToHost, s);
            if (e == cudaSuccess) e = cudaStreamSynchronize(s);
            if (e != cudaSuccess) { c->err = std::string("trace (host buffers): ") + cudaGetErrorString(e); rc = PB_ERR_CUDA; }
        }
    } else {
        cudaEventRecord(e0, s);
        rc = trace_rays(c, rays, n, any ? nullptr : (pb_hit*)out, any ? (uint8_t*)out : nullptr, any, counted, s);
        cudaEventRecord(e1, s);
        if (rc == PB_OK) {
            cudaError_t e = cudaStreamSynchronize(s);
            if (e != cudaSuccess) { c->err = std::string("trace: ") + cudaGetErrorString(e); rc = PB_ERR_CUDA; }
        }
    }
    if (rc == PB_OK) {
        float ms = 0.f;
        cudaEventElapsedTime(&ms, e0, e1);
        c->stats.last_trace_ms = ms;
        if (any) c->stats.rays_any = n; else c->stats.rays_closest = n;
    }
    return rc;
}

extern "C" {

PB_API int pb_trace_closest(pb_ctx* c, const pb_ray* rays, uint64_t n, pb_hit* hits, int host, void* stream) {
    return trace_entry(c, rays, n, hits, host, stream, false, false);
}
PB_API int pb_trace_any(pb_ctx* c, const pb_ray* rays, uint64_t n, uint8_t* occ, int host, void* stream) {
    return trace_entry(c, rays, n, occ, host, stream, true, false);
}
PB_API int pb_trace_closest_counted(pb_ctx* c, const pb_ray* rays, uint64_t n, pb_hit* hits, int host, void* stream) {
    return trace_entry(c, rays, n, hits, host, stream, false, true);
}
PB_API int pb_trace_any_counted(pb_ctx* c, const pb_ray* rays, uint64_t n, uint8_t* occ, int host, void* stream) {
    return trace_entry(c, rays, n, occ, host, stream, true, true);
}

PB_API int pb_render_ex(pb_ctx* c, const pb_render_params* p, const pb_render_targets* tg, void* stream) {
    if (!c) return PB_ERR_INVALID;
    if (!c->committed) { c->err = "scene not committed"; return PB_ERR_NOT_COMMITTED; }
    PB_REQUIRE(p && tg && tg->accum_sum_rgba, "null render params / accumulation buffer");
    PB_REQUIRE(p->width && p->height, "empty film");
    float* accum_sum_rgba = tg->accum_sum_rgba;
    float* albedo_rgb = tg->albedo_rgb;
    float* normal_rgb = tg->normal_rgb;
    float* test = tg->test;
    cudaSetDevice(c->device);
    cudaStream_t s = stream ? (cudaStream_t)stream : c->stream;
    const size_t npx = (size_t)p->width * p->height;
    const bool reduce = (p->flags & PB_RENDER_REDUCE) != 0u;
    if (reduce && !c->nccl_comm) { c->err = "PB_RENDER_REDUCE without a communicator (pb_comm_init_rank / pb_create_multi)"; return PB_ERR_INVALID; }
    // this rank's frames -> its own cleared film -> ncclReduce to rank 0 on the same stream -> the root adds the result
    auto render_reduced = [&](float* d_dst, float* d_alb, float* d_nrm, float* d_tst) -> int {
        int rc = comm_reserve_film(c, npx);
        if (rc != PB_OK) return rc;
        pb_render_params q = *p;
        q.flags = (p->flags & ~(uint32_t)(PB_RENDER_REDUCE | PB_RENDER_HOST_BUFFERS)) | PB_RENDER_CLEAR | PB_RENDER_ASYNC;
        rc = wavefront_render(c, &q, c->d_reduce, d_alb, d_nrm, d_tst, s);
        if (rc == PB_OK) rc = comm_reduce_enqueue(c, c->d_reduce, npx, s);
        if (rc == PB_OK && c->rank == 0) {
            if (p->flags & PB_RENDER_CLEAR) PB_CUDA_TRY(c, cudaMemcpyAsync(d_dst, c->d_reduce, npx * 16, cudaMemcpyDeviceToDevice, s));
            else rc = film_add(c, d_dst, c->d_reduce, (uint32_t)npx, s);
        }
        return rc;
    };
    auto finish = [&](int rc) -> int {     // synchronous calls: wait, then poll the communicator
        if (p->flags & PB_RENDER_ASYNC) return rc;
        const int rw = wavefront_wait(c);
        if (rc == PB_OK) rc = rw;
        if (rc == PB_OK && reduce) rc = comm_poll(c);
        return rc;
    };
    if (!(p->flags & PB_RENDER_HOST_BUFFERS)) {
        if (reduce) return finish(render_reduced(accum_sum_rgba, albedo_rgb, normal_rgb, test));
        return wavefront_render(c, p, accum_sum_rgba, albedo_rgb, normal_rgb, test, s);   // synchronises unless PB_RENDER_ASYNC
    }
    const auto h0 = std::chrono::steady_clock::now();
    // film staging lives in the context (no cudaMalloc / cudaFree - both synchronise the device - per call)
    if (c->stage_px < npx) {
        wavefront_wait(c);   // an asynchronous render may still be writing the old staging film
        cudaFree(c->d_stage); c->d_stage = nullptr; c->stage_px = 0;
        PB_CUDA_TRY(c, cudaMalloc(&c->d_stage, npx * 44));
        c->stage_px = npx;
    }
    float* d_sum = c->d_stage;
    float* d_alb = albedo_rgb ? c->d_stage + npx * 4 : nullptr;
    float* d_nrm = normal_rgb ? c->d_stage + npx * 7 : nullptr;
    float* d_tst = test ? c->d_stage + npx * 10 : nullptr;
    const bool owns_sum = !reduce || c->rank == 0;      // with PB_RENDER_REDUCE only the root's accumulation buffer is read / written
    if (!(p->flags & PB_RENDER_CLEAR) && owns_sum) PB_CUDA_TRY(c, cudaMemcpyAsync(d_sum, accum_sum_rgba, npx * 16, cudaMemcpyHostToDevice, s));
    pb_render_params q = *p;
    q.flags |= PB_RENDER_ASYNC;          // one synchronise for kernels and copies together (wavefront_wait below)
    int rc = reduce ? render_reduced(d_sum, d_alb, d_nrm, d_tst) : wavefront_render(c, &q, d_sum, d_alb, d_nrm, d_tst, s);
    if (rc == PB_OK) {
        cudaError_t e = owns_sum ? cudaMemcpyAsync(accum_sum_rgba, d_sum, npx * 16, cudaMemcpyDeviceToHost, s) : cudaSuccess;
        if (e == cudaSuccess && albedo_rgb) e = cudaMemcpyAsync(albedo_rgb, d_alb, npx * 12, cudaMemcpyDeviceToHost, s);
        if (e == cudaSuccess && normal_rgb) e = cudaMemcpyAsync(normal_rgb, d_nrm, npx * 12, cudaMemcpyDeviceToHost, s);
        if (e == cudaSuccess && test) e = cudaMemcpyAsync(test, d_tst, npx * 4, cudaMemcpyDeviceToHost, s);
        if (e != cudaSuccess) { c->err = std::string("render (host buffers): ") + cudaGetErrorString(e); rc = PB_ERR_CUDA; }
        const auto h1 = std::chrono::steady_clock::now();
        // PB_RENDER_ASYNC with host buffers: upload, kernels and download are all stream-ordered, so consecutive calls
        // may share the staging film; the caller's buffers must stay valid (and should be pinned) until pb_wait
        if (p->flags & PB_RENDER_ASYNC) return rc;
        const int rw = wavefront_wait(c);
        if (rc == PB_OK) rc = rw;
        if (rc == PB_OK && reduce) rc = comm_poll(c);
        if (getenv("PB_DEBUG_TIMING")) {
            const auto h2 = std::chrono::steady_clock::now();
            auto us = [](auto a, auto b) { return (long)std::chrono::duration_cast<std::chrono::microseconds>(b - a).count(); };
            fprintf(stderr, "[pb] render(host buffers): enqueue %ld us, wait %ld us, kernels %.2f ms\n", us(h0, h1), us(h1, h2), c->stats.last_render_ms);
        }
    }
    return rc;
}

PB_API int pb_render(pb_ctx* c, const pb_render_params* p, float* accum_sum_rgba, float* albedo_rgb, float* normal_rgb, void* stream) {
    const pb_render_targets tg{accum_sum_rgba, albedo_rgb, normal_rgb, nullptr};
    return pb_render_ex(c, p, &tg, stream);
}

PB_API int pb_resolve(pb_ctx* c, const float* sum, float* frame, uint32_t n, int host, void* stream) {
    if (!c) return PB_ERR_INVALID;
    cudaSetDevice(c->device);
    cudaStream_t s = stream ? (cudaStream_t)stream : c->stream;
    if (!host) {
        int rc = resolve_frame(c, sum, frame, n, s);
        if (rc != PB_OK) return rc;
        PB_CUDA_TRY(c, cudaStreamSynchronize(s));
        return PB_OK;
    }
    DevBuf b_a, b_b;
    PB_CUDA_TRY(c, b_a.alloc((size_t)n * 16)); PB_CUDA_TRY(c, b_b.alloc((size_t)n * 16));
    float *d_a = b_a.as<float>(), *d_b = b_b.as<float>();
    PB_CUDA_TRY(c, cudaMemcpyAsync(d_a, sum, (size_t)n * 16, cudaMemcpyHostToDevice, s));
    int rc = resolve_frame(c, d_a, d_b, n, s);
    if (rc == PB_OK) {
        PB_CUDA_TRY(c, cudaMemcpyAsync(frame, d_b, (size_t)n * 16, cudaMemcpyDeviceToHost, s));
        PB_CUDA_TRY(c, cudaStreamSynchronize(s));
    }
    return rc;
}

PB_API int pb_tonemap(pb_ctx* c, const float* frame, uint8_t* rgba8, uint32_t n, int tone, int gamma, int host, void* stream) {
    if (!c) return PB_ERR_INVALID;
    PB_REQUIRE(n == 0 || (frame && rgba8), "null frame / output buffer");
    cudaSetDevice(c->device);
    cudaStream_t s = stream ? (cudaStream_t)stream : c->stream;
    if (!host) {
        int rc = tonemap_frame(c, frame, rgba8, n, tone, gamma, s);
        if (rc != PB_OK) return rc;
        PB_CUDA_TRY(c, cudaStreamSynchronize(s));
        return PB_OK;
    }
    DevBuf b_a, b_b;
    PB_CUDA_TRY(c, b_a.alloc((size_t)n * 16));
    PB_CUDA_TRY(c, b_b.alloc((size_t)n * 4));
    float* d_a = b_a.as<float>();
    uint8_t* d_b = b_b.as<uint8_t>();
    cudaError_t e = cudaMemcpyAsync(d_a, frame, (size_t)n * 16, cudaMemcpyHostToDevice, s);
    int rc = e == cudaSuccess ? tonemap_frame(c, d_a, d_b, n, tone, gamma, s) : PB_ERR_CUDA;
    if (rc == PB_OK) {
        e = cudaMemcpyAsync(rgba8, d_b, (size_t)n * 4, cudaMemcpyDeviceToHost, s);
        if (e == cudaSuccess) e = cudaStreamSynchronize(s);
    }
    if (e != cudaSuccess) { c->err = std::string("tonemap: ") + cudaGetErrorString(e); return PB_ERR_CUDA; }
    return rc;
}

PB_API int pb_wait(pb_ctx* c) {
    if (!c) return PB_ERR_INVALID;
    cudaSetDevice(c->device);
    int rc = wavefront_wait(c);
    if (rc == PB_OK) rc = comm_poll(c);      // a peer that died or a failed link shows up here, not as a hang in the caller
    return rc;
}

PB_API int pb_get_stats(pb_ctx* c, pb_stats* out) {
    if (!c || !out) return PB_ERR_INVALID;
    cudaSetDevice(c->device);
    int rc = wavefront_wait(c);
    if (rc == PB_OK) rc = comm_poll(c);
    if (rc != PB_OK) return rc;
    c->stats.refits = c->refits; c->stats.rebuilds = c->rebuilds;
    *out = c->stats;
    return PB_OK;
}

PB_API int pb_export_bvh(pb_ctx* c, void* nodes80, uint64_t* n_nodes, void* prims48, uint64_t* n_prims) {
    if (!c) return PB_ERR_INVALID;
    if (!c->committed) { c->err = "scene not committed"; return PB_ERR_NOT_COMMITTED; }
    cudaSetDevice(c->device);
    if (n_nodes) *n_nodes = c->n_nodes;
    if (n_prims) *n_prims = c->n_prims;
    if (nodes80 && c->n_nodes) PB_CUDA_TRY(c, cudaMemcpy(nodes80, c->d_nodes, (size_t)c->n_nodes * sizeof(Node8), cudaMemcpyDeviceToHost));
    if (prims48 && c->n_prims) PB_CUDA_TRY(c, cudaMemcpy(prims48, c->d_prims, (size_t)c->n_prims * sizeof(PrimRec), cudaMemcpyDeviceToHost));
    return PB_OK;
}

}  // extern "C"
