// pb_comm.cu — multi-GPU plumbing below the C ABI: one context per GPU, frames sharded by sample index,
// ONE ncclReduce(sum, fp32, root 0) of the accumulation film per pb_render, enqueued on the render's own stream so that
// it runs behind the last wave without a host round trip.  The reference is single-GPU (framework/cuda/context.cpp:15-44
// picks one device; example/path_tracer/pt_pass.cpp:39-57 launches on it), so there is no reference call site to mirror;
// the contract is the determinism rule of SURVEY 8(e): frame s of pixel p is seeded Init(4, p, s) on any GPU.
//
// NCCL is bound at run time: dlopen("libnccl.so.2") returns the copy already mapped into the process (torch's bundled
// one under bench.py / pytest) or the system library (plain C++ hosts).  Linking it would load a second, differently
// versioned libnccl next to torch's.  Without NCCL every entry point here returns PB_ERR_NCCL and the single-GPU path is
// unaffected.
#include <dlfcn.h>
#include <nccl.h>   // types and prototypes only

#include <cstring>
#include <mutex>

#include "pb_internal.h"

namespace pb {

namespace {
struct NcclApi {
    void* lib = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommInitAll)(ncclComm_t*, int, const int*) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*CommAbort)(ncclComm_t) = nullptr;
    ncclResult_t (*CommGetAsyncError)(ncclComm_t, ncclResult_t*) = nullptr;
    ncclResult_t (*Reduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
    std::string why;
};

NcclApi* nccl() {
    static NcclApi api;
    static std::once_flag once;
    std::call_once(once, [] {
        const char* names[] = {"libnccl.so.2", "libnccl.so"};
        for (const char* n : names) {
            api.lib = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
            if (api.lib) break;
        }
        if (!api.lib) { api.why = std::string("NCCL is not available: ") + (dlerror() ? dlerror() : "libnccl.so.2 not found"); return; }
        auto sym = [&](const char* s) { void* p = dlsym(api.lib, s); if (!p && api.why.empty()) api.why = std::string("NCCL symbol missing: ") + s; return p; };
        api.GetUniqueId = (decltype(api.GetUniqueId))sym("ncclGetUniqueId");
        api.CommInitRank = (decltype(api.CommInitRank))sym("ncclCommInitRank");
        api.CommInitAll = (decltype(api.CommInitAll))sym("ncclCommInitAll");
        api.CommDestroy = (decltype(api.CommDestroy))sym("ncclCommDestroy");
        api.CommAbort = (decltype(api.CommAbort))sym("ncclCommAbort");
        api.CommGetAsyncError = (decltype(api.CommGetAsyncError))sym("ncclCommGetAsyncError");
        api.Reduce = (decltype(api.Reduce))sym("ncclReduce");
        api.GroupStart = (decltype(api.GroupStart))sym("ncclGroupStart");
        api.GroupEnd = (decltype(api.GroupEnd))sym("ncclGroupEnd");
        api.GetErrorString = (decltype(api.GetErrorString))sym("ncclGetErrorString");
    });
    return api.why.empty() ? &api : nullptr;
}

int fail(pb_ctx* c, const char* what, ncclResult_t r) {
    NcclApi* a = nccl();
    if (c) c->err = std::string(what) + ": " + (a && a->GetErrorString ? a->GetErrorString(r) : "NCCL error");
    return PB_ERR_NCCL;
}
}  // namespace

int comm_reserve_film(pb_ctx* c, size_t n_px) {
    if (c->reduce_px >= n_px) return PB_OK;
    wavefront_wait(c);                       // an enqueued reduce may still use the old film
    cudaFree(c->d_reduce); c->d_reduce = nullptr; c->reduce_px = 0;
    PB_CUDA_TRY(c, cudaMalloc(&c->d_reduce, n_px * 16));
    c->reduce_px = n_px;
    return PB_OK;
}

int comm_reduce_enqueue(pb_ctx* c, float* d_film, size_t n_px, cudaStream_t s) {
    if (!c->nccl_comm) { c->err = "PB_RENDER_REDUCE without a communicator (pb_comm_init_rank / pb_create_multi)"; return PB_ERR_INVALID; }
    NcclApi* a = nccl();
    if (!a) { c->err = "NCCL is not available"; return PB_ERR_NCCL; }
    const ncclResult_t r = a->Reduce(d_film, d_film, n_px * 4, ncclFloat, ncclSum, 0, (ncclComm_t)c->nccl_comm, s);
    if (r != ncclSuccess) return fail(c, "ncclReduce", r);
    return PB_OK;
}

int comm_poll(pb_ctx* c) {
    if (!c->nccl_comm) return PB_OK;
    NcclApi* a = nccl();
    if (!a) return PB_OK;
    ncclResult_t async = ncclSuccess;
    const ncclResult_t r = a->CommGetAsyncError((ncclComm_t)c->nccl_comm, &async);
    if (r != ncclSuccess) return fail(c, "ncclCommGetAsyncError", r);
    if (async != ncclSuccess && async != ncclInProgress) return fail(c, "asynchronous NCCL error on this communicator", async);
    return PB_OK;
}

void comm_free(pb_ctx* c) {
    if (c->nccl_comm) {
        if (NcclApi* a = nccl()) a->CommDestroy((ncclComm_t)c->nccl_comm);
        c->nccl_comm = nullptr;
    }
    c->rank = 0; c->n_ranks = 1;
    cudaFree(c->d_reduce); c->d_reduce = nullptr; c->reduce_px = 0;
}

int comm_group_begin(pb_ctx* c) {
    NcclApi* a = nccl();
    if (!a) { if (c) c->err = "NCCL is not available"; return PB_ERR_NCCL; }
    const ncclResult_t r = a->GroupStart();
    return r == ncclSuccess ? PB_OK : fail(c, "ncclGroupStart", r);
}
int comm_group_end(pb_ctx* c) {
    NcclApi* a = nccl();
    if (!a) { if (c) c->err = "NCCL is not available"; return PB_ERR_NCCL; }
    const ncclResult_t r = a->GroupEnd();
    return r == ncclSuccess ? PB_OK : fail(c, "ncclGroupEnd", r);
}

}  // namespace pb

using namespace pb;

extern "C" {

PB_API int pb_comm_unique_id(unsigned char id[PB_UNIQUE_ID_BYTES]) {
    static_assert(PB_UNIQUE_ID_BYTES == NCCL_UNIQUE_ID_BYTES, "unique id size");
    if (!id) return PB_ERR_INVALID;
    NcclApi* a = nccl();
    if (!a) return PB_ERR_NCCL;
    ncclUniqueId u;
    if (a->GetUniqueId(&u) != ncclSuccess) return PB_ERR_NCCL;
    memcpy(id, u.internal, PB_UNIQUE_ID_BYTES);
    return PB_OK;
}

PB_API int pb_comm_init_rank(pb_ctx* c, const unsigned char id[PB_UNIQUE_ID_BYTES], int n_ranks, int rank) {
    if (!c) return PB_ERR_INVALID;
    if (!id || n_ranks < 1 || rank < 0 || rank >= n_ranks) { c->err = "pb_comm_init_rank: bad rank / size"; return PB_ERR_INVALID; }
    NcclApi* a = nccl();
    if (!a) { c->err = "NCCL is not available (libnccl.so.2 could not be loaded)"; return PB_ERR_NCCL; }
    cudaSetDevice(c->device);
    wavefront_wait(c);
    comm_free(c);
    ncclUniqueId u;
    memcpy(u.internal, id, PB_UNIQUE_ID_BYTES);
    ncclComm_t comm = nullptr;
    const ncclResult_t r = a->CommInitRank(&comm, n_ranks, u, rank);
    if (r != ncclSuccess) return fail(c, "ncclCommInitRank", r);
    c->nccl_comm = comm; c->rank = rank; c->n_ranks = n_ranks;
    return PB_OK;
}

PB_API int pb_comm_info(const pb_ctx* c, int* n_ranks, int* rank) {
    if (!c) return PB_ERR_INVALID;
    if (n_ranks) *n_ranks = c->nccl_comm ? c->n_ranks : 1;
    if (rank) *rank = c->nccl_comm ? c->rank : 0;
    return PB_OK;
}

PB_API int pb_comm_destroy(pb_ctx* c) {
    if (!c) return PB_ERR_INVALID;
    cudaSetDevice(c->device);
    wavefront_wait(c);
    comm_free(c);
    return PB_OK;
}

// n contexts in this process, rank i on devices[i]; on failure nothing is left allocated
PB_API int pb_create_multi(pb_ctx** out, const int* devices, int n) {
    if (!out || !devices || n < 1 || n > 64) return PB_ERR_INVALID;
    for (int i = 0; i < n; ++i) out[i] = nullptr;
    for (int i = 0; i < n; ++i)
        for (int j = 0; j < i; ++j)
            if (devices[i] == devices[j]) return PB_ERR_INVALID;      // NCCL needs distinct devices within one communicator
    int rc = PB_OK;
    for (int i = 0; i < n && rc == PB_OK; ++i) rc = pb_create(&out[i], devices[i]);
    NcclApi* a = rc == PB_OK ? nccl() : nullptr;
    if (rc == PB_OK && !a) { out[0]->err = "NCCL is not available (libnccl.so.2 could not be loaded)"; rc = PB_ERR_NCCL; }
    if (rc == PB_OK) {
        std::vector<ncclComm_t> comms(n, nullptr);
        const ncclResult_t r = a->CommInitAll(comms.data(), n, devices);
        if (r != ncclSuccess) rc = fail(out[0], "ncclCommInitAll", r);
        else for (int i = 0; i < n; ++i) { out[i]->nccl_comm = comms[i]; out[i]->rank = i; out[i]->n_ranks = n; }
    }
    if (rc != PB_OK) { for (int i = 0; i < n; ++i) { pb_destroy(out[i]); out[i] = nullptr; } }   // the code says why: NO_DEVICE / NCCL / CUDA
    return rc;
}

// The whole job [first_sample + j * stride, j < sample_count) dealt round-robin to the n contexts of this process; the
// reduced film lands in ctxs[0]'s targets.  One thread drives all devices, so the n ncclReduce calls form one group.
PB_API int pb_render_multi(pb_ctx* const* ctxs, int n, const pb_render_params* p, const pb_render_targets* tg, void* const* streams) {
    if (!ctxs || n < 1 || !ctxs[0]) return PB_ERR_INVALID;
    pb_ctx* root = ctxs[0];
    if (!p || !tg || !tg->accum_sum_rgba) { root->err = "null render params / accumulation buffer"; return PB_ERR_INVALID; }
    if (p->flags & PB_RENDER_HOST_BUFFERS) { root->err = "pb_render_multi takes device buffers on ctxs[0]'s device"; return PB_ERR_INVALID; }
    if (!p->width || !p->height) { root->err = "empty film"; return PB_ERR_INVALID; }
    for (int r = 0; r < n; ++r) {
        if (!ctxs[r]) return PB_ERR_INVALID;
        if (!ctxs[r]->committed) { root->err = "scene not committed on rank " + std::to_string(r); return PB_ERR_NOT_COMMITTED; }
        if (n > 1 && (!ctxs[r]->nccl_comm || ctxs[r]->n_ranks != n || ctxs[r]->rank != r)) { root->err = "contexts are not the ranks 0..n-1 of one communicator (pb_create_multi)"; return PB_ERR_INVALID; }
    }
    const size_t n_px = (size_t)p->width * p->height;
    const uint32_t stride0 = p->sample_stride ? p->sample_stride : 1u;
    int rc = PB_OK;
    std::vector<cudaStream_t> ss(n);
    // 1. every rank renders its frames into its own cleared film (asynchronous: all GPUs work at once)
    for (int r = 0; r < n && rc == PB_OK; ++r) {
        pb_ctx* c = ctxs[r];
        cudaSetDevice(c->device);
        ss[r] = streams && streams[r] ? (cudaStream_t)streams[r] : c->stream;
        rc = comm_reserve_film(c, n_px);
        if (rc != PB_OK) break;
        pb_render_params q = *p;
        q.first_sample = p->first_sample + (uint32_t)r * stride0;
        q.sample_stride = stride0 * (uint32_t)n;
        q.sample_count = p->sample_count > (uint32_t)r ? (p->sample_count - (uint32_t)r + (uint32_t)n - 1u) / (uint32_t)n : 0u;
        q.flags = (p->flags & ~(uint32_t)PB_RENDER_REDUCE) | PB_RENDER_CLEAR | PB_RENDER_ASYNC;
        rc = wavefront_render(c, &q, c->d_reduce, r == 0 ? tg->albedo_rgb : nullptr, r == 0 ? tg->normal_rgb : nullptr, r == 0 ? tg->test : nullptr, ss[r]);
        if (rc != PB_OK && c != root) root->err = "rank " + std::to_string(r) + ": " + c->err;
    }
    // 2. one reduce to rank 0 (grouped: a single thread issues the collective for every rank)
    if (rc == PB_OK && n > 1) {
        rc = comm_group_begin(root);
        for (int r = 0; r < n && rc == PB_OK; ++r) {
            cudaSetDevice(ctxs[r]->device);
            rc = comm_reduce_enqueue(ctxs[r], ctxs[r]->d_reduce, n_px, ss[r]);
            if (rc != PB_OK && ctxs[r] != root) root->err = "rank " + std::to_string(r) + ": " + ctxs[r]->err;
        }
        const int re = comm_group_end(root);
        if (rc == PB_OK) rc = re;
    }
    // 3. the root adds the reduced film to the caller's accumulation buffer
    if (rc == PB_OK) {
        cudaSetDevice(root->device);
        if (p->flags & PB_RENDER_CLEAR) { if (cudaMemcpyAsync(tg->accum_sum_rgba, root->d_reduce, n_px * 16, cudaMemcpyDeviceToDevice, ss[0]) != cudaSuccess) { root->err = "film copy failed"; rc = PB_ERR_CUDA; } }
        else rc = film_add(root, tg->accum_sum_rgba, root->d_reduce, (uint32_t)n_px, ss[0]);
    }
    if (p->flags & PB_RENDER_ASYNC) return rc;
    for (int r = 0; r < n; ++r) {
        cudaSetDevice(ctxs[r]->device);
        int rw = wavefront_wait(ctxs[r]);
        if (rw == PB_OK) rw = comm_poll(ctxs[r]);
        if (rw != PB_OK && rc == PB_OK) { rc = rw; if (ctxs[r] != root) root->err = "rank " + std::to_string(r) + ": " + ctxs[r]->err; }
    }
    cudaSetDevice(root->device);
    return rc;
}

}  // extern "C"
