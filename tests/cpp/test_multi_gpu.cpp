// tests/cpp/test_multi_gpu.cpp — TEST PROGRAM (C++, no Python, no torch): the multi-GPU path of the library through its C
// ABIs only (include/pupil_b200.h, include/pupil_host.h), as a C++ host of the reference would use it.
//
//   test_multi_gpu <scene.xml> <n_gpus> [frames]
//
// 1. determinism: rank r's frames (first + r + k n) rendered on GPU r are BIT-IDENTICAL to the same frames
//    rendered on GPU 0 — scene and BVH are rebuilt per device from the same host data, seeds depend on (pixel, frame) only;
// 2. pb_render_multi (one thread, n contexts from pb_create_multi, grouped ncclReduce) == the sum of those per-rank films
//    (bit-exact for n <= 2 where the sum has one order, 1e-6 relative otherwise) == the 1-GPU film up to float
//    re-association (1e-5 relative), with the frame counts in .w exact;
// 3. accumulation over calls on the root (no PB_RENDER_CLEAR) and PB_RENDER_ASYNC + pb_wait;
// 4. the same through the pass interface: System + B200PTPass on n devices (ph_session_create_multi);
// 5. PB_RENDER_REDUCE without a communicator is refused, pb_comm_info reports ranks, NCCL errors are polled (pb_wait).
// Exit 0 = pass, 1 = fail, 77 = skipped (fewer GPUs than asked for, or NCCL missing).
#include <cuda_runtime.h>

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/pupil_host.h"

static int g_fail = 0;
#define CHECK(cond, ...) do { if (!(cond)) { fprintf(stderr, "FAIL %s:%d: ", __FILE__, __LINE__); fprintf(stderr, __VA_ARGS__); fprintf(stderr, "\n"); ++g_fail; } } while (0)
#define REQUIRE_OK(call, ctx) do { int _rc = (call); if (_rc != PB_OK) { fprintf(stderr, "FAIL %s:%d: %s -> %d (%s)\n", __FILE__, __LINE__, #call, _rc, (ctx) ? pb_last_error(ctx) : ""); return 1; } } while (0)

struct DevFilm {
    float* d = nullptr; size_t n_px = 0; int dev = 0;
    DevFilm(int device, size_t px) : n_px(px), dev(device) { cudaSetDevice(dev); cudaMalloc(&d, px * 16); cudaMemset(d, 0, px * 16); }
    ~DevFilm() { cudaSetDevice(dev); cudaFree(d); }
    std::vector<float> host() const { std::vector<float> h(n_px * 4); cudaSetDevice(dev); cudaMemcpy(h.data(), d, n_px * 16, cudaMemcpyDeviceToHost); return h; }
};

static double max_rel(const std::vector<float>& a, const std::vector<float>& b) {
    double m = 0;
    for (size_t i = 0; i < a.size(); ++i) m = std::fmax(m, std::fabs((double)a[i] - b[i]) / (std::fabs((double)b[i]) + 1e-3));
    return m;
}

int main(int argc, char** argv) {
    if (argc < 3) { fprintf(stderr, "usage: test_multi_gpu <scene.xml> <n_gpus> [frames]\n"); return 2; }
    const char* xml = argv[1];
    const int n = atoi(argv[2]);
    const uint32_t frames = argc > 3 ? (uint32_t)atoi(argv[3]) : 6u;
    int n_dev = 0;
    if (cudaGetDeviceCount(&n_dev) != cudaSuccess || n_dev < n || n < 1) { printf("SKIP: %d GPU(s) visible, %d asked for\n", n_dev, n); return 77; }

    ph_scene* scene = nullptr;
    if (ph_scene_load_xml(xml, &scene) != PB_OK) { fprintf(stderr, "cannot load %s: %s\n", xml, ph_last_error()); return 1; }
    ph_scene_info info;
    ph_scene_info_get(scene, &info);
    const uint32_t W = (uint32_t)info.film_w, H = (uint32_t)info.film_h, depth = (uint32_t)info.max_depth;
    const size_t n_px = (size_t)W * H;

    // ---- contexts: one plain context on GPU 0, and the n ranks of one communicator
    pb_ctx* solo = nullptr;
    REQUIRE_OK(pb_create(&solo, 0), solo);
    REQUIRE_OK(ph_scene_upload(scene, nullptr, solo), solo);
    std::vector<int> devices(n);
    for (int i = 0; i < n; ++i) devices[i] = i;
    std::vector<pb_ctx*> ctx(n, nullptr);
    int rc = pb_create_multi(ctx.data(), devices.data(), n);
    if (rc == PB_ERR_NCCL) { printf("SKIP: NCCL is not available\n"); return 77; }
    if (rc != PB_OK) { fprintf(stderr, "pb_create_multi -> %d\n", rc); return 1; }
    for (int r = 0; r < n; ++r) REQUIRE_OK(ph_scene_upload(scene, nullptr, ctx[r]), ctx[r]);
    for (int r = 0; r < n; ++r) { int nr = 0, rk = -1; pb_comm_info(ctx[r], &nr, &rk); CHECK(nr == n && rk == r, "pb_comm_info rank %d: %d of %d", r, rk, nr); }
    { int nr = 0, rk = -1; pb_comm_info(solo, &nr, &rk); CHECK(nr == 1 && rk == 0, "pb_comm_info without a communicator"); }

    // ---- 1. per-rank films: GPU r vs GPU 0, bit for bit
    std::vector<std::vector<float>> rank_film(n);
    for (int r = 0; r < n; ++r) {
        pb_render_params q{};
        q.width = W; q.height = H; q.max_depth = depth;
        q.first_sample = 3u + (uint32_t)r; q.sample_stride = (uint32_t)n;
        q.sample_count = frames > (uint32_t)r ? (frames - r + n - 1) / n : 0;
        q.flags = PB_RENDER_CLEAR;
        DevFilm a(devices[r], n_px), b(0, n_px);
        REQUIRE_OK(pb_render(ctx[r], &q, a.d, nullptr, nullptr, nullptr), ctx[r]);
        REQUIRE_OK(pb_render(solo, &q, b.d, nullptr, nullptr, nullptr), solo);
        rank_film[r] = a.host();
        const std::vector<float> hb = b.host();
        CHECK(memcmp(rank_film[r].data(), hb.data(), n_px * 16) == 0, "rank %d: frames rendered on GPU %d differ from the same frames on GPU 0", r, devices[r]);
    }
    std::vector<float> sum_ranks(n_px * 4, 0.f);
    for (int r = 0; r < n; ++r) for (size_t i = 0; i < sum_ranks.size(); ++i) sum_ranks[i] += rank_film[r][i];

    // ---- 2. pb_render_multi vs the per-rank sum vs one GPU
    pb_render_params p{};
    p.width = W; p.height = H; p.max_depth = depth; p.first_sample = 3u; p.sample_count = frames; p.sample_stride = 1u;
    p.flags = PB_RENDER_CLEAR;
    DevFilm multi(0, n_px), one(0, n_px);
    pb_render_targets tg{multi.d, nullptr, nullptr, nullptr};
    REQUIRE_OK(pb_render_multi(ctx.data(), n, &p, &tg, nullptr), ctx[0]);
    REQUIRE_OK(pb_render(solo, &p, one.d, nullptr, nullptr, nullptr), solo);
    const std::vector<float> hm = multi.host(), h1 = one.host();
    if (n <= 2) CHECK(memcmp(hm.data(), sum_ranks.data(), n_px * 16) == 0, "reduced film != sum of the per-rank films (n <= 2: one summation order)");
    else CHECK(max_rel(hm, sum_ranks) <= 1e-6, "reduced film vs sum of the per-rank films: %.3g", max_rel(hm, sum_ranks));
    CHECK(max_rel(hm, h1) <= 1e-5, "reduced film vs the 1-GPU film: max rel %.3g", max_rel(hm, h1));
    for (size_t i = 0; i < n_px; ++i) if (hm[4 * i + 3] != (float)frames) { CHECK(false, "pixel %zu holds %g frames, want %u", i, hm[4 * i + 3], frames); break; }

    // ---- 3. accumulate over calls (no CLEAR) + asynchronous call and pb_wait
    pb_render_params p2 = p;
    p2.first_sample = 3u + frames; p2.flags = PB_RENDER_ASYNC;
    REQUIRE_OK(pb_render_multi(ctx.data(), n, &p2, &tg, nullptr), ctx[0]);
    for (int r = 0; r < n; ++r) REQUIRE_OK(pb_wait(ctx[r]), ctx[r]);
    p2.flags = 0;
    REQUIRE_OK(pb_render(solo, &p2, one.d, nullptr, nullptr, nullptr), solo);
    const std::vector<float> hm2 = multi.host(), h12 = one.host();
    CHECK(max_rel(hm2, h12) <= 1e-5, "accumulated film vs 1 GPU: max rel %.3g", max_rel(hm2, h12));
    CHECK(hm2[3] == 2.f * frames, "accumulated frame count %g", hm2[3]);

    // ---- 5. argument checks
    pb_render_params bad = p;
    bad.flags = PB_RENDER_CLEAR | PB_RENDER_REDUCE;
    CHECK(pb_render(solo, &bad, one.d, nullptr, nullptr, nullptr) == PB_ERR_INVALID, "PB_RENDER_REDUCE without a communicator must be refused");
    if (n == 1) {   // a 1-rank communicator still walks the NCCL path
        DevFilm red(0, n_px);
        REQUIRE_OK(pb_render(ctx[0], &bad, red.d, nullptr, nullptr, nullptr), ctx[0]);
        const std::vector<float> hr = red.host();
        CHECK(memcmp(hr.data(), h1.data(), n_px * 16) == 0, "1-rank PB_RENDER_REDUCE film != plain film");
    }
    for (int r = 0; r < n; ++r) pb_destroy(ctx[r]);
    pb_destroy(solo);
    ph_scene_free(scene);

    // ---- 4. the pass interface on n devices
    ph_session *sa = nullptr, *sb = nullptr;
    const int dev0 = 0;
    if (ph_session_create_multi(xml, devices.data(), n, &sa) != PB_OK || ph_session_create_multi(xml, &dev0, 1, &sb) != PB_OK) {
        fprintf(stderr, "FAIL: ph_session_create_multi: %s\n", ph_last_error());
        return 1;
    }
    ph_session_set_frames_per_run(sa, frames);
    ph_session_set_frames_per_run(sb, frames);
    CHECK(ph_session_run(sa, 2) == PB_OK && ph_session_run(sb, 2) == PB_OK, "ph_session_run: %s", ph_last_error());
    std::vector<float> fa(n_px * 4), fb(n_px * 4);
    CHECK(ph_session_read(sa, "pt accum buffer", fa.data(), n_px * 16) == PB_OK && ph_session_read(sb, "pt accum buffer", fb.data(), n_px * 16) == PB_OK, "read: %s", ph_last_error());
    CHECK(max_rel(fa, fb) <= 1e-5, "B200PTPass on %d GPUs vs 1 GPU: max rel %.3g", n, max_rel(fa, fb));
    CHECK(fa[3] == 2.f * frames && ph_session_sample_count(sa) == 2 * frames, "pass frame count %g / %u", fa[3], ph_session_sample_count(sa));
    ph_session_destroy(sa);
    ph_session_destroy(sb);

    if (g_fail) { printf("test_multi_gpu: %d check(s) FAILED on %d GPU(s)\n", g_fail, n); return 1; }
    printf("test_multi_gpu: ok on %d GPU(s), %ux%u film, %u frames per call\n", n, W, H, frames);
    return 0;
}
