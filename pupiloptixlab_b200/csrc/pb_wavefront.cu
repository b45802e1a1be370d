// pb_wavefront.cu — wavefront pt+MIS integrator.  Replaces the __raygen__main megakernel and its
// miss / closest-hit programs (example/path_tracer/main.cu:36-233) and the per-frame launch loop of
// PTPass::OnRun (example/path_tracer/pt_pass.cpp:39-57).
//
// One WAVE renders K consecutive frames (samples) of every pixel at once: path slot = k * n_pixels
// + pixel.  Per bounce:   extend (closest hit, bins paths by material)  ->  shade<MAT> x 7 + miss
// (emission/MIS at the hit, Russian roulette, NEE sample -> shadow queue, BSDF sample -> next ray)
// ->  connect (any hit; adds the NEE contribution).  After max_depth bounces `accumulate` adds the K
// radiances of each pixel to the sum buffer in frame order (deterministic, no atomics on pixels).
// Path state is SoA in HBM (float4 per field), queues are compacted with warp ballots + one atomic
// per warp, every kernel is persistent over a device-side counter so a wave needs no host sync.
#include <algorithm>
#include <vector>

#include "pb_bsdf.cuh"
#include "pb_emitter.cuh"
#include "pb_geom.cuh"
#include "pb_internal.h"
#include "pb_trace_loop.cuh"
#include "pb_slots.cuh"

namespace pb {

constexpr int kWfThreads = 128;
#ifndef PB_SHADE_MIN_BLOCKS
#define PB_SHADE_MIN_BLOCKS 8   // 64 registers: 28.8 vs 30.0 ms of shading per C4 step at 5 CTAs (80-96 registers); per material, 9 / 10 CTAs for the
                                // smooth ones and 6 for the rough ones are slower as well (profiles/r02_extend.md, call K)
#endif
#ifndef PB_TRACE_MIN_BLOCKS
#define PB_TRACE_MIN_BLOCKS 8   // 64 registers per thread
#endif
enum { C_EXT_A = 0, C_EXT_B = 1, C_SHADOW = 2, C_MAT0 = 3 /* ..10 */, C_CUR_EXT = 11, C_CUR_CON = 12, C_FRAME0 = 13, C_COUNT = 16 };
enum : uint32_t { F_DEPTH_MASK = 0xffffu, F_DELTA = 0x80000000u };   // depth <= kMaxDepthLimit fits 16 bits
constexpr uint32_t kMaxDepthLimit = 1024u;
constexpr float kRayTmin = kPathTmin, kRayTmax = kPathTmax;
// Queue entries are visited in (increasingly random) queue order, so the first touch of a path record is a
// DRAM round trip that nothing in the warp can overlap.  Every consumer therefore asks L2 for the records
// of the entries a fixed distance AHEAD of the ones it is about to use (prefetch.global.L2: no register, no
// scoreboard); kPrefetchAhead entries = a few microseconds of queue progress, far more than the DRAM
// latency and far less than the time L2 (126 MB) keeps a line under this traffic.
constexpr uint32_t kPrefetchAhead = 16384u;
#ifndef PB_TRACE_PREFETCH
#define PB_TRACE_PREFETCH 1
#endif
#ifndef PB_CONNECT_RED
#define PB_CONNECT_RED 1
#endif
// 256-bit accesses to a 32-byte-aligned pair of float4 (sm_100: LDG/STG.E.ENL2.256).  A path record is visited
// by one lane at a random address, so every request costs the L1 one pass per lane; pairing the two halves of
// a sector halves the passes (the shade kernels were bound by exactly that: ~570 L1 wavefronts per warp-vertex).
__device__ __forceinline__ void ld256(const float4* p, float4& a, float4& b) {
    asm volatile("ld.global.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=f"(a.x), "=f"(a.y), "=f"(a.z), "=f"(a.w), "=f"(b.x), "=f"(b.y), "=f"(b.z), "=f"(b.w) : "l"(p) : "memory");
}
__device__ __forceinline__ void st256(float4* p, const float4& a, const float4& b) {
    asm volatile("st.global.v8.f32 [%8], {%0,%1,%2,%3,%4,%5,%6,%7};"
                 :: "f"(a.x), "f"(a.y), "f"(a.z), "f"(a.w), "f"(b.x), "f"(b.y), "f"(b.z), "f"(b.w), "l"(p) : "memory");
}
__device__ __forceinline__ void prefetch_l2(const void* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }
enum { R_ORG = 0, R_DIR = 1, R_HIT = 2, R_OBJ = 3, R_THR = 4, R_RAD = 5, R_COUNT = 6 };

struct WfState {
    // per path slot: one 96-byte record = 3 whole 32-byte sectors (slots are visited in queue order,
    // i.e. at random, so every access costs whole sectors; six separate SoA arrays cost 6 sectors for 88
    // useful bytes and two partial-sector writes, measured 590 B of DRAM traffic per shaded vertex)
    //   [0] o.xyz, depth|flags   [1] d.xyz, -          written by generate / shade
    //   [2] t, u, v, prim        [3] object, -, -, -   written by extend
    //   [4] throughput.xyz, pdf of the BSDF sample that produced the ray   [5] radiance.xyz, rng state
    // tmin / tmax of path rays are the constants of main.cu:77-82,158-163 and are not stored.
    float4* rec;
    // queues
    uint32_t* q_ext[2];
    uint32_t* q_mat;    // 8 queues of `cap` entries: 0 = miss, 1..7 = material type
    float4* sh_ray;     // 2 per shadow entry
    float4* sh_con;     // contribution.xyz, slot bits
    uint32_t* cnt;      // C_* counters
    unsigned long long* totals;  // [0] closest rays, [1] any rays, [2..4] closest node/tri/sphere visits, [5..7] any
    uint32_t cap;
};

// AOV targets of the last frame (main.cu:94-101); any of them may be null
struct AovOut { float* albedo; float* normal; float* test; };

struct Wavefront {
    WfState st{};
    uint32_t cap = 0;
    void* block = nullptr;
    std::vector<cudaEvent_t> events;   // event pool reused by every call
    // renders enqueued (PB_RENDER_ASYNC) or running and not yet folded into pb_stats by wavefront_wait()
    struct Pending {
        std::vector<cudaEvent_t> whole, ext, con, shd;   // begin / end pairs
        size_t ev_used = 0;
        uint32_t launches = 0, calls = 0;
        uint64_t samples = 0;
        cudaStream_t stream = nullptr;
    } pend;
    // One wave (generate, max_depth x {prepare, extend, miss, shade x 7, after_shade, connect}, accumulate: ~195
    // launches at depth 16) captured once as a CUDA graph and replayed with a single cudaGraphLaunch.  The host
    // cost of a step drops from ~200 driver calls to 3, which is what the synchronous host-buffer path is
    // exposed to (measured on a shared box: the enqueue of one step took 0.8 ms normally and 15-57 ms in bursts).
    struct WaveGraph {
        uint64_t epoch; uint32_t width, height, max_depth, stride, kw, k_aov;
        float *d_sum, *d_albedo, *d_normal, *d_test;
        cudaGraphExec_t exec;
    };
    std::vector<WaveGraph> graphs;
};

struct WfScene {
    TraceScene ts;
    TraceScene ts_any;       // k_connect's copy (own batching knobs)
    SceneGeom geom;
    const MatRec* mats;
    EmitterTable em;
    Camera cam;
    uint32_t width, height, n_px, max_depth;
    uint32_t kw;             // frames in this wave (slot_to_kj)
    uint32_t has_emitters;   // 0: no area emitter and no environment - paths end at the primary vertex (NEE would index an empty table)
    uint32_t stride;      // frame index step; the first frame of a wave is s.cnt[C_FRAME0] (set per wave, so that a
                          // captured wave graph can be replayed for any first_sample)
};

// ---- warp-aggregated queue append: every lane of the warp must call ---------------------------------
__device__ __forceinline__ uint32_t warp_append(uint32_t* counter, bool pred) {
    const unsigned m = __ballot_sync(0xFFFFFFFFu, pred);
    if (m == 0u) return 0u;
    const unsigned lane = threadIdx.x & 31u;
    uint32_t base = 0;
    if (lane == (unsigned)(__ffs(m) - 1)) base = atomicAdd(counter, (uint32_t)__popc(m));
    base = __shfl_sync(0xFFFFFFFFu, base, __ffs(m) - 1);
    return base + (uint32_t)__popc(m & ((1u << lane) - 1u));
}

// ---- stage 0: per-iteration bookkeeping (single thread) --------------------------------------------
__global__ void k_prepare(WfState s, int cur, uint32_t set_cur_count, int set) {
    if (set) s.cnt[cur ? C_EXT_B : C_EXT_A] = set_cur_count;
    s.totals[0] += s.cnt[cur ? C_EXT_B : C_EXT_A];
    s.cnt[cur ? C_EXT_A : C_EXT_B] = 0;            // next extend queue
    s.cnt[C_SHADOW] = 0;
    for (int k = 0; k < 8; ++k) s.cnt[C_MAT0 + k] = 0;
    s.cnt[C_CUR_EXT] = 0; s.cnt[C_CUR_CON] = 0;
}
__global__ void k_after_shade(WfState s) { s.totals[1] += s.cnt[C_SHADOW]; }
__global__ void k_set_frame0(WfState s, uint32_t frame0) { s.cnt[C_FRAME0] = frame0; }

// ---- generate: camera rays of K frames (main.cu:37-75) ----------------------------------------------
__global__ void __launch_bounds__(256) k_generate(WfState s, WfScene sc, uint32_t n_paths) {
    const uint32_t slot = blockIdx.x * blockDim.x + threadIdx.x;
    if (slot >= n_paths) return;
    uint32_t k, j;
    slot_to_kj(slot, sc.n_px, sc.kw, k, j);
    const uint32_t pixel = path_pixel(j, sc.width, sc.height);
    const uint32_t x = pixel % sc.width, y = pixel / sc.width;
    Rng rng;
    rng.init(pixel, s.cnt[C_FRAME0] + k * sc.stride);
    const float jx = rng.next(), jy = rng.next();
    const float sx = (static_cast<float>(x) + jx) / static_cast<float>(sc.width);
    const float sy = (static_cast<float>(y) + jy) / static_cast<float>(sc.height);
    const V4 pof = {sx, sy, 0.f, 1.f};
    const float* m = sc.cam.s2c;
    V4 d = {dot(V4{m[0], m[1], m[2], m[3]}, pof), dot(V4{m[4], m[5], m[6], m[7]}, pof), dot(V4{m[8], m[9], m[10], m[11]}, pof),
            dot(V4{m[12], m[13], m[14], m[15]}, pof)};
    { const float inv = 1.0f / d.w; d.x *= inv; d.y *= inv; d.z *= inv; d.w *= inv; }
    d.w = 0.f;
    d = normalize(d);
    const float* w = sc.cam.c2w;
    const V3 dir = normalize(v3(dot(V4{w[0], w[1], w[2], w[3]}, d), dot(V4{w[4], w[5], w[6], w[7]}, d), dot(V4{w[8], w[9], w[10], w[11]}, d)));
    float4* rec = s.rec + (size_t)R_COUNT * slot;
    st256(rec + R_ORG, make_float4(w[3], w[7], w[11], __uint_as_float(0u)), make_float4(dir.x, dir.y, dir.z, 0.f));
    st256(rec + R_THR, make_float4(1.f, 1.f, 1.f, 0.f), make_float4(0.f, 0.f, 0.f, __uint_as_float(rng.s)));
    s.q_ext[0][slot] = slot;
}

// ---- extend: closest hit for every queued path, then bin by material --------------------------------
struct ExtendIO {
    static constexpr bool kFixedRange = true;    // path rays: tmin / tmax are the constants of main.cu:77-82,158-163
    WfState s;
    const ObjRec* __restrict__ objs;
    const MatRec* __restrict__ mats;
    const uint32_t* __restrict__ q;
    uint32_t n;
    uint32_t slot, pend_slot, pend_bin;   // per lane: current path, finished path waiting for its queue slot
    __device__ __forceinline__ bool load(uint32_t i, V3& o, V3& d, float& tmin, float& tmax) {
        slot = q[i];
        if (PB_TRACE_PREFETCH && i + 2u * kPrefetchAhead < n) prefetch_l2(q + i + 2u * kPrefetchAhead);
        if (PB_TRACE_PREFETCH && i + kPrefetchAhead < n) prefetch_l2(s.rec + (size_t)R_COUNT * q[i + kPrefetchAhead]);
        const float4* rec = s.rec + (size_t)R_COUNT * slot;
        float4 a, b;
        ld256(rec + R_ORG, a, b);
        o = {a.x, a.y, a.z}; d = {b.x, b.y, b.z}; tmin = kRayTmin; tmax = kRayTmax;
        return true;
    }
    __device__ __forceinline__ void store(uint32_t, const Hit& h) {
        float4* rec = s.rec + (size_t)R_COUNT * slot;
        st256(rec + R_HIT, make_float4(h.t, h.u, h.v, __uint_as_float(h.prim)), make_float4(__uint_as_float(h.obj), 0.f, 0.f, 0.f));
        pend_slot = slot;
        pend_bin = h.prim == 0xFFFFFFFFu ? 0u : mats[objs[h.obj].material].type;
    }
    // warp-aggregated append of the finished paths to their material queues (all lanes call)
    __device__ __forceinline__ void flush() {
        if (!__any_sync(0xFFFFFFFFu, pend_bin < 8u)) return;
        const unsigned lane = threadIdx.x & 31u;
        const unsigned peers = __match_any_sync(0xFFFFFFFFu, pend_bin);
        const int leader = __ffs(peers) - 1;
        uint32_t base = 0;
        if (pend_bin < 8u && (int)lane == leader) base = atomicAdd(&s.cnt[C_MAT0 + pend_bin], (uint32_t)__popc(peers));
        base = __shfl_sync(0xFFFFFFFFu, base, leader);
        if (pend_bin < 8u) s.q_mat[(size_t)pend_bin * s.cap + base + (uint32_t)__popc(peers & ((1u << lane) - 1u))] = pend_slot;
        pend_bin = 8u;
    }
};

template<bool COUNT>
__global__ void __launch_bounds__(kWfThreads, PB_TRACE_MIN_BLOCKS) k_extend(WfState s, WfScene sc, int cur) {
    __shared__ U2 s_stack[(PB_SMEM_STACK + PB_COLD_ENTRIES) * kWfThreads];
    U2 tail[PB_LOCAL_STACK];
    LaneStack st;
    st.init(s_stack + threadIdx.x, kWfThreads, tail);
    const uint32_t n = s.cnt[cur ? C_EXT_B : C_EXT_A];
    TraceCounters tc = {0ull, 0ull, 0ull};
    ExtendIO io{s, sc.geom.objs, sc.mats, cur ? s.q_ext[1] : s.q_ext[0], n, 0u, 0u, 8u};
    trace_loop<false, COUNT>(sc.ts, io, st, &s.cnt[C_CUR_EXT], n, sc.ts.refill, tc);
    if (COUNT) { atomicAdd(&s.totals[2], tc.nodes); atomicAdd(&s.totals[3], tc.tris); atomicAdd(&s.totals[4], tc.spheres); }
}

// ---- connect: any hit for the shadow queue; unoccluded contributions go to the path radiance --------
struct ConnectIO {
    static constexpr bool kFixedRange = false;
    WfState s;
    uint32_t n;
    __device__ __forceinline__ bool load(uint32_t i, V3& o, V3& d, float& tmin, float& tmax) const {
        if (PB_TRACE_PREFETCH && i + kPrefetchAhead < n) { prefetch_l2(s.sh_ray + 2 * (i + kPrefetchAhead)); prefetch_l2(s.sh_con + i + kPrefetchAhead); }
        float4 a, b;
        ld256(s.sh_ray + 2 * i, a, b);
        o = {a.x, a.y, a.z}; d = {b.x, b.y, b.z}; tmin = a.w; tmax = b.w;
        return true;
    }
    __device__ __forceinline__ void store(uint32_t i, const Hit& h) const {
        if (h.prim != 0xFFFFFFFFu) return;
        const float4 c = s.sh_con[i];
        const uint32_t slot = __float_as_uint(c.w);
        // one shadow ray per path per bounce, so these are plain adds — issued as reductions (RED.ADD.F32, no
        // return value) so that the finishing lane does not hold its warp for a DRAM round trip of the record
        float* rad = (float*)(s.rec + (size_t)R_COUNT * slot + R_RAD);
#if PB_CONNECT_RED
        atomicAdd(rad, c.x); atomicAdd(rad + 1, c.y); atomicAdd(rad + 2, c.z);
#else
        float4 r = *(float4*)rad; r.x += c.x; r.y += c.y; r.z += c.z; *(float4*)rad = r;
#endif
    }
    __device__ __forceinline__ void flush() const {}
};

template<bool COUNT>
__global__ void __launch_bounds__(kWfThreads, PB_TRACE_MIN_BLOCKS) k_connect(WfState s, WfScene sc) {
    __shared__ U2 s_stack[(PB_SMEM_STACK + PB_COLD_ENTRIES) * kWfThreads];
    U2 tail[PB_LOCAL_STACK];
    LaneStack st;
    st.init(s_stack + threadIdx.x, kWfThreads, tail);
    const uint32_t n = s.cnt[C_SHADOW];
    TraceCounters tc = {0ull, 0ull, 0ull};
    ConnectIO io{s, n};
    trace_loop<true, COUNT>(sc.ts_any, io, st, &s.cnt[C_CUR_CON], n, sc.ts_any.refill, tc);
    if (COUNT) { atomicAdd(&s.totals[5], tc.nodes); atomicAdd(&s.totals[6], tc.tris); atomicAdd(&s.totals[7], tc.spheres); }
}


// ---- shade<MAT>: one path vertex.  Follows main.cu:84-183 line by line -----------------------------------
template<int MAT>
__global__ void __launch_bounds__(kWfThreads, PB_SHADE_MIN_BLOCKS) k_shade(WfState s, WfScene sc, int cur) {
    const uint32_t n = s.cnt[C_MAT0 + MAT];
    const uint32_t* q = s.q_mat + (size_t)MAT * s.cap;
    uint32_t* qnext = cur ? s.q_ext[0] : s.q_ext[1];
    uint32_t* next_count = &s.cnt[cur ? C_EXT_A : C_EXT_B];
    const unsigned lane = threadIdx.x & 31u;
    const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, n_warps = (gridDim.x * blockDim.x) >> 5;
    // Requesting the next iteration's record ahead of time (prefetch.global.L2 or .L1, two queue entries ahead) is slower:
    // 32.2 against 27.3 ms of shading per C4 step (round 2, call R; round 1 measured the same on an older layout).  The
    // prefetch pulls whole lines where the loads take three sectors: these kernels run at the DRAM rate of scattered
    // sectors, not on exposed latency (profiles/r01_shade.md, r02_extend.md).
    for (uint32_t base = warp * 32u; base < n; base += n_warps * 32u) {
        const uint32_t i = base + lane;
        bool push_ray = false, push_shadow = false;
        uint32_t slot = 0;
        V3 sh_o = v3(0.f), sh_d = v3(0.f), contrib = v3(0.f);
        float sh_tmax = 0.f;
        if (i < n) {
            slot = q[i];
            float4* rec = s.rec + (size_t)R_COUNT * slot;
            float4 ro4, rd4, h4, ob4, t4, r4;
            ld256(rec + R_ORG, ro4, rd4);
            ld256(rec + R_HIT, h4, ob4);
            ld256(rec + R_THR, t4, r4);
            const V3 ray_o = {ro4.x, ro4.y, ro4.z}, ray_d = {rd4.x, rd4.y, rd4.z};
            const uint32_t prim = __float_as_uint(h4.w);
            const ObjRec& o = sc.geom.objs[__float_as_uint(ob4.x)];
            const MatRec& m = sc.mats[o.material];
            const LocalGeo geo = hit_geometry(sc.geom, o, prim, h4.x, h4.y, h4.z, ray_o, ray_d, m.twosided != 0u);
            const BsdfLocal bsdf = bsdf_local<MAT>(m, geo.texcoord);
            const int emitter_index = o.emitter_offset >= 0 ? o.emitter_offset + (int)(prim - o.prim_offset) : -1;
            V3 throughput = {t4.x, t4.y, t4.z}, radiance = {r4.x, r4.y, r4.z};
            Rng rng; rng.s = __float_as_uint(r4.w);
            const uint32_t fl = __float_as_uint(ro4.w);
            int depth = (int)(fl & F_DEPTH_MASK);
            if (depth == 0) {  // primary hit: main.cu:87-101
                if (emitter_index >= 0) radiance += emitter_radiance(sc.em.rec[emitter_index], geo.texcoord);
                (void)rng.next();  // test[pixel] = rng.Next(), main.cu:101 — the AOVs themselves (main.cu:94-101) are written by
                                   // k_write_aov for the one frame that needs them: no slot -> pixel arithmetic in the shade kernels
            } else if (emitter_index >= 0) {  // emitter hit by the BSDF sample: main.cu:171-182
                const EmitRec& e = sc.em.rec[emitter_index];
                EmitEval ev;
                emitter_eval(e, ev, geo.position, geo.normal, geo.texcoord, ray_o);
                if (!is_zero(ev.pdf)) {
                    const float mis = (fl & F_DELTA) ? 1.f : mis_weight(t4.w, ev.pdf * e.select_probability);
                    radiance += throughput * ev.radiance * mis;
                }
            }
            // ---- loop body, main.cu:103-163
            bool alive = true;
            ++depth;
            if (depth >= (int)sc.max_depth || !sc.has_emitters) alive = false;
            if (alive) {
                const float rr = depth > 2 ? 0.95 : 1.0;
                if (rng.next() > rr) alive = false;
                else throughput /= rr;
            }
            if (alive) {
                {  // direct light sampling, main.cu:113-140
                    const uint32_t ei = select_emitter(sc.em, rng.next());
                    const EmitRec& e = sc.em.rec[ei];
                    const float x1 = rng.next(), x2 = rng.next();
                    EmitterSample es;
                    emitter_sample_direct(e, es, geo.position, geo.normal, x1, x2);
                    if (es.valid_distance) {
                        BsdfRec er;
                        er.wi = to_local(es.wi, geo.normal);
                        er.wo = to_local(-ray_d, geo.normal);
                        bsdf_eval<MAT>(bsdf, er);
                        if (!is_zero(er.f * es.pdf)) {
                            const float NoL = dot(geo.normal, es.wi);
                            if (NoL > 0.f) {
                                const float mis = mis_weight(es.pdf, er.pdf);
                                const float pdf_e = es.pdf * e.select_probability;
                                contrib = throughput * es.radiance * er.f * NoL * mis / pdf_e;
                                sh_o = geo.position; sh_d = es.wi; sh_tmax = es.distance - 0.001f;
                                push_shadow = true;
                            }
                        }
                    }
                }
                {  // bsdf sampling, main.cu:141-163
                    BsdfRec br;
                    br.wo = to_local(-ray_d, geo.normal);
                    bsdf_sample<MAT>(bsdf, br, rng);
                    if (is_zero(br.f * fabsf(br.wi.z)) || is_zero(br.pdf)) {
                        alive = false;
                    } else {
                        throughput *= br.f * fabsf(br.wi.z) / br.pdf;
                        const V3 nd = to_world(br.wi, geo.normal);
                        st256(rec + R_ORG, make_float4(geo.position.x, geo.position.y, geo.position.z,
                                                       __uint_as_float((uint32_t)depth | ((br.lobe & LOBE_DELTA) ? F_DELTA : 0u))),
                              make_float4(nd.x, nd.y, nd.z, 0.f));
                        t4 = make_float4(throughput.x, throughput.y, throughput.z, br.pdf);
                        push_ray = true;
                    }
                }
            }
            // throughput|pdf (unchanged for a finished path) and radiance|rng share a sector: one 256-bit store
            st256(rec + R_THR, t4, make_float4(radiance.x, radiance.y, radiance.z, __uint_as_float(rng.s)));
        }
        const uint32_t pr = warp_append(next_count, push_ray);
        if (push_ray) qnext[pr] = slot;
        const uint32_t ps = warp_append(&s.cnt[C_SHADOW], push_shadow);
        if (push_shadow) {
            st256(s.sh_ray + 2 * ps, make_float4(sh_o.x, sh_o.y, sh_o.z, 0.001f), make_float4(sh_d.x, sh_d.y, sh_d.z, sh_tmax));
            s.sh_con[ps] = make_float4(contrib.x, contrib.y, contrib.z, __uint_as_float(slot));
        }
    }
}

// ---- aov: albedo / normal / test of the primary hits of ONE frame of the wave (main.cu:94-101), between the first extend and
// the first shade (which overwrites the camera ray).  n_px threads, once per pb_render call: the seven shade kernels carry
// no pixel arithmetic for it.
template<int MAT>
__device__ __forceinline__ V3 aov_albedo(const MatRec& m, V2 uv) { return bsdf_albedo<MAT>(bsdf_local<MAT>(m, uv)); }

__global__ void __launch_bounds__(256) k_write_aov(WfState s, WfScene sc, uint32_t k_aov_frame, AovOut aov) {
    const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= sc.n_px) return;
    const uint32_t pixel = path_pixel(j, sc.width, sc.height);
    const float4* rec = s.rec + (size_t)R_COUNT * kj_to_slot(k_aov_frame, j, sc.n_px, sc.kw);
    const float4 ro4 = rec[R_ORG], rd4 = rec[R_DIR], h4 = rec[R_HIT], ob4 = rec[R_OBJ];
    V3 albedo = v3(0.f), normal = v3(0.f);
    const uint32_t prim = __float_as_uint(h4.w);
    if (prim != 0xFFFFFFFFu) {
        const ObjRec& o = sc.geom.objs[__float_as_uint(ob4.x)];
        const MatRec& m = sc.mats[o.material];
        const LocalGeo geo = hit_geometry(sc.geom, o, prim, h4.x, h4.y, h4.z, v3(ro4.x, ro4.y, ro4.z), v3(rd4.x, rd4.y, rd4.z), m.twosided != 0u);
        normal = geo.normal;
        switch (m.type) {   // Material::GetLocalBsdf + LocalBsdf::GetAlbedo, optix_material.h:69-106
            case MAT_DIFFUSE: albedo = aov_albedo<MAT_DIFFUSE>(m, geo.texcoord); break;
            case MAT_DIELECTRIC: albedo = aov_albedo<MAT_DIELECTRIC>(m, geo.texcoord); break;
            case MAT_ROUGH_DIELECTRIC: albedo = aov_albedo<MAT_ROUGH_DIELECTRIC>(m, geo.texcoord); break;
            case MAT_CONDUCTOR: albedo = aov_albedo<MAT_CONDUCTOR>(m, geo.texcoord); break;
            case MAT_ROUGH_CONDUCTOR: albedo = aov_albedo<MAT_ROUGH_CONDUCTOR>(m, geo.texcoord); break;
            case MAT_PLASTIC: albedo = aov_albedo<MAT_PLASTIC>(m, geo.texcoord); break;
            case MAT_ROUGH_PLASTIC: albedo = aov_albedo<MAT_ROUGH_PLASTIC>(m, geo.texcoord); break;
            default: break;
        }
    }
    if (aov.albedo) { aov.albedo[3 * pixel] = albedo.x; aov.albedo[3 * pixel + 1] = albedo.y; aov.albedo[3 * pixel + 2] = albedo.z; }
    if (aov.normal) { aov.normal[3 * pixel] = normal.x; aov.normal[3 * pixel + 1] = normal.y; aov.normal[3 * pixel + 2] = normal.z; }
    if (aov.test) { Rng rng; rng.s = __float_as_uint(rec[R_RAD].w); aov.test[pixel] = rng.next(); }   // the draw behind the two jitter draws, hit or miss
}

// ---- miss: __miss__default (main.cu:196-212) + main.cu:165-169,185 ----------------------------------------
__global__ void __launch_bounds__(256) k_miss(WfState s, WfScene sc) {
    const uint32_t n = s.cnt[C_MAT0];
    const EmitRec& env = sc.em.rec[sc.em.n_area];
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const uint32_t slot = s.q_mat[i];
        float4* rec = s.rec + (size_t)R_COUNT * slot;
        float4 ro4, rd4;
        ld256(rec + R_ORG, ro4, rd4);
        const uint32_t fl = __float_as_uint(ro4.w);
        const bool primary = (fl & F_DEPTH_MASK) == 0u;
        if (env.type == EM_NONE) continue;
        const V3 ray_o = {ro4.x, ro4.y, ro4.z}, rd = normalize(v3(rd4.x, rd4.y, rd4.z));
        EmitEval ev;
        emitter_eval(env, ev, ray_o + rd, v3(0.f), v2(0.f, 0.f), ray_o);
        V3 env_radiance = ev.radiance;
        float4 r4 = rec[R_RAD];
        if (!primary) {
            const float4 t4 = rec[R_THR];
            const float mis = mis_weight(t4.w, ev.pdf);
            env_radiance *= v3(t4.x, t4.y, t4.z) * mis;
        }
        r4.x += env_radiance.x; r4.y += env_radiance.y; r4.z += env_radiance.z;
        rec[R_RAD] = r4;
    }
}

// ---- accumulate: frame-ordered sum of the K radiances of each pixel (main.cu:187-193 as sum/count) ----------
__global__ void __launch_bounds__(256) k_accumulate(WfState s, float4* sum, uint32_t width, uint32_t height, uint32_t k_count) {
    const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x, n_px = width * height;
    if (j >= n_px) return;
    const uint32_t pixel = path_pixel(j, width, height);
    float4 a = sum[pixel];
    for (uint32_t k = 0; k < k_count; ++k) {
        const float4 r = s.rec[(size_t)R_COUNT * kj_to_slot(k, j, n_px, k_count) + R_RAD];
        a.x += r.x; a.y += r.y; a.z += r.z; a.w += 1.f;
    }
    sum[pixel] = a;
}

// dst += src, float4 per pixel: the root's accumulation buffer takes the film NCCL reduced (pb_comm.cu)
__global__ void k_film_add(float4* __restrict__ dst, const float4* __restrict__ src, uint32_t n) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float4 a = dst[i];
    const float4 b = src[i];
    a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w;
    dst[i] = a;
}

__global__ void k_resolve(const float4* __restrict__ sum, float4* __restrict__ frame, uint32_t n) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float4 v = sum[i];
    const float inv = v.w > 0.f ? 1.0f / v.w : 0.f;   // mean of the frames added so far (main.cu:187-193)
    frame[i] = make_float4(v.x * inv, v.y * inv, v.z * inv, 1.f);
}


// ---- host orchestration -------------------------------------------------------------------------------------
// Paths per wave.  Large waves amortise the tail of every persistent kernel and the ~190 launches of a
// wave (measured on C4: 8 Mi paths 324, 16 Mi 361, 32 Mi 382, 64 Mi 391 Msamples/s); state is ~190 B per
// path, so the default 64 Mi paths take ~12.5 GB of the 180 GB and are capped at a quarter of the free
// memory.  PB_WAVE_PATHS overrides.
static uint32_t wave_path_target() {
    const char* e = getenv("PB_WAVE_PATHS");
    long v = e ? atol(e) : 0;
    if (v > 0) return (uint32_t)std::min<long>(v, 0x7fffffffL);
    size_t free_b = 0, total_b = 0;
    uint64_t paths = 64ull << 20;
    if (cudaMemGetInfo(&free_b, &total_b) == cudaSuccess) paths = std::min<uint64_t>(paths, std::max<uint64_t>(free_b / 4 / 192, 1ull << 20));
    return (uint32_t)paths;
}

// Folds every pending render into pb_stats: one stream synchronise, ray / visit totals, summed stage times.
int wavefront_wait(pb_ctx* c) {
    if (!c->wf || c->wf->pend.calls == 0) return PB_OK;
    Wavefront::Pending& pd = c->wf->pend;
    unsigned long long tot[8];
    cudaError_t e = cudaMemcpyAsync(tot, c->wf->st.totals, 64, cudaMemcpyDeviceToHost, pd.stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(pd.stream);
    if (e == cudaSuccess) e = cudaGetLastError();
    auto sum_pairs = [](std::vector<cudaEvent_t>& v) {
        float total = 0.f;
        for (size_t i = 0; i + 1 < v.size(); i += 2) { float t = 0.f; cudaEventElapsedTime(&t, v[i], v[i + 1]); total += t; }
        return total;
    };
    if (e == cudaSuccess) {
        c->stats.last_render_ms = sum_pairs(pd.whole);
        c->stats.rays_closest = tot[0]; c->stats.rays_any = tot[1];
        c->stats.samples = pd.samples;
        c->stats.kernel_launches = pd.launches;
        c->stats.node_visits = tot[2]; c->stats.prim_tests_tri = tot[3]; c->stats.prim_tests_sphere = tot[4];
        c->stats.node_visits_any = tot[5]; c->stats.prim_tests_tri_any = tot[6]; c->stats.prim_tests_sphere_any = tot[7];
        c->stats.extend_launches = (uint32_t)(pd.ext.size() / 2); c->stats.connect_launches = (uint32_t)(pd.con.size() / 2);
        c->stats.shade_groups = (uint32_t)(pd.shd.size() / 2);
        c->stats.extend_ms = sum_pairs(pd.ext); c->stats.connect_ms = sum_pairs(pd.con); c->stats.shade_ms = sum_pairs(pd.shd);
    }
    pd.whole.clear(); pd.ext.clear(); pd.con.clear(); pd.shd.clear();
    pd.ev_used = 0; pd.launches = 0; pd.calls = 0; pd.samples = 0;
    if (e != cudaSuccess) { c->err = std::string("render: ") + cudaGetErrorString(e); return PB_ERR_CUDA; }
    return PB_OK;
}

void wavefront_free(pb_ctx* c) {
    if (!c->wf) return;
    wavefront_wait(c);
    cudaFree(c->wf->block);
    for (cudaEvent_t e : c->wf->events) cudaEventDestroy(e);
    for (auto& g : c->wf->graphs) cudaGraphExecDestroy(g.exec);
    delete c->wf;
    c->wf = nullptr;
}

static int wavefront_reserve(pb_ctx* c, uint32_t cap) {
    if (c->wf && c->wf->cap >= cap) return PB_OK;
    std::vector<cudaEvent_t> keep;
    if (c->wf) { wavefront_wait(c); keep.swap(c->wf->events); }
    wavefront_free(c);
    c->wf = new Wavefront();
    c->wf->events.swap(keep);
    auto al = [](size_t v) { return (v + 255) & ~(size_t)255; };
    const size_t n = cap;
    size_t off = 0;
    auto take = [&](size_t bytes) { size_t o = off; off += al(bytes); return o; };
    const size_t o_rec = take(n * 16 * R_COUNT), o_q0 = take(n * 4), o_q1 = take(n * 4), o_qm = take(n * 4 * 8), o_shr = take(n * 32),
                 o_shc = take(n * 16), o_cnt = take(C_COUNT * 4), o_tot = take(64);
    PB_CUDA_TRY(c, cudaMalloc(&c->wf->block, off));
    char* b = static_cast<char*>(c->wf->block);
    WfState& s = c->wf->st;
    s.rec = (float4*)(b + o_rec);
    s.q_ext[0] = (uint32_t*)(b + o_q0); s.q_ext[1] = (uint32_t*)(b + o_q1); s.q_mat = (uint32_t*)(b + o_qm);
    s.sh_ray = (float4*)(b + o_shr); s.sh_con = (float4*)(b + o_shc);
    s.cnt = (uint32_t*)(b + o_cnt); s.totals = (unsigned long long*)(b + o_tot);
    s.cap = cap;
    c->wf->cap = cap;
    return PB_OK;
}

template<int MAT>
static void launch_shade(const WfState& st, const WfScene& sc, int cur, unsigned grid, cudaStream_t s) {
    k_shade<MAT><<<grid, kWfThreads, 0, s>>>(st, sc, cur);
}

int wavefront_render(pb_ctx* c, const pb_render_params* p, float* d_sum, float* d_albedo, float* d_normal, float* d_test, cudaStream_t s) {
    const uint32_t n_px = p->width * p->height;
    if ((uint64_t)p->width * p->height > (1ull << 30)) { c->err = "film too large"; return PB_ERR_LIMIT; }
    // config.max_depth is an int in the reference (path_tracer/type.h:12; mitsuba writes -1 for "unbounded"): values <= 1
    // end every path at its primary hit (main.cu:104-106), so one bounce is enqueued; the depth field of the path
    // record holds 16 bits
    const int depth_i = (int)p->max_depth;
    if (depth_i > (int)kMaxDepthLimit) { c->err = "max_depth above 1024 is not supported"; return PB_ERR_LIMIT; }
    const uint32_t n_bounces = depth_i <= 1 ? 1u : (uint32_t)depth_i;
    const AovOut aov{d_albedo, d_normal, d_test};
    if (p->flags & PB_RENDER_CLEAR) PB_CUDA_TRY(c, cudaMemsetAsync(d_sum, 0, (size_t)n_px * 16, s));
    if (p->sample_count == 0) return PB_OK;
    // decided on the first render of a context with the state already allocated by its scene, and kept (PB_WAVE_PATHS is
    // still read every time: tests change it): cudaMemGetInfo took up to 95 ms at the head of a pb_render on a busy box
    { const bool env = getenv("PB_WAVE_PATHS") != nullptr;
      if (c->wave_target == 0u || env || c->wave_target_env) { c->wave_target = wave_path_target(); c->wave_target_env = env; } }
    uint32_t K = std::max<uint32_t>(1u, c->wave_target / n_px);
    K = std::min(K, p->sample_count);
    if ((uint64_t)K * n_px > 0x7fffffffull) { c->err = "wave too large"; return PB_ERR_LIMIT; }
    int rc = wavefront_reserve(c, K * n_px);
    if (rc != PB_OK) return rc;
    const WfState& st = c->wf->st;
    WfScene sc{};
    sc.ts = c->trace_scene();
    sc.ts_any = c->trace_scene_any();
    sc.geom = SceneGeom{c->d_pos, c->d_nrm, c->d_uv, c->d_idx, c->d_objs, (uint32_t)c->h_objs.size(), c->n_prims, c->d_trishade};
    sc.mats = c->d_mats;
    sc.em = EmitterTable{c->d_emit, c->d_emit_prefix, c->n_area_emitters};
    sc.cam = c->cam;
    sc.width = p->width; sc.height = p->height; sc.n_px = n_px; sc.max_depth = p->max_depth;
    sc.stride = p->sample_stride ? p->sample_stride : 1u;
    sc.has_emitters = (c->n_area_emitters > 0u || c->has_env) ? 1u : 0u;

    Wavefront::Pending& pd = c->wf->pend;
    if (pd.calls && pd.stream != s) { rc = wavefront_wait(c); if (rc != PB_OK) return rc; }   // one stream at a time
    // graphs captured against an older scene / camera: drop them (after the renders that may still be replaying them)
    if (!c->wf->graphs.empty() && c->wf->graphs.front().epoch != c->scene_epoch) {
        if (pd.calls) { rc = wavefront_wait(c); if (rc != PB_OK) return rc; }
        for (auto& g : c->wf->graphs) cudaGraphExecDestroy(g.exec);
        c->wf->graphs.clear();
    }
    if (pd.calls == 0) PB_CUDA_TRY(c, cudaMemsetAsync(st.totals, 0, 64, s));                   // totals add up over pending calls
    pd.stream = s;
    const bool counted = (p->flags & PB_RENDER_COUNTED) != 0u, timing = (p->flags & PB_RENDER_STAGE_TIMING) != 0u;
    auto take_event = [&]() -> cudaEvent_t {
        std::vector<cudaEvent_t>& pool = c->wf->events;
        if (pd.ev_used == pool.size()) { cudaEvent_t e; if (cudaEventCreate(&e) != cudaSuccess) return nullptr; pool.push_back(e); }
        return pool[pd.ev_used++];
    };
    auto record = [&](std::vector<cudaEvent_t>& v) {
        cudaEvent_t e = take_event();
        if (e) { cudaEventRecord(e, s); v.push_back(e); }
    };
    auto mark = [&](std::vector<cudaEvent_t>& v) { if (timing) record(v); };
    std::vector<cudaEvent_t>&ev_ext = pd.ext, &ev_con = pd.con, &ev_shd = pd.shd;
    record(pd.whole);
    const unsigned g_trace = (unsigned)c->sm_count * 8u, g_shade = (unsigned)c->sm_count * 8u;
    const bool use_graph = c->use_graphs && !counted && !timing;   // event records stay on the plain launch path
    uint32_t launches = 0;
    for (uint32_t k0 = 0; k0 < p->sample_count; k0 += K) {
        const uint32_t kw = std::min(K, p->sample_count - k0);
        const uint32_t n_paths = kw * n_px;
        sc.kw = kw;
        const bool last_wave = k0 + kw >= p->sample_count;
        const uint32_t k_aov = last_wave ? kw - 1 : 0xFFFFFFFFu;   // AOVs of the last frame (main.cu:94-99)
        k_set_frame0<<<1, 1, 0, s>>>(st, p->first_sample + k0 * sc.stride);
        ++launches;
        const bool want_aov = k_aov != 0xFFFFFFFFu && (aov.albedo || aov.normal || aov.test);
        auto enqueue_wave = [&]() {
            k_generate<<<(n_paths + 255) / 256, 256, 0, s>>>(st, sc, n_paths);
            int cur = 0;
            for (uint32_t b = 0; b < n_bounces; ++b) {
                k_prepare<<<1, 1, 0, s>>>(st, cur, n_paths, b == 0 ? 1 : 0);
                mark(ev_ext);
                if (counted) k_extend<true><<<g_trace, kWfThreads, 0, s>>>(st, sc, cur);
                else k_extend<false><<<g_trace, kWfThreads, 0, s>>>(st, sc, cur);
                mark(ev_ext);
                if (b == 0 && want_aov) k_write_aov<<<(n_px + 255) / 256, 256, 0, s>>>(st, sc, k_aov, aov);
                mark(ev_shd);
                k_miss<<<(unsigned)c->sm_count * 4u, 256, 0, s>>>(st, sc);
                launch_shade<1>(st, sc, cur, g_shade, s);
                launch_shade<2>(st, sc, cur, g_shade, s);
                launch_shade<3>(st, sc, cur, g_shade, s);
                launch_shade<4>(st, sc, cur, g_shade, s);
                launch_shade<5>(st, sc, cur, g_shade, s);
                launch_shade<6>(st, sc, cur, g_shade, s);
                launch_shade<7>(st, sc, cur, g_shade, s);
                k_after_shade<<<1, 1, 0, s>>>(st);
                mark(ev_shd);
                mark(ev_con);
                if (counted) k_connect<true><<<g_trace, kWfThreads, 0, s>>>(st, sc);
                else k_connect<false><<<g_trace, kWfThreads, 0, s>>>(st, sc);
                mark(ev_con);
                cur ^= 1;
            }
            k_accumulate<<<(n_px + 255) / 256, 256, 0, s>>>(st, reinterpret_cast<float4*>(d_sum), p->width, p->height, kw);
        };
        launches += 2u + 12u * n_bounces + (want_aov ? 1u : 0u);
        cudaGraphExec_t exec = nullptr;
        if (use_graph) {
            for (auto& g : c->wf->graphs)
                if (g.epoch == c->scene_epoch && g.width == p->width && g.height == p->height && g.max_depth == p->max_depth &&
                    g.stride == sc.stride && g.kw == kw && g.k_aov == k_aov && g.d_sum == d_sum && g.d_albedo == d_albedo && g.d_normal == d_normal && g.d_test == d_test)
                    exec = g.exec;
            if (!exec && cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal) == cudaSuccess) {
                enqueue_wave();
                cudaGraph_t graph = nullptr;
                if (cudaStreamEndCapture(s, &graph) == cudaSuccess && graph && cudaGraphInstantiate(&exec, graph, 0) == cudaSuccess) {
                    if (c->wf->graphs.size() >= 8 && pd.calls == 0) { cudaGraphExecDestroy(c->wf->graphs.front().exec); c->wf->graphs.erase(c->wf->graphs.begin()); }
                    c->wf->graphs.push_back({c->scene_epoch, p->width, p->height, p->max_depth, sc.stride, kw, k_aov, d_sum, d_albedo, d_normal, d_test, exec});
                } else {
                    exec = nullptr;
                }
                if (graph) cudaGraphDestroy(graph);
                (void)cudaGetLastError();
            }
        }
        if (exec) PB_CUDA_TRY(c, cudaGraphLaunch(exec, s));
        else enqueue_wave();
    }
    record(pd.whole);
    pd.launches += launches; pd.calls += 1; pd.samples += (uint64_t)n_px * p->sample_count;
    PB_CUDA_TRY(c, cudaGetLastError());
    if (p->flags & PB_RENDER_ASYNC) return PB_OK;     // folded into pb_stats by the next pb_wait / pb_get_stats
    return wavefront_wait(c);
}

// output.hlsl:30-73 (ACESToneMapping, GammaCorrection, PSMain) + the UNORM8 store of the render target
__global__ void k_tonemap(const float4* __restrict__ frame, uchar4* __restrict__ out, uint32_t n, int tone, int gamma) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float4 v = frame[i];
    float c[3] = {v.x, v.y, v.z};
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        float x = c[k];
        if (tone) x = (x * (2.51f * x + 0.03f)) / (x * (2.43f * x + 0.59f) + 0.14f);
        if (gamma) x = powf(x, 1.f / 2.2f);
        x = fminf(fmaxf(x, 0.f), 1.f);          // NaN -> 0, as the UNORM conversion does
        c[k] = floorf(x * 255.f + 0.5f);
    }
    out[i] = make_uchar4((unsigned char)c[0], (unsigned char)c[1], (unsigned char)c[2], 255);
}

int tonemap_frame(pb_ctx* c, const float* d_frame, uint8_t* d_rgba8, uint32_t n, int tone, int gamma, cudaStream_t s) {
    if (n == 0) return PB_OK;
    k_tonemap<<<(n + 255) / 256, 256, 0, s>>>(reinterpret_cast<const float4*>(d_frame), reinterpret_cast<uchar4*>(d_rgba8), n, tone, gamma);
    PB_CUDA_TRY(c, cudaGetLastError());
    return PB_OK;
}

int film_add(pb_ctx* c, float* d_dst, const float* d_src, uint32_t n_px, cudaStream_t s) {
    if (n_px == 0) return PB_OK;
    k_film_add<<<(n_px + 255) / 256, 256, 0, s>>>(reinterpret_cast<float4*>(d_dst), reinterpret_cast<const float4*>(d_src), n_px);
    PB_CUDA_TRY(c, cudaGetLastError());
    return PB_OK;
}

int resolve_frame(pb_ctx* c, const float* d_sum, float* d_frame, uint32_t n, cudaStream_t s) {
    if (n == 0) return PB_OK;
    k_resolve<<<(n + 255) / 256, 256, 0, s>>>(reinterpret_cast<const float4*>(d_sum), reinterpret_cast<float4*>(d_frame), n);
    PB_CUDA_TRY(c, cudaGetLastError());
    return PB_OK;
}

}  // namespace pb
