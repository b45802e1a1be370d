// pb_traverse.cuh — closest-hit / any-hit traversal of the compressed 8-wide BVH.
//
// Replaces optixTrace (example/path_tracer/main.cu:77-82,158-163: ray type 0, closest hit) and
// Emitter::TraceShadowRay (framework/render/emitter.h:91-100: ray type 1, terminate on first hit).
//
// One ray per lane.  Node = 5 x 128-bit loads, primitive = 3 x 128-bit loads (read-only path).
// Traversal state is two 64-bit "groups" (Ylitie et al. 2017): G = (child base, inner-hit bits |
// imask), P = (prim base, prim-hit bits); deferred node groups go to a per-lane stack whose first
// PB_SMEM_STACK entries live in shared memory ([entry][lane] layout, conflict free) and spill to
// local memory beyond that.  Children are visited front to back through the octant trick: the
// builder places children in octant-ordered slots and bit (24 + (slot ^ octinv)) of the hit mask
// orders them for this ray.
//
// Result = lexicographic min of (t, prim id) over accepted hits (pb_intersect.cuh), which is
// independent of tree shape and visiting order; the node test below may therefore use folded
// arithmetic (one FMA per plane): its rounding error is covered by the delta gap between B_in and B_out.
#pragma once
#include "pb_intersect.cuh"

namespace pb {

struct alignas(16) U4 { uint32_t x, y, z, w; };

struct TraceScene {
    const U4* nodes;      // Node8 as 5 x U4
    const V4* prims;      // PrimRec as 3 x V4
    const ObjRec* objs;
    float delta;
    uint32_t n_nodes;
    uint32_t refill;      // refill idle lanes when fewer than this many lanes of a warp hold a ray
    uint32_t prim_batch;  // run the primitive tests once this many lanes of a warp hold primitives
    uint32_t k47;         // 0x47000000 (see byte_f)
    uint32_t prim_iters;  // primitives a lane tests per primitive phase; the rest waits for the next batch
};

struct Hit { float t, u, v; uint32_t prim; uint32_t obj; };
struct TraceCounters { unsigned long long nodes, tris, spheres; };

#if defined(__CUDA_ARCH__)
PB_D U4 ld_u4(const U4* p) { uint4 r = __ldg(reinterpret_cast<const uint4*>(p)); return {r.x, r.y, r.z, r.w}; }
PB_D V4 ld_v4(const V4* p) { float4 r = __ldg(reinterpret_cast<const float4*>(p)); return {r.x, r.y, r.z, r.w}; }
PB_D int hi_bit(uint32_t x) { return 31 - __clz(x); }
PB_D int lo_bit(uint32_t x) { return __ffs(x) - 1; }
PB_D int pop_cnt(uint32_t x) { return __popc(x); }
#else
inline U4 ld_u4(const U4* p) { return *p; }
inline V4 ld_v4(const V4* p) { return *p; }
inline int hi_bit(uint32_t x) { return 31 - __builtin_clz(x); }
inline int lo_bit(uint32_t x) { return __builtin_ctz(x); }
inline int pop_cnt(uint32_t x) { return __builtin_popcount(x); }
#endif

#ifndef PB_SMEM_STACK
#define PB_SMEM_STACK 12
#endif
#define PB_LOCAL_STACK 52
// PB_COLD_SMEM (default): the per-ray state the node loop never touches - direction, u, v, object of the best hit -
// lives in three more 64-bit shared-memory entries behind the lane's stack instead of in registers, and path rays'
// constant [tmin, tmax] are compile-time constants (Traversal<.., FIXED>).  At 64 registers k_extend then has no
// compiler spills (it had 24 / 32 B in the node loop): 545.4 against 530.7 Msamples/s on the same box (+2.8 %,
// profiles/r02_extend.md, call J).  Capping at 56 registers for a 9th resident CTA still spills 40 / 48 B and loses
// (519 Msamples/s), so kTraceMinBlocks stays 8.  Results are identical either way; 0 = the register-only layout.
#ifndef PB_COLD_SMEM
#define PB_COLD_SMEM 1
#endif
#define PB_COLD_ENTRIES (PB_COLD_SMEM ? 3 : 0)
constexpr float kPathTmin = 0.001f, kPathTmax = 1e16f;   // main.cu:77-82,158-163

struct U2 { uint32_t x, y; };

// Per-lane traversal stack: shared-memory head + local-memory tail.  The tail array is owned by the
// kernel (not a member) so that `sp` is scalar-replaced into a register.
struct LaneStack {
    U2* smem;          // this lane's first entry; entries are `stride` apart
    U2* local;         // PB_LOCAL_STACK entries
    uint32_t stride;
    int sp;
    PB_HD void init(U2* s, uint32_t st, U2* tail) { smem = s; stride = st; local = tail; sp = 0; }
    PB_HD void push(U2 v) {
        if (sp < PB_SMEM_STACK) smem[sp * stride] = v;
        else local[sp - PB_SMEM_STACK] = v;
        ++sp;
    }
    PB_HD U2 pop() {
        --sp;
        return sp < PB_SMEM_STACK ? smem[sp * stride] : local[sp - PB_SMEM_STACK];
    }
    // cold entries behind the stack head (PB_COLD_SMEM): k = 0 (d.x, d.y), 1 (d.z, u), 2 (v, object)
    PB_HD U2& cold(int k) { return smem[(PB_SMEM_STACK + k) * stride]; }
};

// Quantised plane byte -> float without a conversion instruction: one PRMT drops byte J of w into
// byte 1 of 0x47000000, giving the float 32768 + q exactly; the bias is folded into the plane offset
// (b' = b - 32768 a, rounded DOWN for near planes and UP for far planes, so the node box only ever
// grows, by at most a * 2^-8 = 0.4 % of one quantisation step).  q in byte 0 would need a bias of
// 2^23 (half a step of rounding error: not acceptable), and the compiler's own (float)(w >> 8J & 255)
// is a quarter-rate I2F.U8.  The constant 0x47000000 arrives as a kernel parameter (TraceScene::k47) so
// that it sits in a register and the selector is the PRMT immediate; as a literal, ptxas makes the
// constant the immediate and re-materialises the selector with one extra move per PRMT (measured).
// PB_BYTE_DP4A (default): the same bytes through the dot-product unit instead.  IDP.4A.U8.U8 (w, 128 << 8J, k47)
// = 0x47000000 + 128 q = the float 32768 + q / 2, so the plane scale is doubled (plane_scale) and the folded
// bias is 65536 a (error a * 2^-7, still < 1 % of a step).  IDP issues on the FMA-heavy pipe, which the node
// test leaves idle (ncu v6: ALU 60 % busy, FMA 22 %); 48 PRMT per node were 30 % of the node test's ALU work.
#ifndef PB_BYTE_DP4A
#define PB_BYTE_DP4A 1
#endif
// byte J of w, zero-extended: one PRMT (the compiler's shift + mask are two ALU instructions)
#if defined(__CUDA_ARCH__)
template<int J> PB_D uint32_t byte_u(uint32_t w) { return __byte_perm(w, 0u, 0x4440u | J); }
#else
template<int J> inline uint32_t byte_u(uint32_t w) { return (w >> (8 * J)) & 0xffu; }
#endif
#if defined(__CUDA_ARCH__) && PB_BYTE_DP4A
template<int J> PB_D float byte_f(uint32_t w, uint32_t k47) { return __uint_as_float(__dp4a(w, 128u << (8 * J), k47)); }
PB_D float plane_scale(float s) { return __fmul_rn(s, 2.0f); }
PB_D float bias_lo(float a, float b) { return __fmaf_rd(-32768.0f, a, b); }
PB_D float bias_hi(float a, float b) { return __fmaf_ru(-32768.0f, a, b); }
#elif defined(__CUDA_ARCH__)
template<int J> PB_D float byte_f(uint32_t w, uint32_t k47) { return __uint_as_float(__byte_perm(w, k47, 0x7404u | (J << 4))); }
PB_D float plane_scale(float s) { return s; }
PB_D float bias_lo(float a, float b) { return __fmaf_rd(-32768.0f, a, b); }
PB_D float bias_hi(float a, float b) { return __fmaf_ru(-32768.0f, a, b); }
#else
template<int J> inline float byte_f(uint32_t w, uint32_t) { return (float)((w >> (8 * J)) & 0xffu); }
inline float plane_scale(float s) { return s; }
inline float bias_lo(float, float b) { return b; }
inline float bias_hi(float, float b) { return b; }
#endif

#ifndef PB_NODE_FMA2
#define PB_NODE_FMA2 0   // A/B on C4 (scripts/ab_node_test.sh): scalar FFMA 481.3 vs FFMA2 479.0 Msamples/s (the packing MOVs cost what the pairs save)
#endif
#if defined(__CUDA_ARCH__)
// (q0, q1) * a + b on both halves with one FFMA2 (fma.rn.f32x2, sm_100+; ptxas broadcasts the scalar operands)
PB_D float2 fma2(float q0, float q1, float a, float b) {
    unsigned long long q, aa, bb, r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(q) : "f"(q0), "f"(q1));
    asm("mov.b64 %0, {%1, %1};" : "=l"(aa) : "f"(a));
    asm("mov.b64 %0, {%1, %1};" : "=l"(bb) : "f"(b));
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(q), "l"(aa), "l"(bb));
    float2 o;
    asm("mov.b64 {%0, %1}, %2;" : "=f"(o.x), "=f"(o.y) : "l"(r));
    return o;
}
#endif

// Per-ray constants of the octant permutation, packed in one register: byte 0 = octinv, bytes 1..3 = the
// delta-swap masks of the three stages (0x55 / 0x33 / 0x0f when the matching bit of octinv is set, else 0).
// Inner-child hits are collected in slot order (bit 24 + slot) and moved to bit 24 + (slot ^ octinv) by three
// conditional delta swaps IN the high byte of the hit word, so that hi_bit() visits the children front to
// back (Ylitie et al. 2017, octant order).
PB_HD uint32_t oct_perm(uint32_t octinv) {
    return octinv | ((octinv & 1u) ? 0x5500u : 0u) | ((octinv & 2u) ? 0x330000u : 0u) | ((octinv & 4u) ? 0x0f000000u : 0u);
}
#if defined(__CUDA_ARCH__)
template<int B> PB_D uint32_t byte_to_top(uint32_t w) { return __byte_perm(w, 0u, 0x0444u | (B << 12)); }   // byte B -> byte 3, rest 0
// bytes (a0, a1, a2, b3) / (a.B0, b.B1, b.B2 ...) helpers for the group packing below
PB_D uint32_t pack_p(uint32_t hits, uint32_t pm) { return __byte_perm(hits, pm, 0x6210); }        // hits[0..2], pm[2]
PB_D uint32_t pack_g(uint32_t hits, uint32_t pm, uint32_t n0w) {                                   // imask, pm[0], pm[1], hits[3]
    return __byte_perm(__byte_perm(n0w, pm, 0x0543), hits, 0x7210);
}
PB_D uint32_t unpack_pm(uint32_t gy, uint32_t py) { return __byte_perm(gy, py, 0x4721); }   // byte 3 = junk: callers mask below bit 24
#else
template<int B> inline uint32_t byte_to_top(uint32_t w) { return ((w >> (8 * B)) & 0xffu) << 24; }
inline uint32_t pack_p(uint32_t hits, uint32_t pm) { return (hits & 0x00ffffffu) | ((pm >> 16) << 24); }
inline uint32_t pack_g(uint32_t hits, uint32_t pm, uint32_t n0w) { return (hits & 0xff000000u) | ((pm & 0xffffu) << 8) | (n0w >> 24); }
inline uint32_t unpack_pm(uint32_t gy, uint32_t py) { return ((gy >> 8) & 0xffffu) | ((py >> 24) << 16); }
#endif
PB_HD uint32_t delta_swap(uint32_t x, uint32_t m, int k) {
    const uint32_t t = ((x >> k) ^ x) & m;
    return x ^ t ^ (t << k);
}

// hit word of one node for this ray: bits 24..31 = inner children hit, in octant order; bits 0..23 = leaf slots hit
// (3 bits per slot, masked by the node's pmask, i.e. one bit per primitive in slot order)
PB_HD uint32_t node_hits(const U4& n0, const U4& n1, const U4& n2, const U4& n3, const U4& n4, V3 o, V3 id,
                         uint32_t op, float tmin, float tmax, uint32_t k47) {
    // n0 = px py pz [ex ey ez imask]; n1 = child_base prim_base pmask spare
    // n2 = qlox[0..7] qloy[0..7]; n3 = qloz[0..7] qhix[0..7]; n4 = qhiy[0..7] qhiz[0..7]
    const uint32_t e = n0.w;
    const float sx = u2f((e & 0xffu) << 23), sy = u2f(((e >> 8) & 0xffu) << 23), sz = u2f(((e >> 16) & 0xffu) << 23);
    const float ax = f_mul(plane_scale(sx), id.x), ay = f_mul(plane_scale(sy), id.y), az = f_mul(plane_scale(sz), id.z);
    const float bx = f_mul(f_sub(u2f(n0.x), o.x), id.x);
    const float by = f_mul(f_sub(u2f(n0.y), o.y), id.y);
    const float bz = f_mul(f_sub(u2f(n0.z), o.z), id.z);
    const float bnx = bias_lo(ax, bx), bfx = bias_hi(ax, bx);
    const float bny = bias_lo(ay, by), bfy = bias_hi(ay, by);
    const float bnz = bias_lo(az, bz), bfz = bias_hi(az, bz);
    const bool nx = id.x < 0.f, ny = id.y < 0.f, nz = id.z < 0.f;
    uint32_t acc = 0;
#pragma unroll
    for (int half = 0; half < 2; ++half) {
        const uint32_t lx = half ? n2.y : n2.x, ly = half ? n2.w : n2.z, lz = half ? n3.y : n3.x;
        const uint32_t hx = half ? n3.w : n3.z, hy = half ? n4.y : n4.x, hz = half ? n4.w : n4.z;
        const uint32_t qnx = nx ? hx : lx, qfx = nx ? lx : hx;
        const uint32_t qny = ny ? hy : ly, qfy = ny ? ly : hy;
        const uint32_t qnz = nz ? hz : lz, qfz = nz ? lz : hz;
        // slot S hit: its three leaf bits and its inner bit, one constant; pmask / imask keep what exists
#define PB_SLOT_BITS(S) ((7u << (3 * (S))) | (1u << (24 + (S))))
#if defined(__CUDA_ARCH__) && PB_NODE_FMA2
        // two children per instruction: sm_100 packed fp32 FMA (FFMA2, one issue slot for both halves, scalar
        // operands broadcast; same rn rounding as FFMA, so the hit word is bit-identical).  The kernel is bound by
        // issue slots, not by the FMA pipe.
#define PB_CHILD2(J, K)                                                                                          \
        {                                                                                                        \
            const float2 tnx = fma2(byte_f<J>(qnx, k47), byte_f<K>(qnx, k47), ax, bnx);                          \
            const float2 tfx = fma2(byte_f<J>(qfx, k47), byte_f<K>(qfx, k47), ax, bfx);                          \
            const float2 tny = fma2(byte_f<J>(qny, k47), byte_f<K>(qny, k47), ay, bny);                          \
            const float2 tfy = fma2(byte_f<J>(qfy, k47), byte_f<K>(qfy, k47), ay, bfy);                          \
            const float2 tnz = fma2(byte_f<J>(qnz, k47), byte_f<K>(qnz, k47), az, bnz);                          \
            const float2 tfz = fma2(byte_f<J>(qfz, k47), byte_f<K>(qfz, k47), az, bfz);                          \
            const float tn0 = fmaxf(fmaxf(tnx.x, tny.x), fmaxf(tnz.x, tmin)), tn1 = fmaxf(fmaxf(tnx.y, tny.y), fmaxf(tnz.y, tmin)); \
            const float tf0 = fminf(fminf(tfx.x, tfy.x), fminf(tfz.x, tmax)), tf1 = fminf(fminf(tfx.y, tfy.y), fminf(tfz.y, tmax)); \
            if (tn0 <= tf0) acc |= PB_SLOT_BITS(4 * half + J);                                                   \
            if (tn1 <= tf1) acc |= PB_SLOT_BITS(4 * half + K);                                                   \
        }
        PB_CHILD2(0, 1) PB_CHILD2(2, 3)
#undef PB_CHILD2
#else
#define PB_CHILD(J)                                                                                              \
        {                                                                                                        \
            const float tnx = f_fma(byte_f<J>(qnx, k47), ax, bnx), tfx = f_fma(byte_f<J>(qfx, k47), ax, bfx);   \
            const float tny = f_fma(byte_f<J>(qny, k47), ay, bny), tfy = f_fma(byte_f<J>(qfy, k47), ay, bfy);   \
            const float tnz = f_fma(byte_f<J>(qnz, k47), az, bnz), tfz = f_fma(byte_f<J>(qfz, k47), az, bfz);   \
            const float tn = fmaxf(fmaxf(tnx, tny), fmaxf(tnz, tmin));                                           \
            const float tf = fminf(fminf(tfx, tfy), fminf(tfz, tmax));                                           \
            if (tn <= tf) acc |= PB_SLOT_BITS(4 * half + J);                                                     \
        }
        PB_CHILD(0) PB_CHILD(1) PB_CHILD(2) PB_CHILD(3)
#undef PB_CHILD
#endif
#undef PB_SLOT_BITS
    }
    acc &= n1.z | (n0.w & 0xff000000u);                       // pmask | imask << 24
    acc = delta_swap(acc, byte_to_top<1>(op), 1);
    acc = delta_swap(acc, byte_to_top<2>(op), 2);
    acc = delta_swap(acc, op & 0xff000000u, 4);
    return acc;
}

// Resumable traversal of one ray: begin() once, then step() until it returns true.  One step = at
// most one node visit (8 child boxes) and, when the warp decides so, the primitives found so far.
// Kernels interleave steps of 32 rays per warp and refill finished lanes between steps
// (pb_trace_loop.cuh).  On the device step() is called by ALL 32 lanes (alive or not): a lane that
// has found primitives waits (takes no further nodes, so the visiting order is unchanged) until at
// least `prim_batch` lanes of the warp hold primitives or no lane can advance through nodes any
// more; the primitive tests then run with many lanes instead of the 2.6 / 32 measured before
// (profiles/r01_extend.md).
template<bool ANY, bool COUNT, bool FIXED = false>
struct Traversal {
    V3 o, d;
    float tmin_, tmax_;   // tmax = the ray's own bound (ties on t need it; best.t only orders); unused (and therefore not
                          // kept in registers) when FIXED: path rays always run over [kPathTmin, kPathTmax]
    PB_HD float tmin() const { return FIXED ? kPathTmin : tmin_; }
    PB_HD float tmax() const { return FIXED ? kPathTmax : tmax_; }
    V3 id;              // 1 / d (RayPre::id); RayPre::noid is re-derived in the primitive phase (3 registers less)
    uint32_t op;        // oct_perm(octinv): octinv in the low byte
    U2 G, P;            // pending node group / primitive group.  G.y = inner hits << 24 | pmask[0..15] << 8 | imask,
                        // P.y = pmask[16..23] << 24 | primitive hits: the pmask of the node P came from rides in the spare
                        // bits (primitive of hit bit i = P.x + popc(pmask & ((1 << i) - 1))); G on the stack carries stale
                        // pmask bits that nothing reads
    Hit best;

    PB_HD void begin(const TraceScene& sc, V3 o_, V3 d_, float tmin_in, float tmax_in, LaneStack& st) {
        o = o_; d = d_; tmin_ = tmin_in; tmax_ = tmax_in;
        best.t = tmax(); best.u = 0.f; best.v = 0.f; best.prim = 0xFFFFFFFFu; best.obj = 0xFFFFFFFFu;
        id = ray_pre(o, d_).id;
        const uint32_t oct = (d_.x < 0.f ? 4u : 0u) | (d_.y < 0.f ? 2u : 0u) | (d_.z < 0.f ? 1u : 0u);
        if (PB_COLD_SMEM) {
            st.cold(0) = {f2u(d_.x), f2u(d_.y)};
            st.cold(1) = {f2u(d_.z), 0u};
            st.cold(2) = {0u, 0xFFFFFFFFu};
        }
        op = oct_perm(7u - oct);
        G = {0u, sc.n_nodes ? 0x80000000u : 0u};
        P = {0u, 0u};
        st.sp = 0;
    }
    PB_HD bool finished_at_begin() const { return (G.y & 0xff000000u) == 0u; }
    // the result of a finished ray (cold members come back from shared memory)
    PB_HD Hit result(LaneStack& st) const {
        Hit h = best;
        if (PB_COLD_SMEM && !ANY) { const U2 c1 = st.cold(1), c2 = st.cold(2); h.u = u2f(c1.y); h.v = u2f(c2.x); h.obj = c2.y; }
        return h;
    }

    // returns true when the ray is finished (best holds the result); `alive` = this lane holds a ray
    PB_HD bool step(const TraceScene& sc, LaneStack& st, TraceCounters* cnt, bool alive) {
        if (alive && (P.y & 0x00ffffffu) == 0u) {   // invariant: an alive lane without primitives has inner hits in G
            const int bit = hi_bit(G.y);
            const uint32_t imask = G.y & 0xffu;
            G.y &= ~(1u << bit);
            if (G.y & 0xff000000u) st.push(G);
            const uint32_t slot = (((uint32_t)bit - 24u) ^ op) & 7u;
            const uint32_t rel = (uint32_t)pop_cnt(imask & ((1u << slot) - 1u));
            const U4* np = sc.nodes + 5ull * (G.x + rel);
            const U4 n0 = ld_u4(np), n1 = ld_u4(np + 1), n2 = ld_u4(np + 2), n3 = ld_u4(np + 3), n4 = ld_u4(np + 4);
            if (COUNT) cnt->nodes++;
            const uint32_t hits = node_hits(n0, n1, n2, n3, n4, o, id, op, tmin(), best.t, sc.k47);
            G.x = n1.x;
            G.y = pack_g(hits, n1.z, n0.w);
            P.x = n1.y;
            P.y = pack_p(hits, n1.z);
        }
        const bool has_p = alive && (P.y & 0x00ffffffu) != 0u;
#if defined(__CUDA_ARCH__)
        const unsigned pm = __ballot_sync(0xFFFFFFFFu, has_p);
        const bool do_prims = (uint32_t)__popc(pm) >= sc.prim_batch || pm == __ballot_sync(0xFFFFFFFFu, alive);
#else
        const bool do_prims = true;
#endif
        if (has_p && do_prims) {
            uint32_t iters = sc.prim_iters;
            const RayPre pre = {id, {-f_mul(o.x, id.x), -f_mul(o.y, id.y), -f_mul(o.z, id.z)}};   // == ray_pre(o, d), bit for bit
            V3 dd = d;
            if (PB_COLD_SMEM) { const U2 c0 = st.cold(0), c1 = st.cold(1); dd = {u2f(c0.x), u2f(c0.y), u2f(c1.x)}; }
            do {
                const int i = lo_bit(P.y);
                P.y &= P.y - 1u;
                const V4* pp = sc.prims + 3ull * (P.x + (uint32_t)pop_cnt(unpack_pm(G.y, P.y) & ((1u << i) - 1u)));
                PrimRec r;
                r.a = ld_v4(pp); r.b = ld_v4(pp + 1); r.c = ld_v4(pp + 2);
                if (COUNT) { if (f2u(r.c.w) == PRIM_TRI) cnt->tris++; else cnt->spheres++; }
                float t, u, v;
                if (prim_accept(r, sc.objs, o, dd, pre, tmin(), tmax(), best.t, sc.delta, t, u, v, PB_TRI_TEST_FIRST && !ANY)) {
                    const uint32_t id = f2u(r.a.w);
                    if (ANY) { best.t = t; best.u = u; best.v = v; best.prim = id; best.obj = f2u(r.b.w); P.y = 0u; return true; }
                    if (t < best.t || (t == best.t && id < best.prim)) {
                        best.t = t; best.prim = id;
                        if (PB_COLD_SMEM) { st.cold(1).y = f2u(u); st.cold(2) = {f2u(v), f2u(r.b.w)}; }
                        else { best.u = u; best.v = v; best.obj = f2u(r.b.w); }
                    }
                }
            } while ((P.y & 0x00ffffffu) && --iters);
        }
        if (alive && (P.y & 0x00ffffffu) == 0u && !(G.y & 0xff000000u)) {
            if (st.sp == 0) return true;
            G = st.pop();
        }
        return false;
    }
};

template<bool ANY, bool COUNT>
PB_HD void traverse(const TraceScene& sc, V3 o, V3 d, float tmin, float tmax, Hit& best, LaneStack& st, TraceCounters* cnt) {
    Traversal<ANY, COUNT> tr;
    tr.begin(sc, o, d, tmin, tmax, st);
    if (!tr.finished_at_begin())
        while (!tr.step(sc, st, cnt, true)) {}
    best = tr.result(st);
}

}  // namespace pb
