// ORACLE — test infrastructure, not product code.
//
// orc_cwbvh.h — the oracle's own reading of the acceleration structure the product committed (pb_export_bvh:
// 80-byte compressed 8-wide nodes + 48-byte primitive records in leaf order, layout documented in DESIGN.md §3 and
// pupiloptixlab_b200/csrc/pb_types.h) and an ORDERED, t-CULLED traversal of it.  SURVEY §8(d) makes the visit counts
// of exactly this traversal the source of the algorithmic bytes per ray (B_ray = 32 + 16 + 80 n_nodes + 48 n_tri +
// 32 n_sph): they depend on the committed tree and the ray only, not on how the GPU kernel schedules its lanes,
// batches primitive tests or leaves stale entries on its stack.
//
//   * a node is fetched (counted) when it is popped and its entry distance is <= the closest accepted t so far;
//   * its child boxes are decoded exactly (origin + q * 2^e in double) and slab-tested against [tmin, t_best];
//   * leaf slots that are hit are processed first, nearest entry first, each re-checked against the current t_best;
//     every primitive record of a processed slot counts as one test (the record is fetched whatever the test says);
//   * inner children that are hit are pushed far to near, so the nearest is visited next (depth first).
// The primitive test is the oracle's own acceptance (orc_scene.h) on the primitive's GLOBAL id, so the hit must equal
// the oracle's binary-BVH / brute-force result bit for bit — checked by the callers.
#pragma once
#include <cstring>

#include "orc_scene.h"

namespace orc {

struct CwCounts { uint64_t nodes = 0, tris = 0, spheres = 0; };

struct CwBvh {
    std::vector<uint8_t> nodes;      // 80 bytes each
    std::vector<float> prims;        // 12 floats each (3 x float4), leaf order
    uint64_t n_nodes = 0, n_prims = 0, n_leaf_slots = 0;
    bool empty() const { return n_nodes == 0; }
};

struct CwNode {   // decoded view of one 80-byte record
    double p[3], scale[3];
    uint32_t imask, child_base, prim_base, pmask;
    const uint8_t* q;   // 48 bytes: qlox qloy qloz qhix qhiy qhiz, 8 each
};
inline CwNode cw_decode(const uint8_t* r) {
    CwNode n;
    float f[3];
    std::memcpy(f, r, 12);
    for (int k = 0; k < 3; ++k) {
        n.p[k] = f[k];
        uint32_t bits = (uint32_t)r[12 + k] << 23;
        float s; std::memcpy(&s, &bits, 4);
        n.scale[k] = s;
    }
    n.imask = r[15];
    std::memcpy(&n.child_base, r + 16, 4);
    std::memcpy(&n.prim_base, r + 20, 4);
    std::memcpy(&n.pmask, r + 24, 4);
    n.q = r + 32;
    return n;
}
inline uint32_t cw_leaf_count(uint32_t pmask, int s) { return (uint32_t)__builtin_popcount((pmask >> (3 * s)) & 7u); }

// accept(prim id, t, u, v) is the oracle's acceptance test of the global primitive `id`; kind(id) says 0 tri / 1 sphere
template<bool ANY, class Accept>
inline Hit cw_trace(const CwBvh& bvh, const Ray& r, CwCounts& cnt, Accept accept) {
    Hit best{r.tmax, 0.f, 0.f, kMiss};
    if (bvh.empty()) return best;
    const double o[3] = {r.o.x, r.o.y, r.o.z};
    double id[3];
    { const RayPre pre = ray_pre(r); id[0] = pre.id.x; id[1] = pre.id.y; id[2] = pre.id.z; }   // same clamped reciprocal as every other traversal
    struct Entry { double tn; uint32_t node; };
    Entry stack[256];
    int sp = 0;
    stack[sp++] = {0.0, 0u};
    while (sp) {
        const Entry e = stack[--sp];
        if (e.tn > (double)best.t) continue;             // t-culled: ties must still be visited
        ++cnt.nodes;
        const CwNode n = cw_decode(bvh.nodes.data() + 80ull * e.node);
        double tn_s[8];
        bool hit_s[8];
        for (int s = 0; s < 8; ++s) {
            hit_s[s] = false;
            const bool inner = (n.imask >> s) & 1u;
            if (!inner && cw_leaf_count(n.pmask, s) == 0u) continue;
            double tn = r.tmin, tf = best.t;
            for (int a = 0; a < 3; ++a) {
                const double lo = n.p[a] + n.scale[a] * n.q[8 * a + s], hi = n.p[a] + n.scale[a] * n.q[24 + 8 * a + s];
                const double t0 = (lo - o[a]) * id[a], t1 = (hi - o[a]) * id[a];
                tn = std::max(tn, std::min(t0, t1));
                tf = std::min(tf, std::max(t0, t1));
            }
            if (tn <= tf) { hit_s[s] = true; tn_s[s] = tn; }
        }
        // leaf slots, nearest first
        for (;;) {
            int bs = -1;
            for (int s = 0; s < 8; ++s)
                if (hit_s[s] && !((n.imask >> s) & 1u) && (bs < 0 || tn_s[s] < tn_s[bs])) bs = s;
            if (bs < 0) break;
            hit_s[bs] = false;
            if (tn_s[bs] > (double)best.t) continue;
            const uint32_t first = n.prim_base + (uint32_t)__builtin_popcount(n.pmask & ((1u << (3 * bs)) - 1u));
            for (uint32_t j = 0; j < cw_leaf_count(n.pmask, bs); ++j) {
                const float* pr = bvh.prims.data() + 12ull * (first + j);
                uint32_t gid, kind;
                std::memcpy(&gid, pr + 3, 4); std::memcpy(&kind, pr + 11, 4);
                if (kind == 0u) ++cnt.tris; else ++cnt.spheres;
                float t, u, v;
                if (accept(gid, t, u, v)) {
                    if (ANY) return Hit{t, u, v, gid};
                    if (t < best.t || (t == best.t && gid < best.prim)) best = {t, u, v, gid};
                }
            }
        }
        // inner children, pushed far to near
        int order[8], m = 0;
        for (int s = 0; s < 8; ++s) if (hit_s[s]) order[m++] = s;
        for (int a = 1; a < m; ++a) { int x = order[a], b = a; while (b > 0 && tn_s[order[b - 1]] < tn_s[x]) { order[b] = order[b - 1]; --b; } order[b] = x; }
        for (int a = 0; a < m; ++a) {
            const int s = order[a];
            if (sp < 256) stack[sp++] = {tn_s[s], n.child_base + (uint32_t)__builtin_popcount(n.imask & ((1u << s) - 1u))};
        }
    }
    return best;
}

}  // namespace orc
