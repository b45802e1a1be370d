// pb_collapse.cuh — binary radix tree -> compressed 8-wide BVH, one BFS level at a time.
//   plan(w)  : choose up to 8 children of wide node w (greedy opening by surface area), count its
//              inner children and leaf primitives            -> exclusive scans give the layout
//   emit(w)  : octant slot assignment, conservative quantisation, Node8 write, primitive copy into
//              leaf order, next-level work items
// Deterministic: every offset comes from a prefix sum, nothing from atomics.
#pragma once
#include "pb_build.cuh"

namespace pb {

struct WorkItem { uint32_t bin_ref; uint32_t wide_index; };

struct CollapseArrays {
    const WorkItem* work;        // this level
    uint32_t n_work;
    uint32_t* plan;              // 8 refs per work item (kNoRef = empty)
    uint32_t* cnt_inner;         // per work item (scanned in place -> offsets)
    uint32_t* cnt_prims;
    WorkItem* next_work;         // next level
    uint32_t next_wide_base;     // wide index of the first node of the next level
    uint32_t prim_cursor;        // leaf-order position of this level's first primitive
    Node8* nodes;
    PrimRec* prims_out;
};

PB_HD float half_area(V3 lo, V3 hi) {
    V3 e = hi - lo;
    return e.x * e.y + e.y * e.z + e.z * e.x;
}
PB_HD bool ref_is_inner_wide(const BuildArrays& b, uint32_t ref) {  // becomes an inner child of the wide node
    return !(ref & kLeafFlag) && ref_count(b, ref) > kLeafMax;
}

PB_HD void collapse_plan(const BuildArrays& b, const CollapseArrays& c, uint32_t w) {
    uint32_t ch[8];
    int n = 0;
    const uint32_t root = c.work[w].bin_ref;
    if (root & kLeafFlag) { ch[n++] = root; }            // single-primitive scene
    else { ch[n++] = b.left[root]; ch[n++] = b.right[root]; }
    // pass 0 opens nodes that would become inner children, pass 1 opens small subtrees so that
    // leaf slots hold fewer primitives with tighter boxes
    for (int pass = 0; pass < 2; ++pass) {
        while (n < 8) {
            int best = -1;
            float best_area = -1.f;
            for (int k = 0; k < n; ++k) {
                if (ch[k] & kLeafFlag) continue;
                const bool big = ref_count(b, ch[k]) > kLeafMax;
                if ((pass == 0) != big) continue;
                float a = half_area(b.ilo[ch[k]], b.ihi[ch[k]]);
                if (a > best_area) { best_area = a; best = k; }
            }
            if (best < 0) break;
            const uint32_t r = ch[best];
            ch[best] = b.left[r];
            ch[n++] = b.right[r];
        }
    }
    uint32_t ni = 0, np = 0;
    for (int k = 0; k < 8; ++k) {
        c.plan[8 * w + k] = k < n ? ch[k] : kNoRef;
        if (k < n) { if (ref_is_inner_wide(b, ch[k])) ++ni; else np += ref_count(b, ch[k]); }
    }
    c.cnt_inner[w] = ni;
    c.cnt_prims[w] = np;
}

// smallest biased exponent e (1..254) with ceil(ext / 2^(e-127)) <= 255
PB_HD uint32_t quant_exponent(float ext) {
    if (!(ext > 0.f)) return 1u;
    int ex;
    frexpf(ext / 255.0f, &ex);          // ext/255 = m * 2^ex, m in [0.5, 1)  ->  2^ex >= ext/255
    int e = ex + 127;
    if (e < 1) e = 1;
    if (e > 254) e = 254;
    return (uint32_t)e;
}

PB_HD void collapse_emit(const BuildArrays& b, const CollapseArrays& c, uint32_t w, uint32_t off_inner, uint32_t off_prims) {
    uint32_t ch[8];
    V3 clo[8], chi[8];
    int n = 0;
    for (int k = 0; k < 8; ++k) {
        uint32_t r = c.plan[8 * w + k];
        if (r == kNoRef) break;
        ch[n] = r;
        ref_box(b, r, clo[n], chi[n]);
        ++n;
    }
    V3 lo = clo[0], hi = chi[0];
    for (int k = 1; k < n; ++k) {
        lo = {fminf(lo.x, clo[k].x), fminf(lo.y, clo[k].y), fminf(lo.z, clo[k].z)};
        hi = {fmaxf(hi.x, chi[k].x), fmaxf(hi.y, chi[k].y), fmaxf(hi.z, chi[k].z)};
    }
    // ---- octant slot assignment: greedy max of dot(centre offset, octant sign); slot bit set = + side
    const V3 ctr = (lo + hi) * 0.5f;
    int slot_of[8];
    bool slot_used[8] = {false, false, false, false, false, false, false, false};
    bool child_done[8] = {false, false, false, false, false, false, false, false};
    for (int it = 0; it < n; ++it) {
        int bk = -1, bs = -1;
        float bc = -3.0e38f;
        for (int k = 0; k < n; ++k) {
            if (child_done[k]) continue;
            V3 o = (clo[k] + chi[k]) * 0.5f - ctr;
            for (int s = 0; s < 8; ++s) {
                if (slot_used[s]) continue;
                float cost = ((s & 4) ? o.x : -o.x) + ((s & 2) ? o.y : -o.y) + ((s & 1) ? o.z : -o.z);
                if (cost > bc) { bc = cost; bk = k; bs = s; }
            }
        }
        slot_of[bk] = bs; slot_used[bs] = true; child_done[bk] = true;
    }
    // ---- quantisation grid
    Node8 nd;
    nd.px = lo.x; nd.py = lo.y; nd.pz = lo.z;
    uint32_t ex = quant_exponent(hi.x - lo.x), ey = quant_exponent(hi.y - lo.y), ez = quant_exponent(hi.z - lo.z);
    // exact (double) check that every child maximum fits in 8 bits; bump the exponent otherwise
    for (int axis = 0; axis < 3; ++axis) {
        uint32_t& e = axis == 0 ? ex : (axis == 1 ? ey : ez);
        const double p = axis == 0 ? lo.x : (axis == 1 ? lo.y : lo.z);
        const double h = axis == 0 ? hi.x : (axis == 1 ? hi.y : hi.z);
        while (e < 254u && ceil((h - p) / (double)u2f(e << 23)) > 255.0) ++e;
    }
    nd.ex = (uint8_t)ex; nd.ey = (uint8_t)ey; nd.ez = (uint8_t)ez;
    const double sx = u2f(ex << 23), sy = u2f(ey << 23), sz = u2f(ez << 23);
    uint32_t imask = 0;
    for (int s = 0; s < 8; ++s) {
        nd.meta[s] = 0;
        nd.qlox[s] = nd.qloy[s] = nd.qloz[s] = 0;
        nd.qhix[s] = nd.qhiy[s] = nd.qhiz[s] = 0;
    }
    // children in slot order: inner children get consecutive wide indices, leaves consecutive prims
    uint32_t inner_rank = 0, prim_off = 0;
    for (int s = 0; s < 8; ++s) {
        int k = -1;
        for (int q = 0; q < n; ++q) if (slot_of[q] == s) k = q;
        if (k < 0) continue;
        auto qlo = [](float v, double p, double sc) { double q = floor(((double)v - p) / sc); return (uint8_t)(q < 0.0 ? 0.0 : (q > 255.0 ? 255.0 : q)); };
        auto qhi = [](float v, double p, double sc) { double q = ceil(((double)v - p) / sc); return (uint8_t)(q < 0.0 ? 0.0 : (q > 255.0 ? 255.0 : q)); };
        nd.qlox[s] = qlo(clo[k].x, lo.x, sx); nd.qloy[s] = qlo(clo[k].y, lo.y, sy); nd.qloz[s] = qlo(clo[k].z, lo.z, sz);
        nd.qhix[s] = qhi(chi[k].x, lo.x, sx); nd.qhiy[s] = qhi(chi[k].y, lo.y, sy); nd.qhiz[s] = qhi(chi[k].z, lo.z, sz);
        if (ref_is_inner_wide(b, ch[k])) {
            imask |= 1u << s;
            nd.meta[s] = (uint8_t)(0x20u | (24u + (uint32_t)s));
            c.next_work[off_inner + inner_rank] = {ch[k], c.next_wide_base + off_inner + inner_rank};
            ++inner_rank;
        } else {
            const uint32_t cnt = ref_count(b, ch[k]);
            const uint32_t first = (ch[k] & kLeafFlag) ? (ch[k] & ~kLeafFlag) : b.first[ch[k]];
            nd.meta[s] = (uint8_t)((((1u << cnt) - 1u) << 5) | prim_off);
            for (uint32_t j = 0; j < cnt; ++j)
                c.prims_out[c.prim_cursor + off_prims + prim_off + j] = b.prims_in[b.order[first + j]];
            prim_off += cnt;
        }
    }
    nd.imask = (uint8_t)imask;
    nd.child_base = c.next_wide_base + off_inner;
    nd.prim_base = c.prim_cursor + off_prims;
    c.nodes[c.work[w].wide_index] = nd;
}

}  // namespace pb
