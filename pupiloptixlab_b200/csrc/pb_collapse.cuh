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

// Plan entries ("slot refs"): an inner child is the index of a binary internal node (< 2^29, no
// flag); a leaf slot is a RANGE of the sorted primitive order: kLeafFlag | count << 29 | first,
// count = 1..3.
PB_HD uint32_t range_ref(uint32_t first, uint32_t count) { return kLeafFlag | (count << 29) | first; }
PB_HD uint32_t range_first(uint32_t ref) { return ref & 0x1fffffffu; }
PB_HD uint32_t range_count(uint32_t ref) { return (ref >> 29) & 3u; }
PB_HD void slot_box(const BuildArrays& b, uint32_t ref, V3& lo, V3& hi) {
    if (!(ref & kLeafFlag)) { lo = b.ilo[ref]; hi = b.ihi[ref]; return; }
    const uint32_t first = range_first(ref), cnt = range_count(ref);
    uint32_t p = b.order[first];
    lo = b.blo[p]; hi = b.bhi[p];
    for (uint32_t j = 1; j < cnt; ++j) {
        p = b.order[first + j];
        lo = {fminf(lo.x, b.blo[p].x), fminf(lo.y, b.blo[p].y), fminf(lo.z, b.blo[p].z)};
        hi = {fmaxf(hi.x, b.bhi[p].x), fmaxf(hi.y, b.bhi[p].y), fmaxf(hi.z, b.bhi[p].z)};
    }
}

PB_HD void collapse_plan(const BuildArrays& b, const CollapseArrays& c, uint32_t w) {
    uint32_t ch[8];
    int n = 0;
    const uint32_t root = c.work[w].bin_ref;
    uint32_t ni = 0, np = 0;
    if (root & kLeafFlag) {                                    // single-primitive scene
        ch[n++] = range_ref(root & ~kLeafFlag, 1u);
    } else if (ref_count(b, root) <= b.bottom_max) {
        // bottom node: the whole subtree (a contiguous run of the Morton order) in one wide node,
        // cut into <= 8 runs of ceil(count / 8) <= 3 primitives; no sparse inner nodes below it
        const uint32_t cnt = ref_count(b, root), first = b.first[root], g = (cnt + 7u) / 8u;   // g <= leaf_max
        for (uint32_t s0 = 0; s0 < cnt; s0 += g) ch[n++] = range_ref(first + s0, cnt - s0 < g ? cnt - s0 : g);
    } else {
        uint32_t bin[8];
        bin[n++] = b.left[root]; bin[n++] = b.right[root];
        // pass 0 opens nodes that would become inner children, pass 1 opens small subtrees so that
        // leaf slots hold fewer primitives with tighter boxes
        for (int pass = 0; pass < 2; ++pass) {
            while (n < 8) {
                int best = -1;
                float best_area = -1.f;
                for (int k = 0; k < n; ++k) {
                    if (bin[k] & kLeafFlag) continue;
                    const bool big = ref_count(b, bin[k]) > b.leaf_max;
                    if ((pass == 0) != big) continue;
                    float a = half_area(b.ilo[bin[k]], b.ihi[bin[k]]);
                    if (a > best_area) { best_area = a; best = k; }
                }
                if (best < 0) break;
                const uint32_t r = bin[best];
                bin[best] = b.left[r];
                bin[n++] = b.right[r];
            }
        }
        for (int k = 0; k < n; ++k) {
            const uint32_t r = bin[k];
            if (r & kLeafFlag) ch[k] = range_ref(r & ~kLeafFlag, 1u);
            else if (ref_count(b, r) <= b.leaf_max) ch[k] = range_ref(b.first[r], ref_count(b, r));
            else ch[k] = r;
        }
    }
    for (int k = 0; k < 8; ++k) {
        c.plan[8 * w + k] = k < n ? ch[k] : kNoRef;
        if (k < n) { if (ch[k] & kLeafFlag) np += range_count(ch[k]); else ++ni; }
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

// Quantisation frame of a wide node whose children span [lo, hi]: origin = lo, one power-of-two scale per axis such that
// every child maximum fits in 8 bits (checked exactly in double).  Used by the build and by the refit.
PB_HD void node_frame(Node8& nd, V3 lo, V3 hi) {
    nd.px = lo.x; nd.py = lo.y; nd.pz = lo.z;
    uint32_t ex = quant_exponent(hi.x - lo.x), ey = quant_exponent(hi.y - lo.y), ez = quant_exponent(hi.z - lo.z);
    for (int axis = 0; axis < 3; ++axis) {
        uint32_t& e = axis == 0 ? ex : (axis == 1 ? ey : ez);
        const double p = axis == 0 ? lo.x : (axis == 1 ? lo.y : lo.z);
        const double h = axis == 0 ? hi.x : (axis == 1 ? hi.y : hi.z);
        while (e < 254u && ceil((h - p) / (double)u2f(e << 23)) > 255.0) ++e;
    }
    nd.ex = (uint8_t)ex; nd.ey = (uint8_t)ey; nd.ez = (uint8_t)ez;
}
// child box of slot s in the node's frame, rounded outward
PB_HD void node_quantise_slot(Node8& nd, int s, V3 clo, V3 chi) {
    const double sx = u2f((uint32_t)nd.ex << 23), sy = u2f((uint32_t)nd.ey << 23), sz = u2f((uint32_t)nd.ez << 23);
    auto qlo = [](float v, double p, double sc) { double q = floor(((double)v - p) / sc); return (uint8_t)(q < 0.0 ? 0.0 : (q > 255.0 ? 255.0 : q)); };
    auto qhi = [](float v, double p, double sc) { double q = ceil(((double)v - p) / sc); return (uint8_t)(q < 0.0 ? 0.0 : (q > 255.0 ? 255.0 : q)); };
    nd.qlox[s] = qlo(clo.x, nd.px, sx); nd.qloy[s] = qlo(clo.y, nd.py, sy); nd.qloz[s] = qlo(clo.z, nd.pz, sz);
    nd.qhix[s] = qhi(chi.x, nd.px, sx); nd.qhiy[s] = qhi(chi.y, nd.py, sy); nd.qhiz[s] = qhi(chi.z, nd.pz, sz);
}

PB_HD void collapse_emit(const BuildArrays& b, const CollapseArrays& c, uint32_t w, uint32_t off_inner, uint32_t off_prims) {
    uint32_t ch[8];
    V3 clo[8], chi[8];
    int n = 0;
    for (int k = 0; k < 8; ++k) {
        uint32_t r = c.plan[8 * w + k];
        if (r == kNoRef) break;
        ch[n] = r;
        slot_box(b, r, clo[n], chi[n]);
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
    node_frame(nd, lo, hi);
    uint32_t imask = 0, pmask = 0;
    for (int s = 0; s < 8; ++s) {
        nd.qlox[s] = nd.qloy[s] = nd.qloz[s] = 0;
        nd.qhix[s] = nd.qhiy[s] = nd.qhiz[s] = 0;
    }
    // children in slot order: inner children get consecutive wide indices, leaves consecutive prims
    uint32_t inner_rank = 0, prim_off = 0;
    for (int s = 0; s < 8; ++s) {
        int k = -1;
        for (int q = 0; q < n; ++q) if (slot_of[q] == s) k = q;
        if (k < 0) continue;
        node_quantise_slot(nd, s, clo[k], chi[k]);
        if (!(ch[k] & kLeafFlag)) {
            imask |= 1u << s;
            c.next_work[off_inner + inner_rank] = {ch[k], c.next_wide_base + off_inner + inner_rank};
            ++inner_rank;
        } else {
            const uint32_t cnt = range_count(ch[k]), first = range_first(ch[k]);
            pmask |= ((1u << cnt) - 1u) << (3 * s);
            for (uint32_t j = 0; j < cnt; ++j)
                c.prims_out[c.prim_cursor + off_prims + prim_off + j] = b.prims_in[b.order[first + j]];
            prim_off += cnt;
        }
    }
    nd.imask = (uint8_t)imask;
    nd.pmask = pmask; nd.spare = 0u;
    nd.child_base = c.next_wide_base + off_inner;
    nd.prim_base = c.prim_cursor + off_prims;
    c.nodes[c.work[w].wide_index] = nd;
}

// ---- refit: IAS::Update semantics (framework/world/ias_manager.cpp:116-151: optixAccelBuild in UPDATE mode keeps the
// topology and recomputes the bounds) for the flattened 8-wide tree.  One node: child boxes from the leaf slots' (already
// rewritten) primitives and from the inner children's boxes of the level below, new quantisation frame, same slots /
// masks / bases.  `box` holds one (lo, hi) pair per node; levels are processed bottom-up.
PB_HD void refit_wide_node(Node8* nodes, const PrimRec* prims, V3* box_lo, V3* box_hi, float delta, uint32_t i) {
    Node8 nd = nodes[i];
    V3 clo[8], chi[8];
    bool used[8];
    const float fm = 3.402823466e38f;
    V3 lo = v3(fm), hi = v3(-fm);
    uint32_t rank = 0, prim_off = 0;
    for (int s = 0; s < 8; ++s) {
        used[s] = false;
        const uint32_t cnt = (uint32_t)pop_cnt((nd.pmask >> (3 * s)) & 7u);
        if ((nd.imask >> s) & 1u) {
            const uint32_t ch = nd.child_base + rank++;
            clo[s] = box_lo[ch]; chi[s] = box_hi[ch];
            used[s] = true;
        } else if (cnt) {
            V3 l = v3(fm), h = v3(-fm);
            for (uint32_t j = 0; j < cnt; ++j) {
                V3 pl, ph;
                prim_bbox(prims[nd.prim_base + prim_off + j], pl, ph);
                pad_box(pl, ph, 2.0f * delta);            // B_out, as build_prim pads it
                l = {fminf(l.x, pl.x), fminf(l.y, pl.y), fminf(l.z, pl.z)};
                h = {fmaxf(h.x, ph.x), fmaxf(h.y, ph.y), fmaxf(h.z, ph.z)};
            }
            prim_off += cnt;
            clo[s] = l; chi[s] = h;
            used[s] = true;
        }
        if (used[s]) {
            lo = {fminf(lo.x, clo[s].x), fminf(lo.y, clo[s].y), fminf(lo.z, clo[s].z)};
            hi = {fmaxf(hi.x, chi[s].x), fmaxf(hi.y, chi[s].y), fmaxf(hi.z, chi[s].z)};
        }
    }
    box_lo[i] = lo; box_hi[i] = hi;
    node_frame(nd, lo, hi);
    for (int s = 0; s < 8; ++s) {
        nd.qlox[s] = nd.qloy[s] = nd.qloz[s] = 0;
        nd.qhix[s] = nd.qhiy[s] = nd.qhiz[s] = 0;
        if (used[s]) node_quantise_slot(nd, s, clo[s], chi[s]);
    }
    nodes[i] = nd;
}

}  // namespace pb
