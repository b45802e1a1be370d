// tests/hostsim/hostsim.cpp — TEST HARNESS ONLY.  Compiles the per-thread bodies of the CUDA kernels
// (pupiloptixlab_b200/csrc/*.cuh, all PB_HD) with g++ and runs them in plain loops, so the BVH
// build / traversal LOGIC can be checked against the oracle in the GPU-less dev container.  It is
// not part of the product, is never loaded by pupiloptixlab_b200 and is no fallback: the product
// library fails with PB_ERR_NO_DEVICE when no B200 is present.
#include <algorithm>
#include <cstdio>
#include <cstring>
#include <numeric>
#include <vector>

#include "../../pupiloptixlab_b200/csrc/pb_collapse.cuh"
#include "../../pupiloptixlab_b200/csrc/pb_host_math.h"
#include "../../pupiloptixlab_b200/csrc/pb_scene.cuh"
#include "../../pupiloptixlab_b200/csrc/pb_slots.cuh"
#include "../../pupiloptixlab_b200/csrc/pb_host_prep.h"
#include "../devsim/table_body.h"

using namespace pb;

struct HS {
    std::vector<float> pos;
    std::vector<uint32_t> idx;
    std::vector<ObjRec> objs;
    std::vector<PrimRec> prims_in, prims_out;
    std::vector<Node8> nodes;
    float delta = 0;
    uint32_t levels = 0;
    TraceCounters cnt{};
    std::vector<uint32_t> level_first;      // BFS level boundaries of `nodes`
    std::vector<V3> box_lo, box_hi;         // per-node boxes of the refit
};

extern "C" {

static uint32_t hs_bottom_max = 16, hs_leaf_max = 3;
void hs_set_leaf_max(uint32_t v) { hs_leaf_max = v; }
void hs_set_bottom_max(uint32_t v) { hs_bottom_max = v; }

HS* hs_create(const float* pos, uint32_t n_vert, const uint32_t* idx, uint32_t n_tri, const float* mesh_M,
              const float* sphere_M, uint32_t n_sph) {
    HS* h = new HS();
    h->pos.assign(pos, pos + 3 * (size_t)n_vert);
    h->idx.assign(idx, idx + 3 * (size_t)n_tri);
    uint32_t prim = 0;
    if (n_tri) {
        ObjRec o{};
        memcpy(o.M, mesh_M, 48);
        invert_affine(o.M, o.Minv);
        o.prim_offset = 0; o.n_prims = n_tri; o.emitter_offset = -1;
        h->objs.push_back(o);
        prim += n_tri;
    }
    for (uint32_t s = 0; s < n_sph; ++s) {
        ObjRec o{};
        memcpy(o.M, sphere_M + 12 * s, 48);
        invert_affine(o.M, o.Minv);
        o.flags = OBJ_SPHERE; o.prim_offset = prim; o.n_prims = 1; o.emitter_offset = -1;
        h->objs.push_back(o);
        prim += 1;
    }
    const uint32_t n = prim;
    SceneGeom g{h->pos.data(), nullptr, nullptr, h->idx.data(), h->objs.data(), (uint32_t)h->objs.size(), n, nullptr};
    h->prims_in.resize(n);
    const float fm = 3.402823466e38f;
    V3 slo = v3(fm), shi = v3(-fm);
    for (uint32_t i = 0; i < n; ++i) {
        h->prims_in[i] = make_prim(g, i);
        V3 lo, hi;
        prim_bbox(h->prims_in[i], lo, hi);
        slo = {fminf(slo.x, lo.x), fminf(slo.y, lo.y), fminf(slo.z, lo.z)};
        shi = {fmaxf(shi.x, hi.x), fmaxf(shi.y, hi.y), fmaxf(shi.z, hi.z)};
    }
    float ext = n ? fmaxf(fmaxf(shi.x - slo.x, shi.y - slo.y), shi.z - slo.z) : 0.f;
    h->delta = ext * (1.0f / 262144.0f);
    if (n == 0) return h;

    std::vector<V3> blo(n), bhi(n), ilo(n), ihi(n);
    std::vector<unsigned long long> keys(n);
    std::vector<uint32_t> order(n), left(n), right(n), pint(n), pleaf(n), first(n), last(n), flag(n, 0);
    BuildArrays b{h->prims_in.data(), n, h->delta, slo, shi, blo.data(), bhi.data(), keys.data(), order.data(),
                  left.data(), right.data(), pint.data(), pleaf.data(), first.data(), last.data(), ilo.data(), ihi.data(), flag.data(), hs_leaf_max, hs_bottom_max < 8u * hs_leaf_max ? hs_bottom_max : 8u * hs_leaf_max};
    for (uint32_t i = 0; i < n; ++i) build_prim(b, i);
    // sort pairs (stands in for cub::DeviceRadixSort)
    std::vector<uint32_t> perm(n);
    std::iota(perm.begin(), perm.end(), 0u);
    std::stable_sort(perm.begin(), perm.end(), [&](uint32_t x, uint32_t y) { return keys[x] < keys[y]; });
    { std::vector<unsigned long long> k2(n); std::vector<uint32_t> o2(n);
      for (uint32_t i = 0; i < n; ++i) { k2[i] = keys[perm[i]]; o2[i] = order[perm[i]]; }
      keys.swap(k2); order.swap(o2); b.keys = keys.data(); b.order = order.data(); }
    for (uint32_t i = 0; i + 1 < n; ++i) build_radix_node(b, i);
    if (n == 1) pleaf[0] = kNoRef;
    for (uint32_t i = 0; i < n; ++i)
        build_refit_from_leaf(b, i, [](uint32_t* f) { uint32_t old = *f; *f = old + 1; return old; });

    // collapse, level by level
    h->nodes.resize(n);
    h->prims_out.resize(n);
    std::vector<WorkItem> work(1), next;
    work[0] = {n == 1 ? (0u | kLeafFlag) : 0u, 0u};
    uint32_t wide_count = 1, prim_cursor = 0;
    h->level_first.assign(1, 0u);
    while (!work.empty()) {
        ++h->levels;
        h->level_first.push_back(wide_count);
        uint32_t nw = (uint32_t)work.size();
        std::vector<uint32_t> plan(8 * (size_t)nw), ci(nw), cp(nw);
        next.assign(8 * (size_t)nw, WorkItem{0, 0});
        CollapseArrays c{work.data(), nw, plan.data(), ci.data(), cp.data(), next.data(), wide_count, prim_cursor,
                         h->nodes.data(), h->prims_out.data()};
        for (uint32_t w = 0; w < nw; ++w) collapse_plan(b, c, w);
        std::vector<uint32_t> oi(nw), op(nw);
        uint32_t ti = 0, tp = 0;
        for (uint32_t w = 0; w < nw; ++w) { oi[w] = ti; ti += ci[w]; op[w] = tp; tp += cp[w]; }
        for (uint32_t w = 0; w < nw; ++w) collapse_emit(b, c, w, oi[w], op[w]);
        next.resize(ti);
        wide_count += ti; prim_cursor += tp;
        work.swap(next);
    }
    h->nodes.resize(wide_count);
    h->level_first.push_back(wide_count);
    if (prim_cursor != n) fprintf(stderr, "hostsim: leaf prim count %u != %u\n", prim_cursor, n);
    return h;
}
void hs_destroy(HS* h) { delete h; }
uint32_t hs_num_nodes(HS* h) { return (uint32_t)h->nodes.size(); }
uint32_t hs_levels(HS* h) { return h->levels; }
void hs_counters(HS* h, unsigned long long* out) { out[0] = h->cnt.nodes; out[1] = h->cnt.tris; out[2] = h->cnt.spheres; }

static TraceScene scene_of(HS* h) {
    return TraceScene{reinterpret_cast<const U4*>(h->nodes.data()), reinterpret_cast<const V4*>(h->prims_out.data()),
                      h->objs.data(), h->delta, (uint32_t)h->nodes.size(), 32u, 1u, 0x47000000u, 0xFFFFFFFFu};
}
void hs_trace_closest(HS* h, const float* rays, uint64_t n, float* hits /* t,u,v,prim bits */) {
    TraceScene sc = scene_of(h);
    U2 smem[PB_SMEM_STACK + PB_COLD_ENTRIES];
    U2 tail[PB_LOCAL_STACK];
    LaneStack st;
    st.init(smem, 1, tail);
    h->cnt = TraceCounters{};
    for (uint64_t i = 0; i < n; ++i) {
        const float* r = rays + 8 * i;
        Hit hit;
        traverse<false, true>(sc, {r[0], r[1], r[2]}, {r[4], r[5], r[6]}, r[3], r[7], hit, st, &h->cnt);
        hits[4 * i] = hit.t; hits[4 * i + 1] = hit.u; hits[4 * i + 2] = hit.v; hits[4 * i + 3] = u2f(hit.prim);
    }
}
void hs_trace_any(HS* h, const float* rays, uint64_t n, uint8_t* occ) {
    TraceScene sc = scene_of(h);
    U2 smem[PB_SMEM_STACK + PB_COLD_ENTRIES];
    U2 tail[PB_LOCAL_STACK];
    LaneStack st;
    st.init(smem, 1, tail);
    h->cnt = TraceCounters{};
    for (uint64_t i = 0; i < n; ++i) {
        const float* r = rays + 8 * i;
        Hit hit;
        traverse<true, true>(sc, {r[0], r[1], r[2]}, {r[4], r[5], r[6]}, r[3], r[7], hit, st, &h->cnt);
        occ[i] = hit.prim != 0xFFFFFFFFu;
    }
}
// pb_update_instance's refit path (pb_build.cu refit_object) on the CPU: new transform of object `obj`, its leaf records
// rewritten, every level refitted bottom-up with the same PB_HD body.  Returns 0, or 1 when delta would change (the
// product rebuilds then).
int hs_move_object(HS* h, uint32_t obj, const float* M) {
    if (obj >= h->objs.size()) return 2;
    memcpy(h->objs[obj].M, M, 48);
    invert_affine(h->objs[obj].M, h->objs[obj].Minv);
    const uint32_t n = (uint32_t)h->prims_out.size();
    SceneGeom g{h->pos.data(), nullptr, nullptr, h->idx.data(), h->objs.data(), (uint32_t)h->objs.size(), n, nullptr};
    for (uint32_t i = 0; i < n; ++i)
        if (f2u(h->prims_out[i].b.w) == obj) h->prims_out[i] = make_prim(g, f2u(h->prims_out[i].a.w));
    const float fm = 3.402823466e38f;
    V3 slo = v3(fm), shi = v3(-fm);
    for (uint32_t i = 0; i < n; ++i) {
        V3 lo, hi;
        prim_bbox(h->prims_out[i], lo, hi);
        slo = {fminf(slo.x, lo.x), fminf(slo.y, lo.y), fminf(slo.z, lo.z)};
        shi = {fmaxf(shi.x, hi.x), fmaxf(shi.y, hi.y), fmaxf(shi.z, hi.z)};
    }
    const float delta = fmaxf(fmaxf(shi.x - slo.x, shi.y - slo.y), shi.z - slo.z) * (1.0f / 262144.0f);
    if (f2u(delta) != f2u(h->delta)) return 1;
    h->box_lo.resize(h->nodes.size()); h->box_hi.resize(h->nodes.size());
    for (size_t l = h->level_first.size() - 1; l-- > 0;)
        for (uint32_t i = h->level_first[l]; i < h->level_first[l + 1]; ++i)
            refit_wide_node(h->nodes.data(), h->prims_out.data(), h->box_lo.data(), h->box_hi.data(), h->delta, i);
    return 0;
}
// the committed tree as pb_export_bvh hands it out: 80-byte nodes + 48-byte primitive records in leaf order
void hs_export(HS* h, void* nodes80, void* prims48) {
    if (nodes80 && !h->nodes.empty()) memcpy(nodes80, h->nodes.data(), h->nodes.size() * sizeof(Node8));
    if (prims48 && !h->prims_out.empty()) memcpy(prims48, h->prims_out.data(), h->prims_out.size() * sizeof(PrimRec));
}
uint32_t hs_num_prims(HS* h) { return (uint32_t)h->prims_out.size(); }
// per-BSDF known-answer table computed by the product's pb_bsdf.cuh on the CPU (tests/devsim/table_body.h)
int hs_bsdf_table(const pb_material* mat, const float* wo, const uint32_t* seeds, uint32_t n, float* out) {
    if (!mat) return 1;
    const MatRec m = prep::mat_rec(*mat);
    for (uint32_t i = 0; i < n; ++i) bsdf_table_entry_any(m, wo + 3 * i, seeds[i], out + 16 * i);
    return 0;
}
// wave slot -> (frame, pixel) for every slot of a wave of K frames of a w x h film, and the inverse used by k_accumulate
void hs_slot_tables(uint32_t w, uint32_t h, uint32_t K, uint32_t* frame_of_slot, uint32_t* pixel_of_slot, uint32_t* slot_of_kj) {
    const uint32_t n_px = w * h;
    for (uint32_t slot = 0; slot < K * n_px; ++slot) {
        uint32_t k, j;
        slot_to_kj(slot, n_px, K, k, j);
        frame_of_slot[slot] = k;
        pixel_of_slot[slot] = j < n_px ? path_pixel(j, w, h) : 0xFFFFFFFFu;
    }
    for (uint32_t k = 0; k < K; ++k)
        for (uint32_t j = 0; j < n_px; ++j) slot_of_kj[(size_t)k * n_px + j] = kj_to_slot(k, j, n_px, K);
}
}  // extern "C"
