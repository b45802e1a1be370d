// pb_scene.cuh — scene geometry in HBM and the object-space -> world-space primitive kernel body.
//
// All meshes share four arrays (positions, normals, uvs as packed floats; indices as packed
// uint32 triples); ObjRec holds the offsets.  This is the data World::LoadScene/RenderObject hand to
// OptiX as per-shape GAS inputs + per-instance transforms (framework/world/render_object.cpp:10-39,
// framework/world/gas_manager.cpp:79-131, framework/world/ias_manager.cpp:165-185); here instances
// are flattened: one world-space PrimRec per triangle / sphere, global id = insertion order.
#pragma once
#include "pb_intersect.cuh"

namespace pb {

struct SceneGeom {
    const float* pos;        // 3 floats per vertex
    const float* nrm;        // 3 floats per vertex (garbage where the object has no normals)
    const float* uv;         // 2 floats per vertex
    const uint32_t* idx;     // 3 per triangle, object-local vertex numbers
    const ObjRec* objs;
    uint32_t n_objs;
    uint32_t n_prims;
    // shading record per triangle of the index array, 96 B = 3 sectors: object-space p0 p1 p2 (9 floats),
    // n0 n1 n2 (9), t0 t1 t2 (6).  One contiguous fetch per hit instead of ten 8-12 byte gathers through
    // the index (which cost 10-13 32-byte sectors); built on the device at commit (k_make_trishade).
    const float* tri;
};
constexpr uint32_t kTriShadeFloats = 24u;

PB_HD uint32_t object_of_prim(const SceneGeom& g, uint32_t prim) {  // last object with prim_offset <= prim
    uint32_t lo = 0, hi = g.n_objs;
    while (hi - lo > 1) {
        uint32_t mid = (lo + hi) >> 1;
        if (g.objs[mid].prim_offset <= prim) lo = mid; else hi = mid;
    }
    return lo;
}

PB_HD PrimRec make_prim(const SceneGeom& g, uint32_t prim) {
    const uint32_t oi = object_of_prim(g, prim);
    const ObjRec& o = g.objs[oi];
    PrimRec r;
    if (o.flags & OBJ_SPHERE) {
        V3 r0 = {o.M[0], o.M[1], o.M[2]}, r1 = {o.M[4], o.M[5], o.M[6]}, r2 = {o.M[8], o.M[9], o.M[10]};
        r.a = {o.M[3], o.M[7], o.M[11], u2f(prim)};
        r.b = {f_sqrt(dot_fma(r0, r0)), f_sqrt(dot_fma(r1, r1)), f_sqrt(dot_fma(r2, r2)), u2f(oi)};
        r.c = {0.f, 0.f, 0.f, u2f(PRIM_SPHERE)};
        // linear part s * identity: world->object is (1/s) * identity + translation, its off-diagonal terms are +-0 and
        // drop out of the general test without changing a bit (pb_intersect.cuh)
        const float* I = o.Minv;
        const bool uniform = o.M[1] == 0.f && o.M[2] == 0.f && o.M[4] == 0.f && o.M[6] == 0.f && o.M[8] == 0.f && o.M[9] == 0.f &&
                             o.M[0] == o.M[5] && o.M[0] == o.M[10] && o.M[0] != 0.f &&
                             I[1] == 0.f && I[2] == 0.f && I[4] == 0.f && I[6] == 0.f && I[8] == 0.f && I[9] == 0.f &&
                             I[0] == I[5] && I[0] == I[10] && r.b.x == r.b.y && r.b.x == r.b.z;
        if (uniform) {
            r.b = {r.b.x, I[0], 0.f, u2f(oi)};
            r.c = {I[3], I[7], I[11], u2f(PRIM_SPHERE_UNIFORM)};
        }
    } else {
        const uint32_t* t = g.idx + 3ull * (o.index_offset + (prim - o.prim_offset));
        const float* p = g.pos + 3ull * o.vert_offset;
        V3 v0 = xform_point_fma(o.M, {p[3 * t[0]], p[3 * t[0] + 1], p[3 * t[0] + 2]});
        V3 v1 = xform_point_fma(o.M, {p[3 * t[1]], p[3 * t[1] + 1], p[3 * t[1] + 2]});
        V3 v2 = xform_point_fma(o.M, {p[3 * t[2]], p[3 * t[2] + 1], p[3 * t[2] + 2]});
        r.a = {v0.x, v0.y, v0.z, u2f(prim)};
        r.b = {v1.x, v1.y, v1.z, u2f(oi)};
        r.c = {v2.x, v2.y, v2.z, u2f(PRIM_TRI)};
    }
    return r;
}

PB_HD void make_trishade(const SceneGeom& g, uint32_t prim, float* out_all) {
    const uint32_t oi = object_of_prim(g, prim);
    const ObjRec& o = g.objs[oi];
    if (o.flags & OBJ_SPHERE) return;
    const uint32_t ti = o.index_offset + (prim - o.prim_offset);
    const uint32_t* t = g.idx + 3ull * ti;
    float* out = out_all + (size_t)kTriShadeFloats * ti;
    for (int k = 0; k < 3; ++k) {
        const size_t v = (size_t)o.vert_offset + t[k];
        out[3 * k] = g.pos[3 * v]; out[3 * k + 1] = g.pos[3 * v + 1]; out[3 * k + 2] = g.pos[3 * v + 2];
        out[9 + 3 * k] = g.nrm[3 * v]; out[9 + 3 * k + 1] = g.nrm[3 * v + 1]; out[9 + 3 * k + 2] = g.nrm[3 * v + 2];
        out[18 + 2 * k] = g.uv[2 * v]; out[18 + 2 * k + 1] = g.uv[2 * v + 1];
    }
}

// monotone float <-> uint mapping for atomicMin/atomicMax reductions of the scene box
PB_HD uint32_t float_to_ordered(float f) { uint32_t u = f2u(f); return (u & 0x80000000u) ? ~u : (u | 0x80000000u); }
PB_HD float ordered_to_float(uint32_t u) { return u2f((u & 0x80000000u) ? (u & 0x7fffffffu) : ~u); }

}  // namespace pb
