// pb_intersect.cuh — the ray/primitive intersection contract (DESIGN.md §"Intersection contract").
//
// Replaces the closed OptiX triangle / built-in sphere intersectors behind optixTrace
// (example/path_tracer/main.cu:77,158; framework/render/emitter.h:95; sphere module
// framework/optix/module.cpp:98).  Every operation is an explicitly rounded fp32 op (pb::f_*), so
// the CPU oracle — which states the same contract independently in oracle/orc_scene.h — and this
// device code agree bit for bit:
//   accept(t, prim)  <=>  ray.tmin < t < ray.tmax  &&  primitive test succeeds
//                         &&  slabN(B_in(prim)) <= t <= slabF(B_in(prim))
//   B_in = exact bbox +- delta,  delta = 2^-18 * largest scene extent;  BVH boxes cover B_out = bbox +- 2 delta.
#pragma once
#include "pb_types.h"

namespace pb {

struct RayPre { V3 id, noid; };

PB_HD float safe_rcp_dir(float d) {
    if (fabsf(d) < 1e-20f) d = copysignf(1e-20f, d);
    return f_div(1.0f, d);
}
PB_HD RayPre ray_pre(V3 o, V3 d) {
    RayPre p;
    p.id = {safe_rcp_dir(d.x), safe_rcp_dir(d.y), safe_rcp_dir(d.z)};
    p.noid = {-f_mul(o.x, p.id.x), -f_mul(o.y, p.id.y), -f_mul(o.z, p.id.z)};
    return p;
}
PB_HD void slab(const RayPre& p, V3 lo, V3 hi, float& N, float& F) {
    float ax = f_fma(lo.x, p.id.x, p.noid.x), bx = f_fma(hi.x, p.id.x, p.noid.x);
    float ay = f_fma(lo.y, p.id.y, p.noid.y), by = f_fma(hi.y, p.id.y, p.noid.y);
    float az = f_fma(lo.z, p.id.z, p.noid.z), bz = f_fma(hi.z, p.id.z, p.noid.z);
    N = fmaxf(fmaxf(fminf(ax, bx), fminf(ay, by)), fminf(az, bz));
    F = fminf(fminf(fmaxf(ax, bx), fmaxf(ay, by)), fmaxf(az, bz));
}
PB_HD V3 cross_fma(V3 a, V3 b) {
    return {f_fma(a.y, b.z, -f_mul(a.z, b.y)), f_fma(a.z, b.x, -f_mul(a.x, b.z)), f_fma(a.x, b.y, -f_mul(a.y, b.x))};
}
PB_HD float dot_fma(V3 a, V3 b) { return f_fma(a.x, b.x, f_fma(a.y, b.y, f_mul(a.z, b.z))); }
PB_HD V3 sub3(V3 a, V3 b) { return {f_sub(a.x, b.x), f_sub(a.y, b.y), f_sub(a.z, b.z)}; }

// Moeller-Trumbore, no culling; (u, v) = weights of v1, v2
PB_HD bool tri_test(V3 o, V3 d, float tmin, float tmax, V3 v0, V3 v1, V3 v2, float& t, float& u, float& v) {
    V3 e1 = sub3(v1, v0), e2 = sub3(v2, v0);
    V3 p = cross_fma(d, e2);
    float det = dot_fma(e1, p);
    if (det == 0.f) return false;
    float inv = f_div(1.0f, det);
    V3 s = sub3(o, v0);
    u = f_mul(dot_fma(s, p), inv);
    if (u < 0.f || u > 1.f) return false;
    V3 q = cross_fma(s, e1);
    v = f_mul(dot_fma(d, q), inv);
    if (v < 0.f || f_add(u, v) > 1.f) return false;
    t = f_mul(dot_fma(e2, q), inv);
    return t > tmin && t < tmax;
}
PB_HD V3 xform_point_fma(const float* m, V3 p) {
    return {f_fma(m[0], p.x, f_fma(m[1], p.y, f_fma(m[2], p.z, m[3]))),
            f_fma(m[4], p.x, f_fma(m[5], p.y, f_fma(m[6], p.z, m[7]))),
            f_fma(m[8], p.x, f_fma(m[9], p.y, f_fma(m[10], p.z, m[11])))};
}
PB_HD V3 xform_vec_fma(const float* m, V3 p) {
    return {f_fma(m[0], p.x, f_fma(m[1], p.y, f_mul(m[2], p.z))), f_fma(m[4], p.x, f_fma(m[5], p.y, f_mul(m[6], p.z))),
            f_fma(m[8], p.x, f_fma(m[9], p.y, f_mul(m[10], p.z)))};
}
// unit sphere in object space (Minv = world->object rows)
PB_HD bool sphere_test(V3 ro, V3 rd, float tmin, float tmax, const float* Minv, float& t) {
    V3 o = xform_point_fma(Minv, ro);
    V3 d = xform_vec_fma(Minv, rd);
    float a = dot_fma(d, d);
    float b = dot_fma(o, d);
    float c = f_sub(dot_fma(o, o), 1.0f);
    float disc = f_fma(b, b, -f_mul(a, c));
    if (!(disc >= 0.f) || a == 0.f) return false;
    float s = f_sqrt(disc);
    float t0 = f_div(f_sub(-b, s), a);
    if (t0 > tmin && t0 < tmax) { t = t0; return true; }
    float t1 = f_div(f_add(-b, s), a);
    if (t1 > tmin && t1 < tmax) { t = t1; return true; }
    return false;
}

// the same test when world->object is inv * identity + (tx, ty, tz): sphere_test with the terms that multiply the
// zero entries of Minv left out.  fma(0, y, fma(0, z, t)) == t and fma(0, y, 0 * z) == +-0 exactly, so o and d — and with
// them a, b, c, the discriminant and both roots — are the values sphere_test computes, bit for bit (a zero's sign cannot
// reach t: it only enters sums with non-zero terms or squares).
PB_HD bool sphere_test_uniform(V3 ro, V3 rd, float tmin, float tmax, float inv, V3 tr, float& t) {
    V3 o = {f_fma(inv, ro.x, tr.x), f_fma(inv, ro.y, tr.y), f_fma(inv, ro.z, tr.z)};
    V3 d = {f_mul(inv, rd.x), f_mul(inv, rd.y), f_mul(inv, rd.z)};
    float a = dot_fma(d, d);
    float b = dot_fma(o, d);
    float c = f_sub(dot_fma(o, o), 1.0f);
    float disc = f_fma(b, b, -f_mul(a, c));
    if (!(disc >= 0.f) || a == 0.f) return false;
    float s = f_sqrt(disc);
    float t0 = f_div(f_sub(-b, s), a);
    if (t0 > tmin && t0 < tmax) { t = t0; return true; }
    float t1 = f_div(f_add(-b, s), a);
    if (t1 > tmin && t1 < tmax) { t = t1; return true; }
    return false;
}

// exact bbox of a primitive record (before padding)
PB_HD void prim_bbox(const PrimRec& r, V3& lo, V3& hi) {
    if (f2u(r.c.w) == PRIM_SPHERE_UNIFORM) {
        lo = {f_sub(r.a.x, r.b.x), f_sub(r.a.y, r.b.x), f_sub(r.a.z, r.b.x)};
        hi = {f_add(r.a.x, r.b.x), f_add(r.a.y, r.b.x), f_add(r.a.z, r.b.x)};
    } else if (f2u(r.c.w) == PRIM_TRI) {
        lo = {fminf(fminf(r.a.x, r.b.x), r.c.x), fminf(fminf(r.a.y, r.b.y), r.c.y), fminf(fminf(r.a.z, r.b.z), r.c.z)};
        hi = {fmaxf(fmaxf(r.a.x, r.b.x), r.c.x), fmaxf(fmaxf(r.a.y, r.b.y), r.c.y), fmaxf(fmaxf(r.a.z, r.b.z), r.c.z)};
    } else {
        lo = {f_sub(r.a.x, r.b.x), f_sub(r.a.y, r.b.y), f_sub(r.a.z, r.b.z)};
        hi = {f_add(r.a.x, r.b.x), f_add(r.a.y, r.b.y), f_add(r.a.z, r.b.z)};
    }
}
PB_HD void pad_box(V3& lo, V3& hi, float d) {
    lo = {f_sub(lo.x, d), f_sub(lo.y, d), f_sub(lo.z, d)};
    hi = {f_add(hi.x, d), f_add(hi.y, d), f_add(hi.z, d)};
}

// Full acceptance test of one primitive record.  objs = object table (for the sphere's Minv).
// The slab interval [N, F] of B_in is evaluated FIRST: acceptance needs N <= t <= F, so an empty
// interval, F <= tmin, N >= tmax or N > t_best (t_best = closest accepted t so far; ties on t_best
// must still be tested) rejects without touching the primitive test — same accepted set, and the
// sphere's dependent Minv fetch is skipped for most leaf-slot neighbours.
#ifndef PB_SPHERE_TEST_FIRST
#define PB_SPHERE_TEST_FIRST 1   // same-box A/B on C4 (round 2): 516.5 -> 520.6 Msamples/s
#endif
#ifndef PB_TRI_TEST_FIRST
#define PB_TRI_TEST_FIRST 1      // closest-hit rays only (Traversal passes tri_first = !ANY): same-box A/B on C4, twice: 542.0 -> 546.0 Msamples/s;
#endif                           // for any-hit rays too: k_connect 19.4 -> 19.7 ms per step; rejecting t > t_best before the box as well:
                                 // 525.5 (spills) - profiles/r02_extend.md, calls N and O
PB_HD bool prim_accept(const PrimRec& r, const ObjRec* objs, V3 o, V3 d, const RayPre& pre, float tmin, float tmax,
                       float t_best, float delta, float& t, float& u, float& v, bool tri_first = false) {
    V3 lo, hi;
    if (PB_SPHERE_TEST_FIRST && f2u(r.c.w) == PRIM_SPHERE_UNIFORM) {
        // the record holds everything the sphere test needs, and the test is cheaper than the slab prologue it would skip:
        // test first, evaluate the padded box's slab interval only for the (few) rays that hit the sphere.  Same accepted
        // set: tmin < t < tmax and N <= t <= F imply every rejection of the slab-first order except "N > t_best", which
        // only drops hits behind the current best (they lose the (t, id) comparison anyway).
        u = 0.f; v = 0.f;
        if (!sphere_test_uniform(o, d, tmin, tmax, r.b.y, {r.c.x, r.c.y, r.c.z}, t)) return false;
        prim_bbox(r, lo, hi);
        pad_box(lo, hi, delta);
        float N, F;
        slab(pre, lo, hi, N, F);
        return N <= t && t <= F;
    }
    if (tri_first && f2u(r.c.w) == PRIM_TRI) {
        // the same argument for triangles: Moeller-Trumbore rejects on u after about as many instructions as the slab
        // prologue takes, and only hits pay for the padded box (a warp runs this branch with 2-3 lanes)
        if (!tri_test(o, d, tmin, tmax, {r.a.x, r.a.y, r.a.z}, {r.b.x, r.b.y, r.b.z}, {r.c.x, r.c.y, r.c.z}, t, u, v)) return false;
        prim_bbox(r, lo, hi);
        pad_box(lo, hi, delta);
        float N, F;
        slab(pre, lo, hi, N, F);
        return N <= t && t <= F;
    }
    prim_bbox(r, lo, hi);
    pad_box(lo, hi, delta);
    float N, F;
    slab(pre, lo, hi, N, F);
    if (!(N <= F) || !(F > tmin) || !(N < tmax) || N > t_best) return false;
    if (f2u(r.c.w) == PRIM_TRI) {
        if (!tri_test(o, d, tmin, tmax, {r.a.x, r.a.y, r.a.z}, {r.b.x, r.b.y, r.b.z}, {r.c.x, r.c.y, r.c.z}, t, u, v)) return false;
    } else if (f2u(r.c.w) == PRIM_SPHERE_UNIFORM) {
        u = 0.f; v = 0.f;
        if (!sphere_test_uniform(o, d, tmin, tmax, r.b.y, {r.c.x, r.c.y, r.c.z}, t)) return false;
    } else {
        u = 0.f; v = 0.f;
#if defined(__CUDA_ARCH__)
        // the 12 floats of Minv as 3 x LDG.128 (offset 48 of a 128 B record) instead of 12 scalar loads
        const float4* mp = reinterpret_cast<const float4*>(objs[f2u(r.b.w)].Minv);
        const float4 m0 = __ldg(mp), m1 = __ldg(mp + 1), m2 = __ldg(mp + 2);
        const float Minv[12] = {m0.x, m0.y, m0.z, m0.w, m1.x, m1.y, m1.z, m1.w, m2.x, m2.y, m2.z, m2.w};
#else
        const float* Minv = objs[f2u(r.b.w)].Minv;
#endif
        if (!sphere_test(o, d, tmin, tmax, Minv, t)) return false;
    }
    return N <= t && t <= F;
}

}  // namespace pb
