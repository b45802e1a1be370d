// ORACLE — test infrastructure, not product code.
//
// orc_scene.h — host scene records, host-side scene preparation (the reference's World/
// EmitterHelper/Material::LoadMaterial work) and the DEFINITION of the ray/primitive
// intersection that replaces the closed OptiX traversal ("we define the reference
// traversal").  The product (pupiloptixlab_b200/csrc) implements the same definition separately.
#pragma once
#include <algorithm>
#include <cstdio>
#include <limits>
#include <memory>
#include <string>
#include <vector>

#include "restated/orc_shading.h"

namespace orc {

// ------------------------------------------------------------------------------------------------
// Intersection definition (shared wording with DESIGN.md §"Intersection contract")
//
//   * all arithmetic fp32, every product-sum written with explicit fmaf so that CPU and GPU agree
//     bit for bit; division and sqrt IEEE round-to-nearest.
//   * a hit (t, prim) is ACCEPTED iff  ray.tmin < t < ray.tmax,  the primitive test below
//     succeeds, and t lies inside the slab interval of the primitive's padded box B_in
//     (exact bbox ± delta, delta = 2^-18 * largest scene extent).  BVH boxes are built over
//     B_out = bbox ± 2*delta, which makes every accepted hit reachable through any conservative
//     hierarchy: closest hit = lexicographic min of (t, prim id) over accepted hits, for ANY tree.
// ------------------------------------------------------------------------------------------------
struct Ray { f3 o; float tmin; f3 d; float tmax; };
struct Hit { float t, u, v; uint32_t prim; };
constexpr uint32_t kMiss = 0xFFFFFFFFu;

inline f3 cross_fma(f3 a, f3 b) {
    return {fmaf(a.y, b.z, -(a.z * b.y)), fmaf(a.z, b.x, -(a.x * b.z)), fmaf(a.x, b.y, -(a.y * b.x))};
}
inline float dot_fma(f3 a, f3 b) { return fmaf(a.x, b.x, fmaf(a.y, b.y, a.z * b.z)); }

struct RayPre {  // per-ray constants of the slab test
    f3 id;    // 1/d with |d| clamped to >= 1e-20
    f3 noid;  // -(o * id)
};
inline float safe_rcp_dir(float d) {
    if (fabsf(d) < 1e-20f) d = copysignf(1e-20f, d);
    return 1.0f / d;
}
inline RayPre ray_pre(const Ray& r) {
    RayPre p;
    p.id = {safe_rcp_dir(r.d.x), safe_rcp_dir(r.d.y), safe_rcp_dir(r.d.z)};
    p.noid = {-(r.o.x * p.id.x), -(r.o.y * p.id.y), -(r.o.z * p.id.z)};
    return p;
}
// slab interval [N, F] of box [lo, hi]; no clamping to the ray interval
inline void slab(const RayPre& p, f3 lo, f3 hi, float& N, float& F) {
    float ax = fmaf(lo.x, p.id.x, p.noid.x), bx = fmaf(hi.x, p.id.x, p.noid.x);
    float ay = fmaf(lo.y, p.id.y, p.noid.y), by = fmaf(hi.y, p.id.y, p.noid.y);
    float az = fmaf(lo.z, p.id.z, p.noid.z), bz = fmaf(hi.z, p.id.z, p.noid.z);
    N = fmaxf(fmaxf(fminf(ax, bx), fminf(ay, by)), fminf(az, bz));
    F = fminf(fminf(fmaxf(ax, bx), fmaxf(ay, by)), fmaxf(az, bz));
}

// Möller–Trumbore, no culling; (u, v) = weights of v1, v2 (optixGetTriangleBarycentrics layout)
inline bool tri_test(const Ray& r, f3 v0, f3 v1, f3 v2, float& t, float& u, float& v) {
    f3 e1 = v1 - v0, e2 = v2 - v0;
    f3 p = cross_fma(r.d, e2);
    float det = dot_fma(e1, p);
    if (det == 0.f) return false;
    float inv = 1.0f / det;
    f3 s = r.o - v0;
    u = dot_fma(s, p) * inv;
    if (u < 0.f || u > 1.f) return false;
    f3 q = cross_fma(s, e1);
    v = dot_fma(r.d, q) * inv;
    if (v < 0.f || u + v > 1.f) return false;
    t = dot_fma(e2, q) * inv;
    return t > r.tmin && t < r.tmax;
}
// unit sphere in object space; Minv = rows of the world->object 3x4
inline f3 xform_point_fma(const float* m, f3 p) {
    return {fmaf(m[0], p.x, fmaf(m[1], p.y, fmaf(m[2], p.z, m[3]))),
            fmaf(m[4], p.x, fmaf(m[5], p.y, fmaf(m[6], p.z, m[7]))),
            fmaf(m[8], p.x, fmaf(m[9], p.y, fmaf(m[10], p.z, m[11])))};
}
inline f3 xform_vec_fma(const float* m, f3 p) {
    return {fmaf(m[0], p.x, fmaf(m[1], p.y, m[2] * p.z)), fmaf(m[4], p.x, fmaf(m[5], p.y, m[6] * p.z)),
            fmaf(m[8], p.x, fmaf(m[9], p.y, m[10] * p.z))};
}
inline bool sphere_test(const Ray& r, const float* Minv, float& t) {
    f3 o = xform_point_fma(Minv, r.o);
    f3 d = xform_vec_fma(Minv, r.d);
    float a = dot_fma(d, d);
    float b = dot_fma(o, d);
    float c = dot_fma(o, o) - 1.0f;
    float disc = fmaf(b, b, -(a * c));
    if (!(disc >= 0.f) || a == 0.f) return false;
    float s = sqrtf(disc);
    float t0 = (-b - s) / a;
    if (t0 > r.tmin && t0 < r.tmax) { t = t0; return true; }
    float t1 = (-b + s) / a;
    if (t1 > r.tmin && t1 < r.tmax) { t = t1; return true; }
    return false;
}

// ------------------------------------------------------------------------------------------------
struct Material {
    uint32_t type = 0;
    bool twosided = false;
    float int_ior = 0, ext_ior = 0, eta = 0;  // eta = int/ext (optix_material.cpp:64,94,111)
    bool nonlinear = false;
    Texture reflectance, specular_reflectance, specular_transmittance, alpha, eta_tex, k_tex;
    float int_fdr = 0, specular_sampling_weight = 0;  // optix_material.cpp:95-99,112-116
};

struct Object {
    bool sphere = false;
    float M[12]{};     // object -> world rows
    float Minv[12]{};  // world -> object rows
    uint32_t flags = 0, material = 0;
    bool is_emitter = false;
    Texture area_radiance;
    int emitter_offset = -1;
    uint32_t prim_offset = 0, n_prims = 0;
    std::vector<float> pos, nrm, uv;
    std::vector<uint32_t> idx;
};

struct Prim {  // world space
    f3 v0, v1, v2;   // triangle vertices | sphere: v0 = centre, v1 = half extents of the world box
    uint32_t obj;
    uint32_t kind;   // 0 triangle, 1 sphere
};

// world->object of an affine 3x4: adjugate/determinant in double, rounded once to float.
// The product computes the same expression (pupiloptixlab_b200/csrc/pb_host_math.h).
inline void invert_affine(const float* M, float* inv) {
    double a = M[0], b = M[1], c = M[2], d = M[4], e = M[5], f = M[6], g = M[8], h = M[9], i = M[10];
    double A = e * i - f * h, B = -(d * i - f * g), C = d * h - e * g;
    double det = a * A + b * B + c * C;
    double r[9] = {A / det, -(b * i - c * h) / det, (b * f - c * e) / det,
                   B / det, (a * i - c * g) / det, -(a * f - c * d) / det,
                   C / det, -(a * h - b * g) / det, (a * e - b * d) / det};
    double tx = M[3], ty = M[7], tz = M[11];
    for (int k = 0; k < 3; ++k) {
        inv[4 * k + 0] = (float)r[3 * k + 0];
        inv[4 * k + 1] = (float)r[3 * k + 1];
        inv[4 * k + 2] = (float)r[3 * k + 2];
        inv[4 * k + 3] = (float)(-(r[3 * k + 0] * tx + r[3 * k + 1] * ty + r[3 * k + 2] * tz));
    }
}

// Transform::TransformPoint with an affine matrix (framework/util/transform.cpp:100-107; w == 1)
inline f3 transform_point(const float* m, f3 p) {
    float x = m[0] * p.x + m[1] * p.y + m[2] * p.z + m[3];
    float y = m[4] * p.x + m[5] * p.y + m[6] * p.z + m[7];
    float z = m[8] * p.x + m[9] * p.y + m[10] * p.z + m[11];
    float w = 0.f * p.x + 0.f * p.y + 0.f * p.z + 1.f;
    return {x / w, y / w, z / w};
}
// normal by the inverse-transpose: n'_j = sum_i n_i * inv[i][j]
// (optixTransformNormalFromObjectToWorldSpace; host analogue transform.cpp:115-123)
inline f3 transform_normal_raw(const float* inv, f3 n) {
    return {inv[0] * n.x + inv[4] * n.y + inv[8] * n.z, inv[1] * n.x + inv[5] * n.y + inv[9] * n.z,
            inv[2] * n.x + inv[6] * n.y + inv[10] * n.z};
}

}  // namespace orc
