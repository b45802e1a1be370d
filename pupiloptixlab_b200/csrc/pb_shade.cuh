// pb_shade.cuh — device shading arithmetic of the wavefront integrator: RNG, sampling helpers,
// Fresnel, GGX, textures and the seven BSDFs as compile-time specialisations (one shade kernel per
// material, no indirect calls — replaces the 14 OptiX direct callables of
// framework/render/material/bsdf/bsdf.cu:7-14 / optix_material.h:36-45).
// fp32, operation order of the reference headers (cited per function); the TU is built with
// -fmad=false, so results differ from the IEEE host oracle only through libdevice sin/cos/atan2/acos.
#pragma once
#include "pb_types.h"

namespace pb {

constexpr float kPi = 3.14159265358979323846f;
constexpr float kInvPi = 0.318309886183790671538f;
constexpr float kEps = 0.000001f;      // optix/util.h:8
constexpr float kMaxDistance = 1e16f;  // optix/util.h:9

// ---- RNG: framework/cuda/random.h:14-40 (TEA-4 seed, LCG stream, 24-bit floats) -------------------
struct Rng {
    uint32_t s;
    PB_HD void init(uint32_t v0, uint32_t v1) {
        uint32_t s0 = 0;
#pragma unroll
        for (int n = 0; n < 4; ++n) {
            s0 += 0x9e3779b9u;
            v0 += ((v1 << 4) + 0xa341316cu) ^ (v1 + s0) ^ ((v1 >> 5) + 0xc8013ea4u);
            v1 += ((v0 << 4) + 0xad90777du) ^ (v0 + s0) ^ ((v0 >> 5) + 0x7e95761eu);
        }
        s = v0;
    }
    PB_HD float next() {
        s = 1664525u * s + 1013904223u;
        return static_cast<float>(s & 0x00FFFFFFu) / 0x01000000;
    }
};

// ---- framework/optix/util.h ------------------------------------------------------------------------
PB_HD V3 uniform_sample_triangle(float u1, float u2) { float su = sqrtf(u1); return {1.f - su, su * (1.f - u2), u2 * su}; }  // :33
PB_HD V3 uniform_sample_sphere(float u1, float u2) {  // :38
    float z = 1.f - 2.f * u1;
    float st = sqrtf(fmaxf(0.f, 1.f - z * z));
    float phi = 2.f * kPi * u2;
    return {st * cosf(phi), st * sinf(phi), z};
}
PB_HD V3 cosine_sample_hemisphere(float u1, float u2) {  // :45
    float st = sqrtf(u1);
    float phi = 2.0f * kPi * u2;
    V3 p;
    p.x = st * cosf(phi);
    p.y = st * sinf(phi);
    p.z = sqrtf(fmaxf(0.f, 1.f - st * st));
    return p;
}
PB_HD float cosine_hemisphere_pdf(V3 v) { return v.z > 0.f ? kInvPi * v.z : 0.f; }  // :55
PB_HD V3 uniform_sample_hemisphere(float u1, float u2) {  // :59
    float z = 1.f - 2.f * u1;
    float st = sqrtf(fmaxf(0.f, 1.f - z * z));
    float phi = 2.0f * kPi * u2;
    return {st * cosf(phi), st * sinf(phi), fabsf(z)};
}
PB_HD float uniform_hemisphere_pdf(V3 v) { return v.z > 0.f ? kInvPi * 0.5f : 0.f; }  // :70
PB_HD V3 reflect_z(V3 v) { return {-v.x, -v.y, v.z}; }                                 // :74
PB_HD V3 reflect_n(V3 v, V3 n) { return -v + 2 * dot(v, n) * n; }                       // :80
PB_HD V3 refract_z(V3 v, float cos_t, float eta) {                                     // :84
    float scale = -(cos_t < 0.f ? 1.f / eta : eta);
    return normalize(v3(scale * v.x, scale * v.y, cos_t));
}
PB_HD V3 refract_n(V3 v, V3 n, float cos_t, float eta) {  // :89
    if (cos_t < 0) eta = 1 / eta;
    return n * (dot(v, n) * eta + cos_t) - v * eta;
}
PB_HD void build_onb(V3 N, V3& b1, V3& b2) {  // :95
    float sign = copysignf(1.f, N.z);
    float a = -1.f / (sign + N.z);
    float b = N.x * N.y * a;
    b1 = v3(1.f + sign * N.x * N.x * a, sign * b, -sign * N.x);
    b2 = v3(b, sign + N.y * N.y * a, -N.y);
}
PB_HD V3 to_local(V3 v, V3 N) { V3 b1, b2; build_onb(N, b1, b2); return {dot(v, b1), dot(v, b2), dot(v, N)}; }  // :103
PB_HD V3 to_world(V3 v, V3 N) { V3 b1, b2; build_onb(N, b1, b2); return b1 * v.x + b2 * v.y + N * v.z; }        // :110
PB_HD V2 sphere_texcoord(V3 p) {  // :117
    float phi = atan2f(p.y, p.x);
    phi = phi < 0.f ? phi + kPi * 2.f : phi;
    float theta = acosf(p.z);
    return {phi * kInvPi * 0.5f, theta * kInvPi};
}
PB_HD float mis_weight(float x, float y) { return x / (x + y); }                                            // :165
PB_HD bool is_zero(float v) { return fabsf(v) < kEps; }                                                     // :169
PB_HD bool is_zero(V3 v) { return fabsf(v.x) < kEps && fabsf(v.y) < kEps && fabsf(v.z) < kEps; }            // :177

// ---- framework/render/material/fresnel.h -----------------------------------------------------------
PB_HD float fresnel_dielectric(float eta, float cos_i, float& cos_t) {  // :7
    float scale = cos_i > 0.f ? 1.f / eta : eta;
    float cos_t2 = 1.f - (1.f - cos_i * cos_i) * (scale * scale);
    if (cos_t2 <= 0.0f) { cos_t = 0.0f; return 1.0f; }
    float o_cos_i = cos_i;
    cos_i = fabsf(cos_i);
    cos_t = sqrtf(fmaxf(0.f, cos_t2));
    float rs = (cos_i - eta * cos_t) / (cos_i + eta * cos_t);
    float rp = (eta * cos_i - cos_t) / (eta * cos_i + cos_t);
    cos_t = o_cos_i > 0.f ? -cos_t : cos_t;
    return 0.5f * (rs * rs + rp * rp);
}
PB_HD float fresnel_dielectric(float eta, float cos_i) { float ct; return fresnel_dielectric(eta, cos_i, ct); }
PB_HD float fresnel_conductor1(float eta, float k, float cos_i) {  // :31
    float c2 = cos_i * cos_i;
    float s2 = 1.f - c2;
    float s4 = s2 * s2;
    float t1 = eta * eta - k * k - s2;
    float a2pb2 = sqrtf(fmaxf(0.f, t1 * t1 + 4.f * k * k * eta * eta));
    float a = sqrtf(fmaxf(0.f, 0.5f * (a2pb2 + t1)));
    float term1 = a2pb2 + c2;
    float term2 = 2.f * a * cos_i;
    float rs2 = (term1 - term2) / (term1 + term2);
    float term3 = a2pb2 * c2 + s4;
    float term4 = term2 * s2;
    float rp2 = rs2 * (term3 - term4) / (term3 + term4);
    return 0.5f * (rp2 + rs2);
}
PB_HD V3 fresnel_conductor(V3 eta, V3 k, float c) { return {fresnel_conductor1(eta.x, k.x, c), fresnel_conductor1(eta.y, k.y, c), fresnel_conductor1(eta.z, k.z, c)}; }

// ---- framework/render/material/ggx.h (isotropic, visible-area sampling) ------------------------------
PB_HD float ggx_lambda(V3 w, float alpha) { float a2 = alpha * alpha; V3 v2 = w * w; return (-1.f + sqrtf(1.f + (v2.x + v2.y) * a2 / v2.z)) / 2.f; }  // :10
PB_HD float ggx_g1(V3 w, float alpha) { return 1.f / (1.f + ggx_lambda(w, alpha)); }             // :16
PB_HD float ggx_g(V3 wi, V3 wo, float alpha) { return ggx_g1(wi, alpha) * ggx_g1(wo, alpha); }   // :20
PB_HD float ggx_d(V3 wh, float alpha) { float a2 = alpha * alpha; V3 v2 = wh * wh; float t = (v2.x + v2.y) / a2 + v2.z; return 1.f / (kPi * a2 * t * t); }  // :24
PB_HD float ggx_pdf(V3 wo, V3 wh, float alpha) { return ggx_d(wh, alpha) * ggx_g1(wo, alpha) * dot(wo, wh) / fabsf(wo.z); }  // :31
PB_HD V3 ggx_sample(V3 wo, float alpha, float xi_x, float xi_y) {  // :39
    V3 vh = normalize(v3(alpha * wo.x, alpha * wo.y, wo.z));
    V3 T1 = wo.z < 0.9999f ? normalize(cross(v3(0.f, 0.f, 1.f), vh)) : v3(1.f, 0.f, 0.f);
    V3 T2 = cross(vh, T1);
    float r = sqrtf(xi_x);
    float phi = 2.f * kPi * xi_y;
    float t1 = r * cosf(phi);
    float t2 = r * sinf(phi);
    float s = 0.5f * (1.f + vh.z);
    t2 = (1.f - s) * sqrtf(1.f - t1 * t1) + s * t2;
    V3 nh = t1 * T1 + t2 * T2 + sqrtf(fmaxf(0.f, 1.f - t1 * t1 - t2 * t2)) * vh;
    V3 ne = v3(alpha * nh.x, alpha * nh.y, fmaxf(0.f, nh.z));
    return normalize(ne);
}

// ---- textures: framework/cuda/texture.h:33-57 --------------------------------------------------------
// tex2D<float4> on a normalised-coordinate texture (texture.h:52-54; sampler state cuda/texture.cpp:60-96:
// wrap / clamp / mirror / border addressing, point or linear filter) done in software on a plain RGBA float
// array: fp32 bilinear weights instead of the texture unit's 8-bit fixed-point ones, the same function the
// oracle uses (oracle.cpp sample_bitmap), so both sides agree to rounding.
PB_HD float tex_addr(uint32_t mode, float x, uint32_t n) {
    if (mode == 0u) { x = x - floorf(x); return x * (float)n; }
    if (mode == 2u) { const float f = floorf(x); float fr = x - f; if (((int)f) & 1) fr = 1.f - fr; return fr * (float)n; }
    return fminf(fmaxf(x, 0.f), 1.f) * (float)n;
}
PB_HD V3 tex_texel(const TexBitmapView& b, int x, int y) {
    const int w = (int)b.w, h = (int)b.h;
    if (b.address_mode == 0u) { x %= w; if (x < 0) x += w; y %= h; if (y < 0) y += h; }
    else { x = x < 0 ? 0 : (x >= w ? w - 1 : x); y = y < 0 ? 0 : (y >= h ? h - 1 : y); }
    const float* p = b.texels + 4 * ((size_t)y * b.w + (size_t)x);
    return {p[0], p[1], p[2]};
}
PB_HD V3 tex_bitmap_sample(const TexRec& t, float u, float v) {
    const TexBitmapView b = tex_bitmap(t);
    if (!b.texels || b.w == 0u || b.h == 0u) return v3(0.f);
    const float x = tex_addr(b.address_mode, u, b.w), y = tex_addr(b.address_mode, v, b.h);
    if (b.filter_mode == 0u) return tex_texel(b, (int)floorf(x), (int)floorf(y));
    const float xb = x - 0.5f, yb = y - 0.5f;
    const int i = (int)floorf(xb), j = (int)floorf(yb);
    const float fa = xb - (float)i, fb = yb - (float)j;
    return (1.f - fa) * (1.f - fb) * tex_texel(b, i, j) + fa * (1.f - fb) * tex_texel(b, i + 1, j) +
           (1.f - fa) * fb * tex_texel(b, i, j + 1) + fa * fb * tex_texel(b, i + 1, j + 1);
}

PB_HD V3 tex_sample(const TexRec& t, V2 uv) {
    if (t.type == 0u) return {t.a[0], t.a[1], t.a[2]};
    float tx = t.r0[0] * uv.x + t.r0[1] * uv.y + t.r0[2] * 0.f + t.r0[3] * 1.f;
    float ty = t.r1[0] * uv.x + t.r1[1] * uv.y + t.r1[2] * 0.f + t.r1[3] * 1.f;
    if (t.type == 1u) return tex_bitmap_sample(t, tx, ty);
    tx = tx - (tx > 0.f ? floorf(tx) : ceilf(tx));
    ty = ty - (ty > 0.f ? floorf(ty) : ceilf(ty));
    if (tx < 0.f) tx += 1.f;
    if (ty < 0.f) ty += 1.f;
    const bool first = (tx > 0.5f) == (ty > 0.5f);   // patch1 where both halves agree (texture.h:45-48)
    return first ? v3(t.a[0], t.a[1], t.a[2]) : v3(t.b[0], t.b[1], t.b[2]);
}

}  // namespace pb
