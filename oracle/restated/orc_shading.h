// ORACLE — test infrastructure, not product code (see oracle/README.md).
//
// orc_shading.h — CPU restatement of the reference's per-path arithmetic: RNG, sampling helpers,
// Fresnel, GGX, textures, the seven BSDFs and the emitters.  Each function cites the reference
// lines it follows; float operation order is kept so that this file is bit-identical to the
// reference headers on the host (checked by tests/test_oracle_vs_ref.py against oracle/_ref, and
// on any machine against tests/golden/*.npz generated from oracle/_ref).
#pragma once
#include "orc_vec.h"

namespace orc {

// ---------------------------------------------------------------- RNG: framework/cuda/random.h
struct Rng {
    uint32_t s = 0;
    void init(uint32_t n, uint32_t v0, uint32_t v1) {  // random.h:14-28 (TEA, n rounds)
        uint32_t s0 = 0;
        for (uint32_t i = 0; i < n; ++i) {
            s0 += 0x9e3779b9u;
            v0 += ((v1 << 4) + 0xa341316cu) ^ (v1 + s0) ^ ((v1 >> 5) + 0xc8013ea4u);
            v1 += ((v0 << 4) + 0xad90777du) ^ (v0 + s0) ^ ((v0 >> 5) + 0x7e95761eu);
        }
        s = v0;
    }
    float next() {  // random.h:31-36 (LCG, 24-bit mantissa)
        s = 1664525u * s + 1013904223u;
        return static_cast<float>(s & 0x00FFFFFFu) / 0x01000000;
    }
    f2 next2() { float a = next(); float b = next(); return {a, b}; }  // random.h:38-40
};

// ---------------------------------------------------------------- framework/optix/util.h
constexpr float kEps = 0.000001f;        // util.h:8
constexpr float kMaxDistance = 1e16f;    // util.h:9

inline f3 uniform_sample_triangle(float u1, float u2) {  // util.h:33-36
    float su = sqrtf(u1);
    return {1.f - su, su * (1.f - u2), u2 * su};
}
inline f3 uniform_sample_sphere(float u1, float u2) {  // util.h:38-43
    float z = 1.f - 2.f * u1;
    float st = sqrtf(fmaxf(0.f, 1.f - z * z));
    float phi = 2.f * kPi * u2;
    return {st * cosf(phi), st * sinf(phi), z};
}
inline f3 cosine_sample_hemisphere(float u1, float u2) {  // util.h:45-54
    float st = sqrtf(u1);
    float phi = 2.0f * kPi * u2;
    f3 p;
    p.x = st * cosf(phi);
    p.y = st * sinf(phi);
    p.z = sqrtf(fmaxf(0.f, 1.f - st * st));
    return p;
}
inline float cosine_hemisphere_pdf(f3 v) { return v.z > 0.f ? kInvPi * v.z : 0.f; }  // :55-57
inline f3 uniform_sample_hemisphere(float u1, float u2) {  // util.h:59-69
    float z = 1.f - 2.f * u1;
    float st = sqrtf(fmaxf(0.f, 1.f - z * z));
    float phi = 2.0f * kPi * u2;
    return {st * cosf(phi), st * sinf(phi), fabsf(z)};
}
inline float uniform_hemisphere_pdf(f3 v) { return v.z > 0.f ? kInvPi * 0.5f : 0.f; }  // :70-72

inline f3 reflect_z(f3 v) { return {-v.x, -v.y, v.z}; }                          // util.h:74-78
inline f3 reflect_n(f3 v, f3 n) { return -v + 2 * dot(v, n) * n; }                // util.h:80-82
inline f3 refract_z(f3 v, float cos_t, float eta) {                              // util.h:84-87
    float scale = -(cos_t < 0.f ? 1.f / eta : eta);
    return normalize(mk3(scale * v.x, scale * v.y, cos_t));
}
inline f3 refract_n(f3 v, f3 n, float cos_t, float eta) {                        // util.h:89-92
    if (cos_t < 0) eta = 1 / eta;
    return n * (dot(v, n) * eta + cos_t) - v * eta;
}
inline void build_onb(f3 N, f3& b1, f3& b2) {  // util.h:95-101 (Pixar ONB)
    float sign = copysignf(1.f, N.z);
    float a = -1.f / (sign + N.z);
    float b = N.x * N.y * a;
    b1 = mk3(1.f + sign * N.x * N.x * a, sign * b, -sign * N.x);
    b2 = mk3(b, sign + N.y * N.y * a, -N.y);
}
inline f3 to_local(f3 v, f3 N) { f3 b1, b2; build_onb(N, b1, b2); return {dot(v, b1), dot(v, b2), dot(v, N)}; }
inline f3 to_world(f3 v, f3 N) { f3 b1, b2; build_onb(N, b1, b2); return b1 * v.x + b2 * v.y + N * v.z; }
inline f2 sphere_texcoord(f3 p) {  // util.h:117-129
    float phi = atan2f(p.y, p.x);
    phi = phi < 0.f ? phi + kPi * 2.f : phi;
    float theta = acosf(p.z);
    return {phi * kInvPi * 0.5f, theta * kInvPi};
}
inline float luminance(f3 c) { return 0.2126f * c.x + 0.7152f * c.y + 0.0722f * c.z; }  // :161-163
inline float mis_weight(float x, float y) { return x / (x + y); }                       // :165-167
inline bool is_zero(float v) { return fabsf(v) < kEps; }                                // :169-171
inline bool is_zero(f3 v) { return fabsf(v.x) < kEps && fabsf(v.y) < kEps && fabsf(v.z) < kEps; }  // :177-179
inline float lerpf(float a, float b, float t) { return a + t * (b - a); }               // :181-183

// ---------------------------------------------------------------- framework/render/material/fresnel.h
inline float fresnel_dielectric(float eta, float cos_i, float& cos_t) {  // fresnel.h:7-26
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
inline float fresnel_dielectric(float eta, float cos_i) { float ct; return fresnel_dielectric(eta, cos_i, ct); }
inline float fresnel_conductor1(float eta, float k, float cos_i) {  // fresnel.h:31-49
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
inline f3 fresnel_conductor(f3 eta, f3 k, float cos_i) {  // fresnel.h:51-56
    return {fresnel_conductor1(eta.x, k.x, cos_i), fresnel_conductor1(eta.y, k.y, cos_i),
            fresnel_conductor1(eta.z, k.z, cos_i)};
}
inline float fresnel_diffuse_reflectance(float eta) {  // fresnel.h:58-85
    if (eta < 1) {
        return -1.4399f * (eta * eta) + 0.7099f * eta + 0.6681f + 0.0636f / eta;
    } else {
        float i1 = 1.0f / eta, i2 = i1 * i1, i3 = i2 * i1, i4 = i3 * i1, i5 = i4 * i1;
        return 0.919317f - 3.4793f * i1 + 6.75335f * i2 - 7.80989f * i3 + 4.98554f * i4 - 1.36881f * i5;
    }
}

// ---------------------------------------------------------------- framework/render/material/ggx.h (isotropic, VNDF on: ggx.h:6)
inline float ggx_lambda(f3 w, float alpha) {  // ggx.h:10-14
    float a2 = alpha * alpha;
    f3 v2 = w * w;
    return (-1.f + sqrtf(1.f + (v2.x + v2.y) * a2 / v2.z)) / 2.f;
}
inline float ggx_g1(f3 w, float alpha) { return 1.f / (1.f + ggx_lambda(w, alpha)); }          // :16-18
inline float ggx_g(f3 wi, f3 wo, float alpha) { return ggx_g1(wi, alpha) * ggx_g1(wo, alpha); }  // :20-22
inline float ggx_d(f3 wh, float alpha) {  // ggx.h:24-29
    float a2 = alpha * alpha;
    f3 v2 = wh * wh;
    float t = (v2.x + v2.y) / a2 + v2.z;
    return 1.f / (kPi * a2 * t * t);
}
inline float ggx_pdf(f3 wo, f3 wh, float alpha) {  // ggx.h:31-37 (visible-area branch)
    return ggx_d(wh, alpha) * ggx_g1(wo, alpha) * dot(wo, wh) / fabsf(wo.z);
}
inline f3 ggx_sample(f3 wo, float alpha, f2 xi) {  // ggx.h:39-64 (visible-area branch)
    f3 vh = normalize(mk3(alpha * wo.x, alpha * wo.y, wo.z));
    f3 T1 = wo.z < 0.9999f ? normalize(cross(mk3(0.f, 0.f, 1.f), vh)) : mk3(1.f, 0.f, 0.f);
    f3 T2 = cross(vh, T1);
    float r = sqrtf(xi.x);
    float phi = 2.f * kPi * xi.y;
    float t1 = r * cosf(phi);
    float t2 = r * sinf(phi);
    float s = 0.5f * (1.f + vh.z);
    t2 = (1.f - s) * sqrtf(1.f - t1 * t1) + s * t2;
    f3 nh = t1 * T1 + t2 * T2 + sqrtf(fmaxf(0.f, 1.f - t1 * t1 - t2 * t2)) * vh;
    f3 ne = mk3(alpha * nh.x, alpha * nh.y, fmaxf(0.f, nh.z));
    return normalize(ne);
}

// ---------------------------------------------------------------- textures: framework/cuda/texture.h:9-57
struct Texture {
    uint32_t type = 0;  // ETextureType: 0 rgb, 1 bitmap, 2 checkerboard (util/texture.h:22-26)
    f3 a{0, 0, 0}, b{0, 0, 0};
    f4 r0{1, 0, 0, 0}, r1{0, 1, 0, 0};
    const float* bitmap = nullptr;  // RGBA float
    uint32_t w = 0, h = 0, address_mode = 0, filter_mode = 0;
};
f3 sample_bitmap(const Texture& t, float u, float v);  // oracle.cpp (host stand-in for tex2D)
inline f3 tex_sample(const Texture& t, f2 uv) {  // texture.h:33-57
    f4 tc{uv.x, uv.y, 0.f, 1.f};
    float tx = dot(t.r0, tc);
    float ty = dot(t.r1, tc);
    switch (t.type) {
        case 0: return t.a;
        case 2: {
            tx = tx - (tx > 0.f ? floorf(tx) : ceilf(tx));
            ty = ty - (ty > 0.f ? floorf(ty) : ceilf(ty));
            if (tx < 0.f) tx += 1.f;
            if (ty < 0.f) ty += 1.f;
            if (tx > 0.5f) return ty > 0.5f ? t.a : t.b;
            return ty > 0.5f ? t.b : t.a;
        }
        default: return sample_bitmap(t, tx, ty);
    }
}

// ---------------------------------------------------------------- BSDFs: framework/render/material/bsdf/*.h
enum Lobe : uint32_t {  // bsdf/bsdf.h:7-24
    kLobeUnknown = 0, kDiffuseReflection = 1 << 1, kGlossyReflection = 1 << 3, kGlossyTransmission = 1 << 4,
    kDeltaReflection = 1 << 5, kDeltaTransmission = 1 << 6, kDelta = kDeltaReflection | kDeltaTransmission
};
enum MatType : uint32_t { kDiffuse = 1, kDielectric, kRoughDielectric, kConductor, kRoughConductor, kPlastic, kRoughPlastic };

// Texture-resolved BSDF at a hit: the union of the reference's <Material>::Local structs
// (optix_material.h:25-35).
struct BsdfLocal {
    uint32_t type = 0;
    f3 reflectance{};             // diffuse.reflectance / plastic.diffuse_reflectance
    f3 specular_reflectance{};
    f3 specular_transmittance{};
    f3 eta3{}, k3{};              // conductors
    float alpha = 0, eta = 0, int_fdr = 0, specular_sampling_weight = 0;
    bool nonlinear = false;
};
struct BsdfRecord {  // bsdf/bsdf.h:26-37
    f3 wi{}, wo{};
    f3 f{0, 0, 0};
    float pdf = 0.f;
    uint32_t sampled_type = kLobeUnknown;
};

// diffuse.h:12-35
inline void diffuse_f(const BsdfLocal& b, BsdfRecord& r) {
    f3 f = mk3(0.f);
    if (r.wi.z > 0.f && r.wo.z > 0.f) f = b.reflectance * kInvPi;
    r.f = f;
}
inline void diffuse_pdf(const BsdfLocal&, BsdfRecord& r) {
    float pdf = 0.f;
    if (r.wi.z > 0.f && r.wo.z > 0.f) pdf = cosine_hemisphere_pdf(r.wi);
    r.pdf = pdf;
}
inline void diffuse_sample(const BsdfLocal& b, BsdfRecord& r, Rng& rng) {
    f2 xi = rng.next2();
    r.wi = cosine_sample_hemisphere(xi.x, xi.y);
    diffuse_pdf(b, r);
    diffuse_f(b, r);
    r.sampled_type = kDiffuseReflection;
}
// conductor.h:19-35
inline void conductor_sample(const BsdfLocal& b, BsdfRecord& r, Rng&) {
    r.wi = reflect_z(r.wo);
    r.pdf = 1.f;
    f3 fr = fresnel_conductor(b.eta3, b.k3, r.wo.z);
    r.f = b.specular_reflectance * fr / fabsf(r.wi.z);
    r.sampled_type = kDeltaReflection;
}
// dielectric.h:20-45
inline void dielectric_sample(const BsdfLocal& b, BsdfRecord& r, Rng& rng) {
    float cos_t;
    float fr = fresnel_dielectric(b.eta, r.wo.z, cos_t);
    if (rng.next() < fr) {
        r.wi = reflect_z(r.wo);
        r.pdf = fr;
        r.f = b.specular_reflectance * fr / fabsf(r.wi.z);
        r.sampled_type = kDeltaReflection;
    } else {
        r.wi = refract_z(r.wo, cos_t, b.eta);
        r.pdf = 1.f - fr;
        float factor = cos_t < 0.f ? 1.f / b.eta : b.eta;
        r.f = b.specular_transmittance * (1.f - fr) * factor * factor / fabsf(r.wi.z);
        r.sampled_type = kDeltaTransmission;
    }
}
// rough_conductor.h:21-47
inline void rough_conductor_f(const BsdfLocal& b, BsdfRecord& r) {
    r.f = mk3(0.f);
    if (r.wi.z <= 0.f || r.wo.z <= 0.f) return;
    f3 wh = normalize(r.wi + r.wo);
    f3 fo = fresnel_conductor(b.eta3, b.k3, dot(r.wo, wh));
    r.f = b.specular_reflectance * ggx_d(wh, b.alpha) * fo * ggx_g(r.wi, r.wo, b.alpha) / (4.f * r.wi.z * r.wo.z);
}
inline void rough_conductor_pdf(const BsdfLocal& b, BsdfRecord& r) {
    r.pdf = 0.f;
    if (r.wi.z <= 0.f || r.wo.z <= 0.f) return;
    f3 wh = normalize(r.wi + r.wo);
    wh = normalize(wh);
    r.pdf = ggx_pdf(r.wo, wh, b.alpha) / (4.f * dot(r.wo, wh));
}
inline void rough_conductor_sample(const BsdfLocal& b, BsdfRecord& r, Rng& rng) {
    f2 xi = rng.next2();
    r.wi = reflect_n(r.wo, ggx_sample(r.wo, b.alpha, xi));
    rough_conductor_pdf(b, r);
    rough_conductor_f(b, r);
    r.sampled_type = kDiffuseReflection;  // sic: rough_conductor.h:45
}
// rough_dielectric.h:21-97
inline void rough_dielectric_f(const BsdfLocal& b, BsdfRecord& r) {
    r.f = mk3(0.f);
    if (is_zero(r.wo.z)) return;
    f3 wh;
    bool refl = r.wo.z * r.wi.z > 0.f;
    if (refl) wh = normalize(r.wo + r.wi);
    else wh = normalize(r.wo + r.wi * (r.wo.z > 0.f ? b.eta : 1.f / b.eta));
    wh = wh * (wh.z > 0.f ? 1.f : -1.f);
    float F = fresnel_dielectric(b.eta, dot(r.wo, wh));
    float G = ggx_g(r.wi, r.wo, b.alpha);
    float D = ggx_d(wh, b.alpha);
    if (refl) {
        r.f = b.specular_reflectance * F * G * D / (4.f * fabsf(r.wi.z) * fabsf(r.wo.z));
    } else {
        float e = r.wo.z > 0.f ? b.eta : 1.f / b.eta;
        float sd = dot(r.wo, wh) + e * dot(r.wi, wh);
        r.f = b.specular_transmittance *
              fabsf((1.f - F) * D * G * dot(r.wi, wh) * dot(r.wo, wh) / (sd * sd * r.wi.z * r.wo.z));
    }
}
inline void rough_dielectric_pdf(const BsdfLocal& b, BsdfRecord& r) {
    r.pdf = 0.f;
    bool refl = r.wo.z * r.wi.z > 0.f;
    f3 wh;
    float dwh_dwo;
    if (refl) {
        wh = normalize(r.wo + r.wi);
        dwh_dwo = 1.f / (4.f * dot(r.wi, wh));
    } else {
        float e = r.wo.z > 0.f ? b.eta : 1.f / b.eta;
        wh = normalize(r.wo + r.wi * e);
        float sd = dot(r.wo, wh) + e * dot(r.wi, wh);
        dwh_dwo = (e * e * dot(r.wi, wh)) / (sd * sd);
    }
    wh = wh * (wh.z > 0.f ? 1.f : -1.f);
    f3 wo = r.wo * (r.wo.z > 0.f ? 1.f : -1.f);
    float F = fresnel_dielectric(b.eta, dot(r.wo, wh));
    r.pdf = fabsf(ggx_pdf(wo, wh, b.alpha) * (refl ? F : 1.f - F) * dwh_dwo);
}
inline void rough_dielectric_sample(const BsdfLocal& b, BsdfRecord& r, Rng& rng) {
    f2 xi = rng.next2();
    f3 wo = r.wo * (r.wo.z > 0.f ? 1.f : -1.f);
    f3 wh = ggx_sample(wo, b.alpha, xi);
    float cos_t = 0.f;
    float F = fresnel_dielectric(b.eta, dot(r.wo, wh), cos_t);
    if (rng.next() < F) {
        r.wi = reflect_n(r.wo, wh);
        r.sampled_type = kGlossyReflection;
    } else {
        if (is_zero(cos_t)) return;
        r.wi = refract_n(r.wo, wh, cos_t, b.eta);
        r.sampled_type = kGlossyTransmission;
        if (r.wi.z * r.wo.z >= 0.f) return;
    }
    rough_dielectric_pdf(b, r);
    rough_dielectric_f(b, r);
}
// plastic.h:32-81
inline f3 plastic_diff(const BsdfLocal& b) {
    return b.reflectance / (1.f - (b.nonlinear ? b.reflectance * b.int_fdr : mk3(b.int_fdr)));
}
inline float plastic_spec_prob(const BsdfLocal& b, float fo) {
    return (fo * b.specular_sampling_weight) /
           (fo * b.specular_sampling_weight + (1 - fo) * (1.f - b.specular_sampling_weight));
}
inline void plastic_f(const BsdfLocal& b, BsdfRecord& r) {
    r.f = mk3(0.f);
    if (r.wi.z <= 0.f || r.wo.z <= 0.f) return;
    float fo = fresnel_dielectric(b.eta, r.wo.z);
    float fi = fresnel_dielectric(b.eta, r.wi.z);
    f3 diff = plastic_diff(b);
    r.f = diff * (1.f - fi) * (1.f - fo) * cosine_hemisphere_pdf(r.wi) / (b.eta * b.eta * r.wi.z);
}
inline void plastic_pdf(const BsdfLocal& b, BsdfRecord& r) {
    r.pdf = 0.f;
    if (r.wi.z <= 0.f || r.wo.z <= 0.f) return;
    float fo = fresnel_dielectric(b.eta, r.wo.z);
    float sp = plastic_spec_prob(b, fo);
    r.pdf = cosine_hemisphere_pdf(r.wi) * (1.f - sp);
}
inline void plastic_sample(const BsdfLocal& b, BsdfRecord& r, Rng& rng) {
    if (r.wo.z <= 0.f) return;
    float fo = fresnel_dielectric(b.eta, r.wo.z);
    f2 xi = rng.next2();
    float sp = plastic_spec_prob(b, fo);
    if (xi.x < sp) {
        r.sampled_type = kDeltaReflection;
        r.wi = reflect_z(r.wo);
        r.f = b.specular_reflectance * fo / r.wi.z;
        r.pdf = sp;
    } else {
        r.sampled_type = kDiffuseReflection;
        r.wi = cosine_sample_hemisphere((xi.x - sp) / (1.f - sp), xi.y);
        float fi = fresnel_dielectric(b.eta, r.wi.z);
        f3 diff = plastic_diff(b);
        r.f = diff * (1.f - fi) * (1.f - fo) * cosine_hemisphere_pdf(r.wi) / (b.eta * b.eta * r.wi.z);
        r.pdf = cosine_hemisphere_pdf(r.wi) * (1.f - sp);
    }
}
// rough_plastic.h:31-86
inline void rough_plastic_f(const BsdfLocal& b, BsdfRecord& r) {
    r.f = mk3(0.f);
    if (r.wi.z <= 0.f || r.wo.z <= 0.f) return;
    float fo = fresnel_dielectric(b.eta, r.wo.z);
    f3 wh = normalize(r.wi + r.wo);
    r.f = b.specular_reflectance * fresnel_dielectric(b.eta, dot(wh, r.wo)) * ggx_d(wh, b.alpha) *
          ggx_g(r.wi, r.wo, b.alpha) / (4.f * r.wo.z * r.wi.z);
    float fi = fresnel_dielectric(b.eta, r.wi.z);
    f3 diff = plastic_diff(b);
    r.f += diff * (1.f - fi) * (1.f - fo) * kInvPi / (b.eta * b.eta);
}
inline void rough_plastic_pdf(const BsdfLocal& b, BsdfRecord& r) {
    r.pdf = 0.f;
    if (r.wi.z <= 0.f || r.wo.z <= 0.f) return;
    float fo = fresnel_dielectric(b.eta, r.wo.z);
    float sp = plastic_spec_prob(b, fo);
    float dp = 1.f - sp;
    f3 wh = normalize(r.wi + r.wo);
    r.pdf = sp * ggx_pdf(r.wo, wh, b.alpha) / (4.f * dot(r.wi, wh));
    r.pdf += dp * cosine_hemisphere_pdf(r.wi);
}
inline void rough_plastic_sample(const BsdfLocal& b, BsdfRecord& r, Rng& rng) {
    r.wi = mk3(0.f);
    if (r.wo.z <= 0.f) return;
    float fo = fresnel_dielectric(b.eta, r.wo.z);
    float sp = plastic_spec_prob(b, fo);
    f2 xi = rng.next2();
    if (xi.y < sp) {
        xi.y /= sp;
        f3 wh = ggx_sample(r.wo, b.alpha, xi);
        r.wi = reflect_n(r.wo, wh);
        r.sampled_type = kGlossyReflection;
    } else {
        xi.y = (xi.y - sp) / (1.f - sp);
        r.wi = cosine_sample_hemisphere(xi.x, xi.y);
        r.sampled_type = kDiffuseReflection;
    }
    rough_plastic_pdf(b, r);
    rough_plastic_f(b, r);
}

// LocalBsdf::Sample / Eval dispatch: optix_material.h:36-66 (Eval = GetBsdf then GetPdf)
inline void bsdf_sample(const BsdfLocal& b, BsdfRecord& r, Rng& rng) {
    switch (b.type) {
        case kDiffuse: diffuse_sample(b, r, rng); break;
        case kDielectric: dielectric_sample(b, r, rng); break;
        case kRoughDielectric: rough_dielectric_sample(b, r, rng); break;
        case kConductor: conductor_sample(b, r, rng); break;
        case kRoughConductor: rough_conductor_sample(b, r, rng); break;
        case kPlastic: plastic_sample(b, r, rng); break;
        case kRoughPlastic: rough_plastic_sample(b, r, rng); break;
    }
}
inline void bsdf_eval(const BsdfLocal& b, BsdfRecord& r) {
    switch (b.type) {
        case kDiffuse: diffuse_f(b, r); diffuse_pdf(b, r); break;
        case kDielectric: r.f = mk3(0.f); r.pdf = 0.f; break;   // dielectric.h:12-18
        case kRoughDielectric: rough_dielectric_f(b, r); rough_dielectric_pdf(b, r); break;
        case kConductor: r.f = mk3(0.f); r.pdf = 0.f; break;    // conductor.h:11-17
        case kRoughConductor: rough_conductor_f(b, r); rough_conductor_pdf(b, r); break;
        case kPlastic: plastic_f(b, r); plastic_pdf(b, r); break;
        case kRoughPlastic: rough_plastic_f(b, r); rough_plastic_pdf(b, r); break;
    }
}
inline f3 bsdf_albedo(const BsdfLocal& b) {  // optix_material.h:69-88
    switch (b.type) {
        case kDiffuse: case kPlastic: case kRoughPlastic: return b.reflectance;
        case kDielectric: case kRoughDielectric: case kConductor: case kRoughConductor: return b.specular_reflectance;
    }
    return mk3(0.f);
}

// ---------------------------------------------------------------- emitters: framework/render/emitter*.h
enum EmitterType : uint32_t { kEmNone = 0, kEmTriArea, kEmSphere, kEmConstEnv, kEmEnvMap };  // emitter/types.h:7-15
struct EmitterRec {
    uint32_t type = kEmNone;
    float weight = 0, select_probability = 0;
    Texture radiance;
    float area = 0;
    // TriArea (emitter/area.h:9-15)
    f3 p0{}, p1{}, p2{}, n0{}, n1{}, n2{};
    f2 t0{}, t1{}, t2{};
    // Sphere (emitter/sphere.h:9-12)
    f3 center{};
    float radius = 0;
    // env (emitter/env.h:6-23, :67-69)
    f3 color{};
    f3 env_center{};
    uint32_t map_w = 0, map_h = 0;
    const float* row_cdf = nullptr; uint32_t n_row_cdf = 0;
    const float* col_cdf = nullptr;
    const float* row_weight = nullptr;
    f3 tw0{}, tw1{}, tw2{}, tl0{}, tl1{}, tl2{};
    float normalization = 0, scale = 1;
};
struct EmitterSample {  // emitter/types.h:17-26; is_delta normalised to false
    f3 radiance{}, wi{}, pos{}, normal{};
    float distance = 0.f;
    float pdf = 0.f;
    bool is_delta = false;
    bool valid_distance = false;  // false when the reference leaves `distance` unwritten (quirk 3)
};
struct EmitEval { f3 radiance{0, 0, 0}; float pdf = 0.f; };  // types.h:27-30, zero-initialised (quirk 2)

inline void emitter_sample_direct(const EmitterRec& e, EmitterSample& ret, f3 hit_pos, f3 hit_n, f2 xi) {
    switch (e.type) {
        case kEmTriArea: {  // emitter/area.h:17-35
            f3 t = uniform_sample_triangle(xi.x, xi.y);
            f3 position = e.p0 * t.x + e.p1 * t.y + e.p2 * t.z;
            f3 normal = normalize(e.n0 * t.x + e.n1 * t.y + e.n2 * t.z);
            f2 tex = e.t0 * t.x + e.t1 * t.y + e.t2 * t.z;
            ret.radiance = tex_sample(e.radiance, tex);
            ret.wi = normalize(position - hit_pos);
            float NoL = dot(hit_n, ret.wi);
            float LNoL = dot(normal, -ret.wi);
            if (NoL > 0.f && LNoL > 0.f) {
                float distance = length(position - hit_pos);
                ret.pdf = distance * distance / (LNoL * e.area);
                ret.distance = distance;
                ret.valid_distance = true;
            }
            ret.pos = position;
            ret.normal = normal;
        } break;
        case kEmSphere: {  // emitter/sphere.h:14-32
            f3 t = uniform_sample_sphere(xi.x, xi.y);
            f3 position = t * e.radius + e.center;
            f3 normal = normalize(t);
            f2 tex = sphere_texcoord(t);
            ret.radiance = tex_sample(e.radiance, tex);
            ret.wi = normalize(position - hit_pos);
            float NoL = dot(hit_n, ret.wi);
            float LNoL = dot(normal, -ret.wi);
            if (NoL > 0.f && LNoL > 0.f) {
                float distance = length(position - hit_pos);
                ret.pdf = distance * distance / (LNoL * e.area);
                ret.distance = distance;
                ret.valid_distance = true;
            }
            ret.pos = position;
            ret.normal = normal;
        } break;
        case kEmConstEnv: {  // emitter/env.h:70-80
            f3 lw = uniform_sample_hemisphere(xi.x, xi.y);
            ret.wi = to_world(lw, hit_n);
            ret.pdf = uniform_hemisphere_pdf(lw);
            ret.distance = kMaxDistance;
            ret.valid_distance = true;
            ret.radiance = e.color;
            ret.is_delta = false;
            ret.pos = hit_pos + ret.wi * ret.distance;
            ret.normal = normalize(e.env_center - ret.pos);
        } break;
        case kEmEnvMap: {  // emitter/env.h:25-50
            uint32_t row = 0;
            for (; row < e.n_row_cdf - 1; ++row) if (xi.x <= e.row_cdf[row]) break;
            uint32_t col = 0;
            for (int i = row * (e.map_w + 1); col < e.map_w - 1; ++i, ++col) if (xi.y <= e.col_cdf[i]) break;
            const float phi = col * kPi * 2.f / e.map_w;
            const float theta = row * kPi / e.map_h;
            f3 lw = mk3(sinf(theta) * sinf(kPi - phi), cosf(theta), sinf(theta) * cosf(kPi - phi));
            ret.wi = mk3(dot(e.tw0, lw), dot(e.tw1, lw), dot(e.tw2, lw));
            ret.distance = kMaxDistance;
            ret.valid_distance = true;
            f2 tex = mk2(phi * 0.5f * kInvPi, theta * kInvPi);
            ret.radiance = tex_sample(e.radiance, tex) * e.scale;
            ret.is_delta = false;
            ret.pdf = luminance(ret.radiance) * e.row_weight[row] * e.normalization / fmaxf(1e-4f, fabsf(sinf(theta)));
            if (ret.pdf < 0.f) ret.pdf = 0.f;
            ret.pos = hit_pos + ret.wi * ret.distance;
            ret.normal = normalize(e.env_center - ret.pos);
        } break;
        default: break;
    }
}
inline void emitter_eval(const EmitterRec& e, EmitEval& ret, f3 emit_pos, f3 emit_n, f2 emit_uv, f3 scatter_pos) {
    switch (e.type) {
        case kEmTriArea:   // emitter/area.h:37-45
        case kEmSphere: {  // emitter/sphere.h:34-42
            f3 dir = normalize(scatter_pos - emit_pos);
            float LNoL = dot(emit_n, dir);
            if (LNoL > 0.f) {
                float distance = length(scatter_pos - emit_pos);
                ret.pdf = distance * distance / (LNoL * e.area);
                ret.radiance = tex_sample(e.radiance, emit_uv);
            }
        } break;
        case kEmConstEnv:  // emitter/env.h:82-85
            ret.pdf = 0.25f * kInvPi;
            ret.radiance = e.color;
            break;
        case kEmEnvMap: {  // emitter/env.h:52-64
            f3 dir = normalize(emit_pos - scatter_pos);
            dir = mk3(dot(e.tl0, dir), dot(e.tl1, dir), dot(e.tl2, dir));
            const float phi = kPi - atan2f(dir.x, dir.z);
            const float theta = acosf(dir.y);
            f2 tex = mk2(phi * 0.5f * kInvPi, theta * kInvPi);
            uint32_t row = static_cast<uint32_t>(tex.y * e.map_h);
            uint32_t hi = e.map_h - 2u;
            row = row < hi ? row : hi;  // clamp(row, 0u, map_size.y - 2u)
            ret.radiance = tex_sample(e.radiance, tex) * e.scale;
            ret.pdf = luminance(ret.radiance) *
                      lerpf(e.row_weight[row], e.row_weight[row + 1], tex.y * e.map_h - 1.f * row) *
                      e.normalization / fmaxf(1e-4f, fabsf(sinf(theta)));
        } break;
        default: break;
    }
}
inline f3 emitter_radiance(const EmitterRec& e, f2 uv) {  // emitter.h:53-71
    switch (e.type) {
        case kEmTriArea: case kEmSphere: case kEmEnvMap: return tex_sample(e.radiance, uv);
        case kEmConstEnv: return e.color;
        default: return mk3(0.f);
    }
}

}  // namespace orc
