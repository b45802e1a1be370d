// pb_host_prep.h — host-side scene preparation of the product: what the reference does on the CPU
// between scene load and the first launch.  fp32, same operation order as the reference.
//   Material::LoadMaterial          framework/render/material/optix_material.cpp:9-132
//   EmitterHelper::{SetMeshAreaEmitter,SetSphereAreaEmitter,ComputeProbability}
//                                   framework/world/emitter.cpp:169-243,321-337
#pragma once
#include <cmath>
#include <vector>

#include "../../include/pupil_b200.h"
#include "pb_host_math.h"
#include "pb_types.h"

namespace pb {
namespace prep {

inline float luminance(const float c[3]) { return 0.2126f * c[0] + 0.7152f * c[1] + 0.0722f * c[2]; }  // optix/util.h:161

inline float diffuse_fresnel(float eta) {  // fresnel::DiffuseReflectance, fresnel.h:58-85
    if (eta < 1) return -1.4399f * (eta * eta) + 0.7099f * eta + 0.6681f + 0.0636f / eta;
    float i1 = 1.0f / eta, i2 = i1 * i1, i3 = i2 * i1, i4 = i3 * i1, i5 = i4 * i1;
    return 0.919317f - 3.4793f * i1 + 6.75335f * i2 - 7.80989f * i3 + 4.98554f * i4 - 1.36881f * i5;
}

inline void pixel_average(const pb_texture& t, float out[3]) {  // optix_material.cpp:9-39
    out[0] = out[1] = out[2] = 0.f;
    if (t.type == PB_TEX_RGB) { out[0] = t.a[0]; out[1] = t.a[1]; out[2] = t.a[2]; }
    else if (t.type == PB_TEX_CHECKERBOARD) {
        for (int k = 0; k < 3; ++k) out[k] = (t.a[k] + t.b[k]) * 0.5f;
    } else if (t.bitmap) {
        float r = 0.f, g = 0.f, b = 0.f;
        size_t idx = 0;
        for (size_t i = 0; i < (size_t)t.bitmap_w * t.bitmap_h; ++i) { r += t.bitmap[idx++]; g += t.bitmap[idx++]; b += t.bitmap[idx++]; idx++; }
        float inv = 1.0f / (1.f * t.bitmap_h * t.bitmap_w);
        out[0] = r * inv; out[1] = g * inv; out[2] = b * inv;
    }
}

inline float max3(float r, float g, float b) { return (r > g ? (r > b ? r : b) : (g > b ? g : b)); }
inline float emitter_weight(const pb_texture& t) {  // GetWeight, world/emitter.cpp:73-104
    if (t.type == PB_TEX_RGB) return max3(t.a[0], t.a[1], t.a[2]);
    if (t.type == PB_TEX_CHECKERBOARD) return (max3(t.a[0], t.a[1], t.a[2]) + max3(t.b[0], t.b[1], t.b[2])) * 0.5f;
    float w = 0.f;
    if (t.bitmap) {
        for (size_t i = 0; i < t.bitmap_w; i++)
            for (size_t j = 0; j < t.bitmap_h; j++) {
                size_t k = (i * t.bitmap_w + j) * 4;   // (sic) emitter.cpp:93-97
                if (k + 2 < (size_t)4 * t.bitmap_w * t.bitmap_h) w += max3(t.bitmap[k], t.bitmap[k + 1], t.bitmap[k + 2]);
            }
        w /= 1.f * t.bitmap_w * t.bitmap_h;
    }
    return w;
}

// `texels` = device copy of t.bitmap (bitmap textures only; see TexBitmapView)
inline TexRec tex_rec(const pb_texture& t, const float* texels = nullptr) {
    TexRec r{};
    r.type = t.type;
    for (int k = 0; k < 3; ++k) { r.a[k] = t.a[k]; r.b[k] = t.b[k]; }
    for (int k = 0; k < 4; ++k) { r.r0[k] = t.uv_r0[k]; r.r1[k] = t.uv_r1[k]; }
    r.bitmap_w = t.bitmap_w;
    if (t.type == PB_TEX_BITMAP) {
        const unsigned long long p = (unsigned long long)(uintptr_t)texels;
        r.a[0] = u2f(t.address_mode); r.a[1] = u2f(t.filter_mode); r.a[2] = u2f(texels ? t.bitmap_h : 0u);
        r.b[0] = u2f((uint32_t)p); r.b[1] = u2f((uint32_t)(p >> 32)); r.b[2] = 0.f;
    }
    return r;
}

// maps the host texels of a pb_texture to their device copy
typedef const float* (*texel_lookup)(void* user, const pb_texture& t);

inline MatRec mat_rec(const pb_material& m, texel_lookup dev = nullptr, void* user = nullptr) {
    auto tex_rec = [&](const pb_texture& t) { return prep::tex_rec(t, dev && t.type == PB_TEX_BITMAP ? dev(user, t) : nullptr); };
    MatRec r{};
    r.type = m.type;
    r.twosided = m.twosided;
    r.eta = m.int_ior / m.ext_ior;
    r.int_ior_over_ext_dielectric = m.int_ior / m.ext_ior;
    r.nonlinear = m.nonlinear;
    r.tex[TEX_REFLECTANCE] = tex_rec(m.reflectance);
    r.tex[TEX_SPEC_REFL] = tex_rec(m.specular_reflectance);
    r.tex[TEX_SPEC_TRANS] = tex_rec(m.specular_transmittance);
    r.tex[TEX_ALPHA] = tex_rec(m.alpha);
    r.tex[TEX_ETA] = tex_rec(m.eta);
    r.tex[TEX_K] = tex_rec(m.k);
    if (m.type == PB_MAT_PLASTIC || m.type == PB_MAT_ROUGH_PLASTIC) {  // optix_material.cpp:95-99,112-116
        float d[3], s[3];
        pixel_average(m.reflectance, d);
        pixel_average(m.specular_reflectance, s);
        float dl = luminance(d), sl = luminance(s);
        r.specular_sampling_weight = sl / (sl + dl);
        r.int_fdr = diffuse_fresnel(1.f / r.eta);
    }
    return r;
}

// Transform::TransformPoint (framework/util/transform.cpp:100-107) with an affine matrix (w == 1)
inline void transform_point(const float* m, const float p[3], float out[3]) {
    float x = m[0] * p[0] + m[1] * p[1] + m[2] * p[2] + m[3];
    float y = m[4] * p[0] + m[5] * p[1] + m[6] * p[2] + m[7];
    float z = m[8] * p[0] + m[9] * p[1] + m[10] * p[2] + m[11];
    float w = 0.f * p[0] + 0.f * p[1] + 0.f * p[2] + 1.f;
    out[0] = x / w; out[1] = y / w; out[2] = z / w;
}
// Transform::TransformNormal with the inverse-transpose (transform.cpp:115-123)
inline void transform_normal(const float* inv, const float n[3], float out[3]) {
    float x = inv[0] * n[0] + inv[4] * n[1] + inv[8] * n[2];
    float y = inv[1] * n[0] + inv[5] * n[1] + inv[9] * n[2];
    float z = inv[2] * n[0] + inv[6] * n[1] + inv[10] * n[2];
    float len = sqrtf(x * x + y * y + z * z);
    out[0] = x / len; out[1] = y / len; out[2] = z / len;
}

}  // namespace prep
}  // namespace pb
