// pb_emitter.cuh — emitter selection, direct sampling and evaluation on the device.
// Follows framework/render/emitter.h:110-136 (SelectOneEmiiter), emitter/area.h:17-45,
// emitter/sphere.h:14-42, emitter/env.h:6-65 (environment map), :67-86 (constant environment).
#pragma once
#include "pb_shade.cuh"

namespace pb {

struct EmitterTable {
    const EmitRec* rec;      // [0, n_area) area emitters, rec[n_area] = env emitter (type EM_NONE if absent)
    const float* prefix;     // prefix[i] = running float sum of select_probability after emitter i
    uint32_t n_area;
};

// First i with p <= running_sum_i — the linear scan of emitter.h:113-121 done as a binary search
// over the identical float prefix sums; falls back to env, else the last emitter (emitter.h:135).
PB_HD uint32_t select_emitter(const EmitterTable& t, float p) {
    uint32_t lo = 0, hi = t.n_area;
    while (lo < hi) {
        uint32_t mid = (lo + hi) >> 1;
        if (p <= t.prefix[mid]) hi = mid; else lo = mid + 1;
    }
    if (lo < t.n_area) return lo;
    // never called on an emitter-less scene (k_shade's has_emitters); the slot n_area always exists, so index it then
    return (t.rec[t.n_area].type != EM_NONE || t.n_area == 0u) ? t.n_area : t.n_area - 1;
}

struct EmitterSample { V3 radiance, wi; float distance, pdf; bool valid_distance; };
struct EmitEval { V3 radiance; float pdf; };

PB_HD V3 ld3(const float* p) { return {p[0], p[1], p[2]}; }
PB_HD float luminance(V3 c) { return 0.2126f * c.x + 0.7152f * c.y + 0.0722f * c.z; }   // optix/util.h:161-163

PB_HD void emitter_sample_direct(const EmitRec& e, EmitterSample& ret, V3 hit_pos, V3 hit_n, float xi_x, float xi_y) {
    ret.pdf = 0.f; ret.distance = 0.f; ret.valid_distance = false; ret.radiance = v3(0.f); ret.wi = v3(0.f);
    if (e.type == EM_TRI || e.type == EM_SPHERE) {
        V3 position, normal;
        V2 tex;
        if (e.type == EM_TRI) {  // area.h:17-22
            V3 t = uniform_sample_triangle(xi_x, xi_y);
            position = ld3(e.p0) * t.x + ld3(e.p1) * t.y + ld3(e.p2) * t.z;
            normal = normalize(ld3(e.n0) * t.x + ld3(e.n1) * t.y + ld3(e.n2) * t.z);
            tex = v2(e.t0[0], e.t0[1]) * t.x + v2(e.t1[0], e.t1[1]) * t.y + v2(e.t2[0], e.t2[1]) * t.z;
        } else {  // sphere.h:14-19
            V3 t = uniform_sample_sphere(xi_x, xi_y);
            position = t * e.p1[0] + ld3(e.p0);
            normal = normalize(t);
            tex = sphere_texcoord(t);
        }
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
    } else if (e.type == EM_CONST_ENV) {  // env.h:70-80
        V3 lw = uniform_sample_hemisphere(xi_x, xi_y);
        ret.wi = to_world(lw, hit_n);
        ret.pdf = uniform_hemisphere_pdf(lw);
        ret.distance = kMaxDistance;
        ret.valid_distance = true;
        ret.radiance = ld3(e.p0);
    } else if (e.type == EM_ENV_MAP) {  // env.h:25-50: linear scans of the row / column CDFs as in the reference
        uint32_t row = 0;
        for (; row < e.n_row_cdf - 1u; ++row) if (xi_x <= e.row_cdf[row]) break;
        uint32_t col = 0;
        for (uint32_t i = row * (e.map_w + 1u); col < e.map_w - 1u; ++i, ++col) if (xi_y <= e.col_cdf[i]) break;
        const float phi = (float)col * kPi * 2.f / (float)e.map_w;
        const float theta = (float)row * kPi / (float)e.map_h;
        const V3 lw = v3(sinf(theta) * sinf(kPi - phi), cosf(theta), sinf(theta) * cosf(kPi - phi));
        ret.wi = v3(dot(ld3(e.to_world), lw), dot(ld3(e.to_world + 3), lw), dot(ld3(e.to_world + 6), lw));
        ret.distance = kMaxDistance;
        ret.valid_distance = true;
        ret.radiance = tex_sample(e.radiance, v2(phi * 0.5f * kInvPi, theta * kInvPi)) * e.scale;
        ret.pdf = luminance(ret.radiance) * e.row_weight[row] * e.normalization / fmaxf(1e-4f, fabsf(sinf(theta)));
        if (ret.pdf < 0.f) ret.pdf = 0.f;
    }
}

// area / sphere: emit_pos, emit_n, emit_uv = hit geometry on the light, scatter_pos = previous vertex
PB_HD void emitter_eval(const EmitRec& e, EmitEval& ret, V3 emit_pos, V3 emit_n, V2 emit_uv, V3 scatter_pos) {
    ret.radiance = v3(0.f); ret.pdf = 0.f;   // EmitEvalRecord normalised to zero
    if (e.type == EM_TRI || e.type == EM_SPHERE) {  // area.h:37-45, sphere.h:34-42
        V3 dir = normalize(scatter_pos - emit_pos);
        float LNoL = dot(emit_n, dir);
        if (LNoL > 0.f) {
            float distance = length(scatter_pos - emit_pos);
            ret.pdf = distance * distance / (LNoL * e.area);
            ret.radiance = tex_sample(e.radiance, emit_uv);
        }
    } else if (e.type == EM_CONST_ENV) {  // env.h:82-85
        ret.pdf = 0.25f * kInvPi;
        ret.radiance = ld3(e.p0);
    } else if (e.type == EM_ENV_MAP) {  // env.h:52-64
        V3 dir = normalize(emit_pos - scatter_pos);
        dir = v3(dot(ld3(e.to_local), dir), dot(ld3(e.to_local + 3), dir), dot(ld3(e.to_local + 6), dir));
        const float phi = kPi - atan2f(dir.x, dir.z);
        const float theta = acosf(dir.y);
        const V2 tex = v2(phi * 0.5f * kInvPi, theta * kInvPi);
        uint32_t row = (uint32_t)(tex.y * (float)e.map_h);
        const uint32_t hi = e.map_h - 2u;
        row = row < hi ? row : hi;   // clamp(row, 0u, map_size.y - 2u)
        ret.radiance = tex_sample(e.radiance, tex) * e.scale;
        const float w0 = e.row_weight[row], w1 = e.row_weight[row + 1u], ft = tex.y * (float)e.map_h - 1.f * (float)row;
        ret.pdf = luminance(ret.radiance) * (w0 + ft * (w1 - w0)) * e.normalization / fmaxf(1e-4f, fabsf(sinf(theta)));
    }
}

PB_HD V3 emitter_radiance(const EmitRec& e, V2 uv) {  // emitter.h:53-71
    if (e.type == EM_TRI || e.type == EM_SPHERE || e.type == EM_ENV_MAP) return tex_sample(e.radiance, uv);
    if (e.type == EM_CONST_ENV) return ld3(e.p0);
    return v3(0.f);
}

}  // namespace pb
