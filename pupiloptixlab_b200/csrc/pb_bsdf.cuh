// pb_bsdf.cuh — the seven BSDFs, specialised at compile time on the material type MAT.
// Follows framework/render/material/bsdf/{diffuse,conductor,dielectric,rough_conductor,
// rough_dielectric,plastic,rough_plastic}.h (Local::GetBsdf / GetPdf / Sample); Eval = GetBsdf then
// GetPdf (optix_material.h:57-66).
#pragma once
#include "pb_shade.cuh"

namespace pb {

enum : uint32_t { LOBE_UNKNOWN = 0, LOBE_DIFFUSE_REFL = 1u << 1, LOBE_GLOSSY_REFL = 1u << 3, LOBE_GLOSSY_TRANS = 1u << 4,
                  LOBE_DELTA_REFL = 1u << 5, LOBE_DELTA_TRANS = 1u << 6, LOBE_DELTA = LOBE_DELTA_REFL | LOBE_DELTA_TRANS };
enum : int { MAT_DIFFUSE = 1, MAT_DIELECTRIC, MAT_ROUGH_DIELECTRIC, MAT_CONDUCTOR, MAT_ROUGH_CONDUCTOR, MAT_PLASTIC, MAT_ROUGH_PLASTIC };

// texture-resolved parameters at a hit (the union of the reference's <Material>::Local structs)
struct BsdfLocal {
    V3 reflectance, specular_reflectance, specular_transmittance, eta3, k3;
    float alpha, eta, int_fdr, specular_sampling_weight;
    bool nonlinear;
};
struct BsdfRec { V3 wi, wo, f; float pdf; uint32_t lobe; };

// <Mat>::GetLocal (bsdf/*.h) — only the fields MAT uses are fetched
template<int MAT>
PB_HD BsdfLocal bsdf_local(const MatRec& m, V2 uv) {
    BsdfLocal b{};
    if (MAT == MAT_DIFFUSE) b.reflectance = tex_sample(m.tex[TEX_REFLECTANCE], uv);
    if (MAT == MAT_DIELECTRIC) {
        b.eta = m.int_ior_over_ext_dielectric;
        b.specular_reflectance = tex_sample(m.tex[TEX_SPEC_REFL], uv);
        b.specular_transmittance = tex_sample(m.tex[TEX_SPEC_TRANS], uv);
    }
    if (MAT == MAT_ROUGH_DIELECTRIC) {
        b.alpha = tex_sample(m.tex[TEX_ALPHA], uv).x; b.eta = m.eta;
        b.specular_reflectance = tex_sample(m.tex[TEX_SPEC_REFL], uv);
        b.specular_transmittance = tex_sample(m.tex[TEX_SPEC_TRANS], uv);
    }
    if (MAT == MAT_CONDUCTOR || MAT == MAT_ROUGH_CONDUCTOR) {
        if (MAT == MAT_ROUGH_CONDUCTOR) b.alpha = tex_sample(m.tex[TEX_ALPHA], uv).x;
        b.eta3 = tex_sample(m.tex[TEX_ETA], uv); b.k3 = tex_sample(m.tex[TEX_K], uv);
        b.specular_reflectance = tex_sample(m.tex[TEX_SPEC_REFL], uv);
    }
    if (MAT == MAT_PLASTIC || MAT == MAT_ROUGH_PLASTIC) {
        b.eta = m.eta; b.nonlinear = m.nonlinear != 0u; b.int_fdr = m.int_fdr;
        if (MAT == MAT_ROUGH_PLASTIC) b.alpha = tex_sample(m.tex[TEX_ALPHA], uv).x;
        b.reflectance = tex_sample(m.tex[TEX_REFLECTANCE], uv);
        b.specular_reflectance = tex_sample(m.tex[TEX_SPEC_REFL], uv);
        b.specular_sampling_weight = m.specular_sampling_weight;
    }
    return b;
}
template<int MAT>
PB_HD V3 bsdf_albedo(const BsdfLocal& b) {  // optix_material.h:69-88
    return (MAT == MAT_DIFFUSE || MAT == MAT_PLASTIC || MAT == MAT_ROUGH_PLASTIC) ? b.reflectance : b.specular_reflectance;
}

PB_HD V3 plastic_diff(const BsdfLocal& b) {  // plastic.h:38, rough_plastic.h:44
    return b.reflectance / (1.f - (b.nonlinear ? b.reflectance * b.int_fdr : v3(b.int_fdr)));
}
PB_HD float plastic_spec_prob(const BsdfLocal& b, float fo) {  // plastic.h:47-49
    return (fo * b.specular_sampling_weight) / (fo * b.specular_sampling_weight + (1 - fo) * (1.f - b.specular_sampling_weight));
}

// GetBsdf
template<int MAT>
PB_HD void bsdf_f(const BsdfLocal& b, BsdfRec& r) {
    r.f = v3(0.f);
    if (MAT == MAT_DIFFUSE) {  // diffuse.h:12-17
        if (r.wi.z > 0.f && r.wo.z > 0.f) r.f = b.reflectance * kInvPi;
    } else if (MAT == MAT_ROUGH_CONDUCTOR) {  // rough_conductor.h:21-30
        if (r.wi.z <= 0.f || r.wo.z <= 0.f) return;
        V3 wh = normalize(r.wi + r.wo);
        V3 fo = fresnel_conductor(b.eta3, b.k3, dot(r.wo, wh));
        r.f = b.specular_reflectance * ggx_d(wh, b.alpha) * fo * ggx_g(r.wi, r.wo, b.alpha) / (4.f * r.wi.z * r.wo.z);
    } else if (MAT == MAT_ROUGH_DIELECTRIC) {  // rough_dielectric.h:21-50
        if (is_zero(r.wo.z)) return;
        V3 wh;
        const bool refl = r.wo.z * r.wi.z > 0.f;
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
            r.f = b.specular_transmittance * fabsf((1.f - F) * D * G * dot(r.wi, wh) * dot(r.wo, wh) / (sd * sd * r.wi.z * r.wo.z));
        }
    } else if (MAT == MAT_PLASTIC) {  // plastic.h:32-41
        if (r.wi.z <= 0.f || r.wo.z <= 0.f) return;
        float fo = fresnel_dielectric(b.eta, r.wo.z);
        float fi = fresnel_dielectric(b.eta, r.wi.z);
        V3 diff = plastic_diff(b);
        r.f = diff * (1.f - fi) * (1.f - fo) * cosine_hemisphere_pdf(r.wi) / (b.eta * b.eta * r.wi.z);
    } else if (MAT == MAT_ROUGH_PLASTIC) {  // rough_plastic.h:31-46
        if (r.wi.z <= 0.f || r.wo.z <= 0.f) return;
        float fo = fresnel_dielectric(b.eta, r.wo.z);
        V3 wh = normalize(r.wi + r.wo);
        r.f = b.specular_reflectance * fresnel_dielectric(b.eta, dot(wh, r.wo)) * ggx_d(wh, b.alpha) * ggx_g(r.wi, r.wo, b.alpha) /
              (4.f * r.wo.z * r.wi.z);
        float fi = fresnel_dielectric(b.eta, r.wi.z);
        V3 diff = plastic_diff(b);
        r.f += diff * (1.f - fi) * (1.f - fo) * kInvPi / (b.eta * b.eta);
    }
    // MAT_DIELECTRIC / MAT_CONDUCTOR: delta lobes evaluate to zero (dielectric.h:12-14, conductor.h:11-13)
}

// GetPdf
template<int MAT>
PB_HD void bsdf_pdf(const BsdfLocal& b, BsdfRec& r) {
    r.pdf = 0.f;
    if (MAT == MAT_DIFFUSE) {  // diffuse.h:19-24
        if (r.wi.z > 0.f && r.wo.z > 0.f) r.pdf = cosine_hemisphere_pdf(r.wi);
    } else if (MAT == MAT_ROUGH_CONDUCTOR) {  // rough_conductor.h:32-38
        if (r.wi.z <= 0.f || r.wo.z <= 0.f) return;
        V3 wh = normalize(r.wi + r.wo);
        wh = normalize(wh);
        r.pdf = ggx_pdf(r.wo, wh, b.alpha) / (4.f * dot(r.wo, wh));
    } else if (MAT == MAT_ROUGH_DIELECTRIC) {  // rough_dielectric.h:52-72
        const bool refl = r.wo.z * r.wi.z > 0.f;
        V3 wh;
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
        V3 wo = r.wo * (r.wo.z > 0.f ? 1.f : -1.f);
        float F = fresnel_dielectric(b.eta, dot(r.wo, wh));
        r.pdf = fabsf(ggx_pdf(wo, wh, b.alpha) * (refl ? F : 1.f - F) * dwh_dwo);
    } else if (MAT == MAT_PLASTIC) {  // plastic.h:43-53
        if (r.wi.z <= 0.f || r.wo.z <= 0.f) return;
        float fo = fresnel_dielectric(b.eta, r.wo.z);
        float sp = plastic_spec_prob(b, fo);
        r.pdf = cosine_hemisphere_pdf(r.wi) * (1.f - sp);
    } else if (MAT == MAT_ROUGH_PLASTIC) {  // rough_plastic.h:48-62
        if (r.wi.z <= 0.f || r.wo.z <= 0.f) return;
        float fo = fresnel_dielectric(b.eta, r.wo.z);
        float sp = plastic_spec_prob(b, fo);
        float dp = 1.f - sp;
        V3 wh = normalize(r.wi + r.wo);
        r.pdf = sp * ggx_pdf(r.wo, wh, b.alpha) / (4.f * dot(r.wi, wh));
        r.pdf += dp * cosine_hemisphere_pdf(r.wi);
    }
}

template<int MAT>
PB_HD void bsdf_eval(const BsdfLocal& b, BsdfRec& r) { bsdf_f<MAT>(b, r); bsdf_pdf<MAT>(b, r); }

// Sample.  r.wo set by the caller; r.f / r.pdf / r.lobe start at 0 / 0 / unknown (bsdf.h:26-37).
template<int MAT>
PB_HD void bsdf_sample(const BsdfLocal& b, BsdfRec& r, Rng& rng) {
    r.f = v3(0.f); r.pdf = 0.f; r.lobe = LOBE_UNKNOWN; r.wi = v3(0.f);
    if (MAT == MAT_DIFFUSE) {  // diffuse.h:26-32
        float x = rng.next(), y = rng.next();
        r.wi = cosine_sample_hemisphere(x, y);
        bsdf_pdf<MAT>(b, r); bsdf_f<MAT>(b, r);
        r.lobe = LOBE_DIFFUSE_REFL;
    } else if (MAT == MAT_CONDUCTOR) {  // conductor.h:19-26
        r.wi = reflect_z(r.wo);
        r.pdf = 1.f;
        V3 fr = fresnel_conductor(b.eta3, b.k3, r.wo.z);
        r.f = b.specular_reflectance * fr / fabsf(r.wi.z);
        r.lobe = LOBE_DELTA_REFL;
    } else if (MAT == MAT_DIELECTRIC) {  // dielectric.h:20-37
        float cos_t;
        float fr = fresnel_dielectric(b.eta, r.wo.z, cos_t);
        if (rng.next() < fr) {
            r.wi = reflect_z(r.wo);
            r.pdf = fr;
            r.f = b.specular_reflectance * fr / fabsf(r.wi.z);
            r.lobe = LOBE_DELTA_REFL;
        } else {
            r.wi = refract_z(r.wo, cos_t, b.eta);
            r.pdf = 1.f - fr;
            float factor = cos_t < 0.f ? 1.f / b.eta : b.eta;
            r.f = b.specular_transmittance * (1.f - fr) * factor * factor / fabsf(r.wi.z);
            r.lobe = LOBE_DELTA_TRANS;
        }
    } else if (MAT == MAT_ROUGH_CONDUCTOR) {  // rough_conductor.h:40-46
        float x = rng.next(), y = rng.next();
        r.wi = reflect_n(r.wo, ggx_sample(r.wo, b.alpha, x, y));
        bsdf_pdf<MAT>(b, r); bsdf_f<MAT>(b, r);
        r.lobe = LOBE_DIFFUSE_REFL;  // sic, rough_conductor.h:45
    } else if (MAT == MAT_ROUGH_DIELECTRIC) {  // rough_dielectric.h:74-97
        float x = rng.next(), y = rng.next();
        V3 wo = r.wo * (r.wo.z > 0.f ? 1.f : -1.f);
        V3 wh = ggx_sample(wo, b.alpha, x, y);
        float cos_t = 0.f;
        float F = fresnel_dielectric(b.eta, dot(r.wo, wh), cos_t);
        if (rng.next() < F) {
            r.wi = reflect_n(r.wo, wh);
            r.lobe = LOBE_GLOSSY_REFL;
        } else {
            if (is_zero(cos_t)) return;
            r.wi = refract_n(r.wo, wh, cos_t, b.eta);
            r.lobe = LOBE_GLOSSY_TRANS;
            if (r.wi.z * r.wo.z >= 0.f) return;
        }
        bsdf_pdf<MAT>(b, r); bsdf_f<MAT>(b, r);
    } else if (MAT == MAT_PLASTIC) {  // plastic.h:55-81
        if (r.wo.z <= 0.f) return;
        float fo = fresnel_dielectric(b.eta, r.wo.z);
        float x = rng.next(), y = rng.next();
        float sp = plastic_spec_prob(b, fo);
        if (x < sp) {
            r.lobe = LOBE_DELTA_REFL;
            r.wi = reflect_z(r.wo);
            r.f = b.specular_reflectance * fo / r.wi.z;
            r.pdf = sp;
        } else {
            r.lobe = LOBE_DIFFUSE_REFL;
            r.wi = cosine_sample_hemisphere((x - sp) / (1.f - sp), y);
            float fi = fresnel_dielectric(b.eta, r.wi.z);
            V3 diff = plastic_diff(b);
            r.f = diff * (1.f - fi) * (1.f - fo) * cosine_hemisphere_pdf(r.wi) / (b.eta * b.eta * r.wi.z);
            r.pdf = cosine_hemisphere_pdf(r.wi) * (1.f - sp);
        }
    } else if (MAT == MAT_ROUGH_PLASTIC) {  // rough_plastic.h:64-86
        if (r.wo.z <= 0.f) return;
        float fo = fresnel_dielectric(b.eta, r.wo.z);
        float sp = plastic_spec_prob(b, fo);
        float x = rng.next(), y = rng.next();
        if (y < sp) {
            y /= sp;
            V3 wh = ggx_sample(r.wo, b.alpha, x, y);
            r.wi = reflect_n(r.wo, wh);
            r.lobe = LOBE_GLOSSY_REFL;
        } else {
            y = (y - sp) / (1.f - sp);
            r.wi = cosine_sample_hemisphere(x, y);
            r.lobe = LOBE_DIFFUSE_REFL;
        }
        bsdf_pdf<MAT>(b, r); bsdf_f<MAT>(b, r);
    }
}

}  // namespace pb
