// ORACLE — test infrastructure, not product code.
//
// backend_ref.h — the "reference" shading backend: the reference's OWN headers
// (framework/cuda/random.h, optix/util.h, render/material/{fresnel,ggx}.h,
// render/material/bsdf/*.h, render/emitter.h + emitter/*.h, cuda/texture.h) are #included
// UNMODIFIED from /root/reference at build time and called through thin adapters.  Only built
// into oracle/_ref/ (see oracle/Makefile); never copied into this repository.
#pragma once
#ifdef ORC_WITH_REFERENCE

#include <stdlib.h>  // first: makes unqualified abs(float) resolve to the float overload
#include <cmath>
// Unqualified cos/sin/atan2/acos on a float must be the fp32 overload, as in device code (nvcc maps
// cos(float) to cosf); without these using-declarations g++ picks ::cos(double).  (<math.h> would
// do the same but also injects C++20 std::lerp, which collides with vec_math.h's lerp.)
using std::acos; using std::atan2; using std::cos; using std::sin; using std::tan; using std::atan;
#include <cuda_runtime.h>

// host stand-in for the CUDA texture fetch used by cuda::Texture::Sample (cuda/texture.h:52)
namespace orc { struct Texture; f3 sample_bitmap(const Texture& t, float u, float v); }
template<typename T>
inline T tex2D(cudaTextureObject_t obj, float u, float v) {
    const orc::Texture* t = reinterpret_cast<const orc::Texture*>(obj);
    orc::f3 c = orc::sample_bitmap(*t, u, v);
    return make_float4(c.x, c.y, c.z, 1.f);
}

#include "cuda/random.h"
#include "optix/util.h"
#include "render/material/fresnel.h"
#include "render/material/ggx.h"
#include "render/material/bsdf/bsdf.h"
#include "render/geometry.h"
#include "render/emitter.h"

namespace orc::refb {
namespace P = Pupil::optix;
namespace PM = Pupil::optix::material;

inline float3 cv(f3 v) { return make_float3(v.x, v.y, v.z); }
inline f3 cv(float3 v) { return {v.x, v.y, v.z}; }
inline float2 cv(f2 v) { return make_float2(v.x, v.y); }

inline Pupil::cuda::Texture cv_tex(const Texture& t) {
    Pupil::cuda::Texture r;
    r.type = static_cast<Pupil::util::ETextureType>(t.type);
    if (t.type == 0) r.rgb = cv(t.a);
    else if (t.type == 2) { r.patch1 = cv(t.a); r.patch2 = cv(t.b); }
    else r.bitmap = reinterpret_cast<cudaTextureObject_t>(&t);
    r.transform.r0 = make_float4(t.r0.x, t.r0.y, t.r0.z, t.r0.w);
    r.transform.r1 = make_float4(t.r1.x, t.r1.y, t.r1.z, t.r1.w);
    r.transform.r2 = make_float4(0, 0, 1, 0);
    r.transform.r3 = make_float4(0, 0, 0, 1);
    return r;
}

struct RngBridge {  // keeps one LCG state shared between orc::Rng and cuda::Random
    Pupil::cuda::Random r;
    Rng& o;
    explicit RngBridge(Rng& rng) : o(rng) { r.SetSeed(rng.s); }
    ~RngBridge() { o.s = r.GetSeed(); }
};

inline P::BsdfSamplingRecord to_ref(const BsdfRecord& r) {
    P::BsdfSamplingRecord q;
    q.wi = cv(r.wi); q.wo = cv(r.wo); q.f = cv(r.f); q.pdf = r.pdf;
    q.sampled_type = static_cast<P::EBsdfLobeType>(r.sampled_type);
    return q;
}
inline void from_ref(const P::BsdfSamplingRecord& q, BsdfRecord& r) {
    r.wi = cv(q.wi); r.f = cv(q.f); r.pdf = q.pdf; r.sampled_type = static_cast<uint32_t>(q.sampled_type);
}

template<typename F>
inline void with_local(const BsdfLocal& b, F&& fn) {
    switch (b.type) {
        case kDiffuse: { PM::Diffuse::Local l; l.reflectance = cv(b.reflectance); fn(l); } break;
        case kDielectric: { PM::Dielectric::Local l; l.eta = b.eta; l.specular_reflectance = cv(b.specular_reflectance);
            l.specular_transmittance = cv(b.specular_transmittance); fn(l); } break;
        case kRoughDielectric: { PM::RoughDielectric::Local l; l.alpha = b.alpha; l.eta = b.eta;
            l.specular_reflectance = cv(b.specular_reflectance); l.specular_transmittance = cv(b.specular_transmittance); fn(l); } break;
        case kConductor: { PM::Conductor::Local l; l.eta = cv(b.eta3); l.k = cv(b.k3);
            l.specular_reflectance = cv(b.specular_reflectance); fn(l); } break;
        case kRoughConductor: { PM::RoughConductor::Local l; l.alpha = b.alpha; l.eta = cv(b.eta3); l.k = cv(b.k3);
            l.specular_reflectance = cv(b.specular_reflectance); fn(l); } break;
        case kPlastic: { PM::Plastic::Local l; l.eta = b.eta; l.int_fdr = b.int_fdr;
            l.specular_sampling_weight = b.specular_sampling_weight; l.nonlinear = b.nonlinear;
            l.diffuse_reflectance = cv(b.reflectance); l.specular_reflectance = cv(b.specular_reflectance); fn(l); } break;
        case kRoughPlastic: { PM::RoughPlastic::Local l; l.eta = b.eta; l.int_fdr = b.int_fdr;
            l.specular_sampling_weight = b.specular_sampling_weight; l.alpha = b.alpha; l.nonlinear = b.nonlinear;
            l.diffuse_reflectance = cv(b.reflectance); l.specular_reflectance = cv(b.specular_reflectance); fn(l); } break;
    }
}

inline P::Emitter to_ref(const EmitterRec& e) {
    P::Emitter r;
    r.type = static_cast<P::EEmitterType>(e.type);
    r.weight = e.weight;
    r.select_probability = e.select_probability;
    switch (e.type) {
        case kEmTriArea:
            r.area.radiance = cv_tex(e.radiance); r.area.area = e.area;
            r.area.geo.v0.pos = cv(e.p0); r.area.geo.v1.pos = cv(e.p1); r.area.geo.v2.pos = cv(e.p2);
            r.area.geo.v0.normal = cv(e.n0); r.area.geo.v1.normal = cv(e.n1); r.area.geo.v2.normal = cv(e.n2);
            r.area.geo.v0.tex = cv(e.t0); r.area.geo.v1.tex = cv(e.t1); r.area.geo.v2.tex = cv(e.t2);
            break;
        case kEmSphere:
            r.sphere.radiance = cv_tex(e.radiance); r.sphere.area = e.area;
            r.sphere.geo.center = cv(e.center); r.sphere.geo.radius = e.radius;
            break;
        case kEmConstEnv:
            r.const_env.color = cv(e.color); r.const_env.center = cv(e.env_center);
            break;
        case kEmEnvMap:
            r.env_map.radiance = cv_tex(e.radiance); r.env_map.center = cv(e.env_center);
            r.env_map.map_size = make_uint2(e.map_w, e.map_h);
            r.env_map.row_cdf.SetData(reinterpret_cast<CUdeviceptr>(e.row_cdf), e.n_row_cdf);
            r.env_map.col_cdf.SetData(reinterpret_cast<CUdeviceptr>(e.col_cdf), (size_t)(e.map_w + 1) * e.map_h);
            r.env_map.row_weight.SetData(reinterpret_cast<CUdeviceptr>(e.row_weight), e.map_h);
            r.env_map.to_world.r0 = cv(e.tw0); r.env_map.to_world.r1 = cv(e.tw1); r.env_map.to_world.r2 = cv(e.tw2);
            r.env_map.to_local.r0 = cv(e.tl0); r.env_map.to_local.r1 = cv(e.tl1); r.env_map.to_local.r2 = cv(e.tl2);
            r.env_map.normalization = e.normalization; r.env_map.scale = e.scale;
            break;
        default: break;
    }
    return r;
}
}  // namespace orc::refb

namespace orc {
// The reference-headers backend: same interface as PortBackend (oracle.cpp).
struct RefBackend {
    static constexpr const char* name = "reference-headers";
    static void rng_init(Rng& r, uint32_t n, uint32_t a, uint32_t b) { Pupil::cuda::Random q; q.Init(n, a, b); r.s = q.GetSeed(); }
    static float rng_next(Rng& r) { refb::RngBridge br(r); return br.r.Next(); }
    static void bsdf_sample(const BsdfLocal& b, BsdfRecord& r, Rng& rng) {
        refb::RngBridge br(rng);
        auto q = refb::to_ref(r);
        q.sampler = &br.r;
        refb::with_local(b, [&](auto& l) { l.Sample(q); });
        refb::from_ref(q, r);
    }
    static void bsdf_eval(const BsdfLocal& b, BsdfRecord& r) {
        auto q = refb::to_ref(r);
        refb::with_local(b, [&](auto& l) { l.GetBsdf(q); l.GetPdf(q); });  // optix_material.h:57-66
        refb::from_ref(q, r);
    }
    static f3 tex(const Texture& t, f2 uv) { return refb::cv(refb::cv_tex(t).Sample(refb::cv(uv))); }
    static void emitter_sample_direct(const EmitterRec& e, EmitterSample& s, f3 hp, f3 hn, f2 xi) {
        auto re = refb::to_ref(e);
        Pupil::optix::EmitterSampleRecord rec;
        rec.is_delta = false;                              // never written by area/sphere (quirk 1)
        rec.distance = std::numeric_limits<float>::quiet_NaN();  // detect "left unwritten" (quirk 3)
        Pupil::optix::LocalGeometry g; g.position = refb::cv(hp); g.normal = refb::cv(hn); g.texcoord = make_float2(0, 0);
        re.SampleDirect(rec, g, refb::cv(xi));
        s.radiance = refb::cv(rec.radiance); s.wi = refb::cv(rec.wi); s.pos = refb::cv(rec.pos); s.normal = refb::cv(rec.normal);
        s.pdf = rec.pdf; s.is_delta = rec.is_delta;
        s.valid_distance = !(rec.distance != rec.distance);
        s.distance = s.valid_distance ? rec.distance : 0.f;
    }
    static void emitter_eval(const EmitterRec& e, EmitEval& ev, f3 ep, f3 en, f2 euv, f3 sp) {
        auto re = refb::to_ref(e);
        Pupil::optix::EmitEvalRecord rec; rec.radiance = make_float3(0, 0, 0); rec.pdf = 0.f;  // quirk 2
        Pupil::optix::LocalGeometry g; g.position = refb::cv(ep); g.normal = refb::cv(en); g.texcoord = refb::cv(euv);
        re.Eval(rec, g, refb::cv(sp));
        ev.radiance = refb::cv(rec.radiance); ev.pdf = rec.pdf;
    }
    static f3 emitter_radiance(const EmitterRec& e, f2 uv) { return refb::cv(refb::to_ref(e).GetRadiance(refb::cv(uv))); }
    static f3 to_local(f3 v, f3 n) { return refb::cv(Pupil::optix::ToLocal(refb::cv(v), refb::cv(n))); }
    static f3 to_world(f3 v, f3 n) { return refb::cv(Pupil::optix::ToWorld(refb::cv(v), refb::cv(n))); }
    static bool is_zero(f3 v) { return Pupil::optix::IsZero(refb::cv(v)); }
    static bool is_zero(float v) { return Pupil::optix::IsZero(v); }
    static float mis_weight(float a, float b) { return Pupil::optix::MISWeight(a, b); }
    static f2 sphere_texcoord(f3 p) { float2 t = Pupil::optix::GetSphereTexcoord(refb::cv(p)); return {t.x, t.y}; }
    static float fresnel_diffuse_reflectance(float eta) { return Pupil::optix::material::fresnel::DiffuseReflectance(eta); }
    static float luminance(f3 c) { return Pupil::optix::GetLuminance(refb::cv(c)); }
};
}  // namespace orc
#endif  // ORC_WITH_REFERENCE
