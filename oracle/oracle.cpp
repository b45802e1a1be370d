// ORACLE — test infrastructure, not product code.  See oracle/README.md.
//
// oracle.cpp — CPU restatement of the reference path tracer's hot path behind the SAME C ABI as
// the product (include/pupil_b200.h with the prefix orc_ instead of pb_):
//   * scene preparation: World::LoadScene / RenderObject / EmitterHelper / Material::LoadMaterial
//     (framework/world/world.cpp:101-139, world/emitter.cpp:169-337,
//      render/material/optix_material.cpp:9-132)
//   * closest-hit / any-hit queries that stand in for optixTrace (example/path_tracer/main.cu:77,158,
//     framework/render/emitter.h:91-100) with the intersection contract of orc_scene.h over an
//     independent binary BVH, plus a brute-force ground truth
//   * __raygen__main / __miss__default / __closesthit__default (example/path_tracer/main.cu:36-233),
//     Geometry::GetHitLocalGeometry (framework/render/geometry.h:48-96) and Material::GetLocalBsdf
//     (framework/render/material/optix_material.h:95-106)
// Shading arithmetic comes from a Backend: PortBackend = oracle/restated/orc_shading.h (always),
// RefBackend = the reference's own headers (only in oracle/_ref builds, backend_ref.h).
#include <omp.h>

#include <atomic>
#include <chrono>
#include <cstring>

#include "orc_scene.h"
#include "orc_cwbvh.h"
#include "backend_ref.h"
#include "../include/pupil_b200.h"

namespace orc {

// ---------------------------------------------------------------------------------- backends
struct PortBackend {
    static constexpr const char* name = "restated-port";
    static void rng_init(Rng& r, uint32_t n, uint32_t a, uint32_t b) { r.init(n, a, b); }
    static float rng_next(Rng& r) { return r.next(); }
    static void bsdf_sample(const BsdfLocal& b, BsdfRecord& r, Rng& rng) { orc::bsdf_sample(b, r, rng); }
    static void bsdf_eval(const BsdfLocal& b, BsdfRecord& r) { orc::bsdf_eval(b, r); }
    static f3 tex(const Texture& t, f2 uv) { return tex_sample(t, uv); }
    static void emitter_sample_direct(const EmitterRec& e, EmitterSample& s, f3 hp, f3 hn, f2 xi) { orc::emitter_sample_direct(e, s, hp, hn, xi); }
    static void emitter_eval(const EmitterRec& e, EmitEval& ev, f3 ep, f3 en, f2 euv, f3 sp) { orc::emitter_eval(e, ev, ep, en, euv, sp); }
    static f3 emitter_radiance(const EmitterRec& e, f2 uv) { return orc::emitter_radiance(e, uv); }
    static f3 to_local(f3 v, f3 n) { return orc::to_local(v, n); }
    static f3 to_world(f3 v, f3 n) { return orc::to_world(v, n); }
    static bool is_zero(f3 v) { return orc::is_zero(v); }
    static bool is_zero(float v) { return orc::is_zero(v); }
    static float mis_weight(float a, float b) { return orc::mis_weight(a, b); }
    static f2 sphere_texcoord(f3 p) { return orc::sphere_texcoord(p); }
    static float fresnel_diffuse_reflectance(float eta) { return orc::fresnel_diffuse_reflectance(eta); }
    static float luminance(f3 c) { return orc::luminance(c); }
};

// host stand-in for tex2D<float4> on a normalized-coordinate CUDA texture (cuda/texture.cpp:60-96)
f3 sample_bitmap(const Texture& t, float u, float v) {
    if (!t.bitmap || t.w == 0 || t.h == 0) return mk3(0.f);
    auto addr = [&](float x, int n) -> float {  // returns coordinate in [0, n)
        switch (t.address_mode) {
            case 0: x = x - floorf(x); return x * n;                                   // wrap
            case 2: { float f = floorf(x); float fr = x - f; if (((int)f) & 1) fr = 1.f - fr; return fr * n; }  // mirror
            default: return fminf(fmaxf(x, 0.f), 1.f) * n;                             // clamp / border
        }
    };
    auto texel = [&](int x, int y) -> f3 {
        auto wrapi = [&](int i, int n) {
            if (t.address_mode == 0) { i %= n; if (i < 0) i += n; return i; }
            return i < 0 ? 0 : (i >= n ? n - 1 : i);
        };
        x = wrapi(x, (int)t.w); y = wrapi(y, (int)t.h);
        const float* p = t.bitmap + 4 * ((size_t)y * t.w + x);
        return {p[0], p[1], p[2]};
    };
    float x = addr(u, (int)t.w), y = addr(v, (int)t.h);
    if (t.filter_mode == 0) return texel((int)floorf(x), (int)floorf(y));
    float xb = x - 0.5f, yb = y - 0.5f;
    int i = (int)floorf(xb), j = (int)floorf(yb);
    float a = xb - i, b = yb - j;
    return (1 - a) * (1 - b) * texel(i, j) + a * (1 - b) * texel(i + 1, j) + (1 - a) * b * texel(i, j + 1) +
           a * b * texel(i + 1, j + 1);
}

// ---------------------------------------------------------------------------------- binary BVH
struct BNode {
    f3 lo, hi;
    uint32_t left, right;  // inner: children; leaf: first, count
    uint32_t leaf;
};

struct Counters { uint64_t rays_closest = 0, rays_any = 0; CwCounts cw_closest, cw_any; uint64_t cw_mismatch = 0; };

struct Ctx {
    std::string err;
    int backend = 0;  // 0 port, 1 reference headers
    bool in_scene = false, committed = false;
    std::vector<Material> materials;
    std::vector<Object> objects;
    std::vector<std::unique_ptr<std::vector<float>>> bitmaps;
    pb_env env_desc{};
    bool has_env = false;
    Texture env_tex;
    float s2c[16]{}, c2w[16]{};
    // committed
    std::vector<Prim> prims;
    std::vector<f3> blo, bhi;         // B_in boxes (acceptance)
    f3 scene_lo{}, scene_hi{};
    float delta = 0;
    std::vector<BNode> nodes;
    std::vector<uint32_t> order;      // leaf prim order
    std::vector<EmitterRec> areas;
    std::vector<float> area_prefix;   // area_prefix[i] = running float sum after emitter i (emitter.h:115-121)
    EmitterRec env;
    std::vector<float> env_row_cdf, env_col_cdf, env_row_weight;
    pb_stats stats{};
    CwBvh cw;                         // the product's committed BVH, imported for the visit-count traversal (orc_cwbvh.h)
    uint64_t cw_mismatch = 0;         // rays whose CWBVH-traversal hit differed from the oracle's own BVH (must stay 0)
};

static Texture make_tex(Ctx* c, const pb_texture& t) {
    Texture r;
    r.type = t.type;
    r.a = {t.a[0], t.a[1], t.a[2]};
    r.b = {t.b[0], t.b[1], t.b[2]};
    r.r0 = {t.uv_r0[0], t.uv_r0[1], t.uv_r0[2], t.uv_r0[3]};
    r.r1 = {t.uv_r1[0], t.uv_r1[1], t.uv_r1[2], t.uv_r1[3]};
    if (t.type == PB_TEX_BITMAP && t.bitmap) {
        c->bitmaps.emplace_back(new std::vector<float>(t.bitmap, t.bitmap + (size_t)4 * t.bitmap_w * t.bitmap_h));
        r.bitmap = c->bitmaps.back()->data();
        r.w = t.bitmap_w; r.h = t.bitmap_h; r.address_mode = t.address_mode; r.filter_mode = t.filter_mode;
    }
    return r;
}

// optix_material.cpp:9-39 GetPixelAverage
static f3 pixel_average(const Texture& t) {
    switch (t.type) {
        case PB_TEX_RGB: return t.a;
        case PB_TEX_CHECKERBOARD: return mk3(t.a.x + t.b.x, t.a.y + t.b.y, t.a.z + t.b.z) * 0.5f;
        case PB_TEX_BITMAP: {
            float r = 0.f, g = 0.f, b = 0.f;
            for (size_t i = 0, idx = 0; i < (size_t)t.h * t.w; ++i) { r += t.bitmap[idx++]; g += t.bitmap[idx++]; b += t.bitmap[idx++]; idx++; }
            return mk3(r, g, b) / (1.f * t.h * t.w);
        }
    }
    return mk3(0.f);
}
// world/emitter.cpp:73-104 GetWeight
static float tex_weight(const Texture& t) {
    auto mx = [](float r, float g, float b) { return (r > g ? (r > b ? r : b) : (g > b ? g : b)); };
    switch (t.type) {
        case PB_TEX_RGB: return mx(t.a.x, t.a.y, t.a.z);
        case PB_TEX_CHECKERBOARD: return (mx(t.a.x, t.a.y, t.a.z) + mx(t.b.x, t.b.y, t.b.z)) * 0.5f;
        case PB_TEX_BITMAP: {
            float w = 0.f;  // (sic) the reference indexes (i*w + j) with i<w, j<h: emitter.cpp:91-99
            for (size_t i = 0; i < t.w; i++)
                for (size_t j = 0; j < t.h; j++) {
                    size_t k = (i * t.w + j) * 4;
                    if (k + 2 < (size_t)4 * t.w * t.h) w += mx(t.bitmap[k], t.bitmap[k + 1], t.bitmap[k + 2]);
                }
            return w / (1.f * t.w * t.h);
        }
    }
    return 0.f;
}

template<class B>
static void prepare_material(Material& m) {
    m.eta = m.int_ior / m.ext_ior;
    if (m.type == kPlastic || m.type == kRoughPlastic) {  // optix_material.cpp:95-99,112-116
        float dl = B::luminance(pixel_average(m.reflectance));
        float sl = B::luminance(pixel_average(m.specular_reflectance));
        m.specular_sampling_weight = sl / (sl + dl);
        m.int_fdr = B::fresnel_diffuse_reflectance(1.f / m.eta);
    }
}

// ---- BVH build (median split on the longest centroid axis, leaves <= 4) -------------------------
struct Builder {
    Ctx* c;
    std::vector<f3> cen, olo, ohi;  // centroids, B_out boxes
    uint32_t build(uint32_t first, uint32_t count, int depth) {
        uint32_t me;
#pragma omp critical(orc_nodes)
        { me = (uint32_t)c->nodes.size(); c->nodes.emplace_back(); }
        f3 lo = mk3(std::numeric_limits<float>::max()), hi = mk3(-std::numeric_limits<float>::max());
        f3 clo = lo, chi = hi;
        for (uint32_t i = first; i < first + count; ++i) {
            uint32_t p = c->order[i];
            lo = {fminf(lo.x, olo[p].x), fminf(lo.y, olo[p].y), fminf(lo.z, olo[p].z)};
            hi = {fmaxf(hi.x, ohi[p].x), fmaxf(hi.y, ohi[p].y), fmaxf(hi.z, ohi[p].z)};
            clo = {fminf(clo.x, cen[p].x), fminf(clo.y, cen[p].y), fminf(clo.z, cen[p].z)};
            chi = {fmaxf(chi.x, cen[p].x), fmaxf(chi.y, cen[p].y), fmaxf(chi.z, cen[p].z)};
        }
        BNode n;
        n.lo = lo; n.hi = hi;
        if (count <= 4) {
            n.leaf = 1; n.left = first; n.right = count;
#pragma omp critical(orc_nodes)
            c->nodes[me] = n;
            return me;
        }
        f3 e = chi - clo;
        int ax = e.x >= e.y ? (e.x >= e.z ? 0 : 2) : (e.y >= e.z ? 1 : 2);
        uint32_t mid = count / 2;
        auto key = [&](uint32_t p) { return ax == 0 ? cen[p].x : (ax == 1 ? cen[p].y : cen[p].z); };
        std::nth_element(c->order.begin() + first, c->order.begin() + first + mid, c->order.begin() + first + count,
                         [&](uint32_t a, uint32_t b) { float ka = key(a), kb = key(b); return ka < kb || (ka == kb && a < b); });
        uint32_t l = 0, r = 0;
        if (count > 200000 && depth < 6) {
#pragma omp task shared(l)
            l = build(first, mid, depth + 1);
#pragma omp task shared(r)
            r = build(first + mid, count - mid, depth + 1);
#pragma omp taskwait
        } else {
            l = build(first, mid, depth + 1);
            r = build(first + mid, count - mid, depth + 1);
        }
        n.leaf = 0; n.left = l; n.right = r;
#pragma omp critical(orc_nodes)
        c->nodes[me] = n;
        return me;
    }
};

static inline bool prim_accept(const Ctx* c, const Ray& r, const RayPre& pre, uint32_t p, float& t, float& u, float& v) {
    const Prim& pr = c->prims[p];
    if (pr.kind == 0) {
        if (!tri_test(r, pr.v0, pr.v1, pr.v2, t, u, v)) return false;
    } else {
        u = v = 0.f;
        if (!sphere_test(r, c->objects[pr.obj].Minv, t)) return false;
    }
    float N, F;
    slab(pre, c->blo[p], c->bhi[p], N, F);
    return N <= t && t <= F;
}

static Hit brute_closest(const Ctx* c, const Ray& r) {
    Hit best{r.tmax, 0, 0, kMiss};
    RayPre pre = ray_pre(r);
    for (uint32_t p = 0; p < c->prims.size(); ++p) {
        float t, u, v;
        if (prim_accept(c, r, pre, p, t, u, v) && (t < best.t || (t == best.t && p < best.prim))) best = {t, u, v, p};
    }
    return best;
}

template<bool ANY>
static Hit bvh_trace(const Ctx* c, const Ray& r) {
    Hit best{r.tmax, 0, 0, kMiss};
    if (c->nodes.empty()) return best;
    RayPre pre = ray_pre(r);
    uint32_t stack[128];
    int sp = 0;
    stack[sp++] = 0;
    while (sp) {
        const BNode& n = c->nodes[stack[--sp]];
        float N, F;
        slab(pre, n.lo, n.hi, N, F);
        if (!(fmaxf(N, r.tmin) <= fminf(F, best.t))) continue;
        if (n.leaf) {
            for (uint32_t i = n.left; i < n.left + n.right; ++i) {
                uint32_t p = c->order[i];
                float t, u, v;
                if (prim_accept(c, r, pre, p, t, u, v)) {
                    if (ANY) return Hit{t, u, v, p};
                    if (t < best.t || (t == best.t && p < best.prim)) best = {t, u, v, p};
                }
            }
        } else {
            const BNode& a = c->nodes[n.left];
            const BNode& b = c->nodes[n.right];
            float Na, Fa, Nb, Fb;
            slab(pre, a.lo, a.hi, Na, Fa);
            slab(pre, b.lo, b.hi, Nb, Fb);
            if (Na <= Nb) { stack[sp++] = n.right; stack[sp++] = n.left; }
            else { stack[sp++] = n.left; stack[sp++] = n.right; }
        }
    }
    return best;
}

// ordered, t-culled traversal of the imported 8-wide BVH with this scene's own acceptance test
template<bool ANY>
static Hit cw_trace_ctx(const Ctx* c, const Ray& r, CwCounts& cnt) {
    const RayPre pre = ray_pre(r);
    return cw_trace<ANY>(c->cw, r, cnt, [&](uint32_t gid, float& t, float& u, float& v) {
        return gid < c->prims.size() && prim_accept(c, r, pre, gid, t, u, v);
    });
}
static inline bool same_hit(const Hit& a, const Hit& b) {
    if (a.prim != b.prim) return false;
    return a.prim == kMiss || (f2u(a.t) == f2u(b.t) && f2u(a.u) == f2u(b.u) && f2u(a.v) == f2u(b.v));
}

// ---------------------------------------------------------------------------------- commit
template<class B>
static int commit(Ctx* c) {
    for (auto& m : c->materials) prepare_material<B>(m);
    // world-space primitives, global ids in insertion order
    c->prims.clear();
    c->areas.clear();
    uint32_t pid = 0;
    int emitter_offset = 0;
    for (uint32_t oi = 0; oi < c->objects.size(); ++oi) {
        Object& o = c->objects[oi];
        invert_affine(o.M, o.Minv);
        o.prim_offset = pid;
        if (o.sphere) {
            o.n_prims = 1;
            Prim p;
            p.v0 = {o.M[3], o.M[7], o.M[11]};
            p.v1 = {sqrtf(dot_fma(mk3(o.M[0], o.M[1], o.M[2]), mk3(o.M[0], o.M[1], o.M[2]))),
                    sqrtf(dot_fma(mk3(o.M[4], o.M[5], o.M[6]), mk3(o.M[4], o.M[5], o.M[6]))),
                    sqrtf(dot_fma(mk3(o.M[8], o.M[9], o.M[10]), mk3(o.M[8], o.M[9], o.M[10])))};
            p.v2 = mk3(0.f);
            p.obj = oi; p.kind = 1;
            c->prims.push_back(p);
        } else {
            o.n_prims = (uint32_t)o.idx.size() / 3;
            for (uint32_t f = 0; f < o.n_prims; ++f) {
                Prim p;
                f3 v[3];
                for (int k = 0; k < 3; ++k) {
                    uint32_t vi = o.idx[3 * f + k];
                    v[k] = xform_point_fma(o.M, mk3(o.pos[3 * vi], o.pos[3 * vi + 1], o.pos[3 * vi + 2]));
                }
                p.v0 = v[0]; p.v1 = v[1]; p.v2 = v[2]; p.obj = oi; p.kind = 0;
                c->prims.push_back(p);
            }
        }
        pid += o.n_prims;
        // emitters: pt_pass.cpp:159-170 (offset per emissive object), world/emitter.cpp:169-243
        o.emitter_offset = -1;
        if (o.is_emitter) {
            o.emitter_offset = emitter_offset;
            float select_weight = tex_weight(o.area_radiance);
            if (o.sphere) {
                EmitterRec e;
                e.type = kEmSphere;
                f3 ctr = transform_point(o.M, mk3(0.f, 0.f, 0.f));
                f3 px = transform_point(o.M, mk3(0.f + 1.f, 0.f, 0.f));
                e.center = ctr;
                e.radius = length(e.center - px);
                e.area = 4 * 3.14159265358979323846f * e.radius * e.radius;
                e.radiance = o.area_radiance;
                e.weight = select_weight * e.area;
                c->areas.push_back(e);
                emitter_offset += 1;
            } else {
                for (uint32_t f = 0; f < o.n_prims; ++f) {
                    EmitterRec e;
                    e.type = kEmTriArea;
                    f3 P[3], Nn[3];
                    f2 T[3];
                    for (int k = 0; k < 3; ++k) {
                        uint32_t vi = o.idx[3 * f + k];
                        P[k] = transform_point(o.M, mk3(o.pos[3 * vi], o.pos[3 * vi + 1], o.pos[3 * vi + 2]));
                        f3 n = o.nrm.empty() ? mk3(0.f) : mk3(o.nrm[3 * vi], o.nrm[3 * vi + 1], o.nrm[3 * vi + 2]);
                        f3 tn = transform_normal_raw(o.Minv, n);  // Transform::TransformNormal (transform.cpp:115-123)
                        float len = sqrtf(tn.x * tn.x + tn.y * tn.y + tn.z * tn.z);
                        Nn[k] = {tn.x / len, tn.y / len, tn.z / len};
                        T[k] = o.uv.empty() ? mk2(0.f, 0.f) : mk2(o.uv[2 * vi], o.uv[2 * vi + 1]);
                    }
                    e.p0 = P[0]; e.p1 = P[1]; e.p2 = P[2];
                    e.n0 = Nn[0]; e.n1 = Nn[1]; e.n2 = Nn[2];
                    e.t0 = T[0]; e.t1 = T[1]; e.t2 = T[2];
                    e.area = length(cross(e.p1 - e.p0, e.p2 - e.p0)) * 0.5f;
                    e.radiance = o.area_radiance;
                    e.weight = select_weight * e.area;
                    c->areas.push_back(e);
                }
                emitter_offset += (int)o.n_prims;
            }
        }
    }
    // exact scene box, delta, padded boxes
    const float fm = std::numeric_limits<float>::max();
    f3 slo = mk3(fm), shi = mk3(-fm);
    size_t n = c->prims.size();
    std::vector<f3> elo(n), ehi(n);
    for (size_t i = 0; i < n; ++i) {
        const Prim& p = c->prims[i];
        if (p.kind == 0) {
            elo[i] = {fminf(fminf(p.v0.x, p.v1.x), p.v2.x), fminf(fminf(p.v0.y, p.v1.y), p.v2.y), fminf(fminf(p.v0.z, p.v1.z), p.v2.z)};
            ehi[i] = {fmaxf(fmaxf(p.v0.x, p.v1.x), p.v2.x), fmaxf(fmaxf(p.v0.y, p.v1.y), p.v2.y), fmaxf(fmaxf(p.v0.z, p.v1.z), p.v2.z)};
        } else {
            elo[i] = p.v0 - p.v1;
            ehi[i] = p.v0 + p.v1;
        }
        slo = {fminf(slo.x, elo[i].x), fminf(slo.y, elo[i].y), fminf(slo.z, elo[i].z)};
        shi = {fmaxf(shi.x, ehi[i].x), fmaxf(shi.y, ehi[i].y), fmaxf(shi.z, ehi[i].z)};
    }
    c->scene_lo = slo; c->scene_hi = shi;
    float ext = n ? fmaxf(fmaxf(shi.x - slo.x, shi.y - slo.y), shi.z - slo.z) : 0.f;
    c->delta = ext * (1.0f / 262144.0f);  // 2^-18
    float d1 = c->delta, d2 = 2.0f * c->delta;
    c->blo.resize(n); c->bhi.resize(n);
    Builder b;
    b.c = c;
    b.cen.resize(n); b.olo.resize(n); b.ohi.resize(n);
    for (size_t i = 0; i < n; ++i) {
        c->blo[i] = {elo[i].x - d1, elo[i].y - d1, elo[i].z - d1};
        c->bhi[i] = {ehi[i].x + d1, ehi[i].y + d1, ehi[i].z + d1};
        b.olo[i] = {elo[i].x - d2, elo[i].y - d2, elo[i].z - d2};
        b.ohi[i] = {ehi[i].x + d2, ehi[i].y + d2, ehi[i].z + d2};
        b.cen[i] = (elo[i] + ehi[i]) * 0.5f;
    }
    c->order.resize(n);
    for (size_t i = 0; i < n; ++i) c->order[i] = (uint32_t)i;
    c->nodes.clear();
    c->nodes.reserve(n ? n : 1);
    if (n) {
#pragma omp parallel
#pragma omp single
        b.build(0, (uint32_t)n, 0);
    }
    // environment: world/emitter.cpp:280-313; World::GetAABB = union of transformed shape boxes
    // (world/world.cpp:223-229) — here the union of the exact primitive boxes (documented deviation
    // only relevant to env `center`, which never influences radiance).
    c->env = EmitterRec{};
    if (c->has_env && c->env_desc.type != PB_ENV_NONE) {
        f3 center = (shi + slo) * 0.5f;
        if (c->env_desc.type == PB_ENV_CONSTANT) {
            c->env.type = kEmConstEnv;
            c->env.color = {c->env_desc.radiance[0], c->env_desc.radiance[1], c->env_desc.radiance[2]};
            c->env.env_center = center;
            c->env.weight = 1.f;
        } else {
            c->env.type = kEmEnvMap;
            c->env.radiance = c->env_tex;
            c->env.scale = c->env_desc.scale;
            c->env.env_center = center;
            c->env.weight = 1.f;
            const float* tw = c->env_desc.to_world;
            c->env.tw0 = {tw[0], tw[1], tw[2]}; c->env.tw1 = {tw[4], tw[5], tw[6]}; c->env.tw2 = {tw[8], tw[9], tw[10]};
            float inv[12];
            invert_affine(tw, inv);
            c->env.tl0 = {inv[0], inv[1], inv[2]}; c->env.tl1 = {inv[4], inv[5], inv[6]}; c->env.tl2 = {inv[8], inv[9], inv[10]};
            // BuildEnvMapCdfTable, world/emitter.cpp:107-149
            size_t w = c->env_tex.w, h = c->env_tex.h;
            c->env_col_cdf.assign((w + 1) * h, 0.f);
            c->env_row_cdf.assign(h + 1, 0.f);
            c->env_row_weight.assign(h, 0.f);
            size_t ci = 0, ri = 0;
            float row_sum = 0.f;
            c->env_row_cdf[ri++] = 0.f;
            for (size_t y = 0; y < h; ++y) {
                float col_sum = 0.f;
                c->env_col_cdf[ci++] = 0.f;
                for (size_t x = 0; x < w; ++x) {
                    const float* px = c->env_tex.bitmap + 4 * (y * w + x);
                    col_sum += B::luminance(mk3(px[0], px[1], px[2]));
                    c->env_col_cdf[ci++] = col_sum;
                }
                for (size_t x = 1; x < w; ++x) c->env_col_cdf[ci - x - 1] /= col_sum;
                c->env_col_cdf[ci - 1] = 1.f;
                float weight = std::sin((y + 0.5f) * kPi / h);
                c->env_row_weight[y] = weight;
                row_sum += col_sum * weight;
                c->env_row_cdf[ri++] = row_sum;
            }
            for (size_t y = 1; y < h; ++y) c->env_row_cdf[ri - y - 1] /= row_sum;
            c->env_row_cdf[ri - 1] = 1.f;
            c->env.normalization = 1.f / (row_sum * (2.f * kPi / w) * (kPi / h));
            c->env.map_w = (uint32_t)w; c->env.map_h = (uint32_t)h;
            c->env.row_cdf = c->env_row_cdf.data(); c->env.n_row_cdf = (uint32_t)c->env_row_cdf.size();
            c->env.col_cdf = c->env_col_cdf.data();
            c->env.row_weight = c->env_row_weight.data();
        }
    }
    // EmitterHelper::ComputeProbability, world/emitter.cpp:321-337
    float area_weight_sum = 0.f;
    for (auto& e : c->areas) area_weight_sum += e.weight;
    if (!c->areas.empty())
        for (auto& e : c->areas) e.select_probability = e.weight / area_weight_sum * c->areas.size();
    size_t emitter_num = (c->env.type == kEmNone ? 0 : 1) + c->areas.size();
    for (auto& e : c->areas) e.select_probability = e.select_probability / emitter_num;
    c->env.select_probability = c->env.weight / emitter_num;
    // running float sums of EmitterGroup::SelectOneEmiiter (emitter.h:110-121)
    c->area_prefix.resize(c->areas.size());
    float sum_p = 0.f;
    for (size_t i = 0; i < c->areas.size(); ++i) { c->area_prefix[i] = sum_p + c->areas[i].select_probability; sum_p += c->areas[i].select_probability; }

    c->stats = pb_stats{};
    c->stats.n_prims = n; c->stats.n_nodes = c->nodes.size(); c->stats.n_emitters = c->areas.size() + (c->env.type != kEmNone);
    c->stats.n_objects = (uint32_t)c->objects.size();
    c->committed = true;
    return PB_OK;
}

// ---------------------------------------------------------------------------------- integrator
struct LocalGeo { f3 position, normal; f2 texcoord; };
struct HitInfo { LocalGeo geo; BsdfLocal bsdf; int emitter_index = -1; };

// Material::GetLocalBsdf + <Mat>::GetLocal (optix_material.h:95-106, bsdf/*.h GetLocal)
template<class B>
static BsdfLocal local_bsdf(const Material& m, f2 uv) {
    BsdfLocal b;
    b.type = m.type;
    switch (m.type) {
        case kDiffuse: b.reflectance = B::tex(m.reflectance, uv); break;
        case kDielectric:
            b.eta = m.int_ior / m.ext_ior;  // dielectric.h:50
            b.specular_reflectance = B::tex(m.specular_reflectance, uv);
            b.specular_transmittance = B::tex(m.specular_transmittance, uv);
            break;
        case kRoughDielectric:
            b.alpha = B::tex(m.alpha, uv).x; b.eta = m.eta;
            b.specular_reflectance = B::tex(m.specular_reflectance, uv);
            b.specular_transmittance = B::tex(m.specular_transmittance, uv);
            break;
        case kConductor:
            b.eta3 = B::tex(m.eta_tex, uv); b.k3 = B::tex(m.k_tex, uv);
            b.specular_reflectance = B::tex(m.specular_reflectance, uv);
            break;
        case kRoughConductor:
            b.alpha = B::tex(m.alpha, uv).x; b.eta3 = B::tex(m.eta_tex, uv); b.k3 = B::tex(m.k_tex, uv);
            b.specular_reflectance = B::tex(m.specular_reflectance, uv);
            break;
        case kPlastic:
            b.eta = m.eta; b.nonlinear = m.nonlinear; b.int_fdr = m.int_fdr;
            b.reflectance = B::tex(m.reflectance, uv); b.specular_reflectance = B::tex(m.specular_reflectance, uv);
            b.specular_sampling_weight = m.specular_sampling_weight;
            break;
        case kRoughPlastic:
            b.eta = m.eta; b.nonlinear = m.nonlinear; b.int_fdr = m.int_fdr;
            b.alpha = B::tex(m.alpha, uv).x;
            b.reflectance = B::tex(m.reflectance, uv); b.specular_reflectance = B::tex(m.specular_reflectance, uv);
            b.specular_sampling_weight = m.specular_sampling_weight;
            break;
    }
    return b;
}

// __closesthit__default (main.cu:216-230) + Geometry::GetHitLocalGeometry (geometry.h:48-96)
template<class B>
static void closest_hit_program(const Ctx* c, const Hit& h, f3 ray_o, f3 ray_d, HitInfo& out) {
    const Prim& pr = c->prims[h.prim];
    const Object& o = c->objects[pr.obj];
    const Material& m = c->materials[o.material];
    LocalGeo g;
    g.texcoord = mk2(0.f, 0.f);  // normalised: the reference leaves it stale when a mesh has no uvs
    uint32_t local_prim = h.prim - o.prim_offset;
    if (!o.sphere) {
        uint32_t i0 = o.idx[3 * local_prim], i1 = o.idx[3 * local_prim + 1], i2 = o.idx[3 * local_prim + 2];
        auto P = [&](uint32_t i) { return mk3(o.pos[3 * i], o.pos[3 * i + 1], o.pos[3 * i + 2]); };
        f3 p0 = P(i0), p1 = P(i1), p2 = P(i2);
        float bx = h.u, by = h.v;
        f3 pos = (1.f - bx - by) * p0 + bx * p1 + by * p2;
        g.position = {o.M[0] * pos.x + o.M[1] * pos.y + o.M[2] * pos.z + o.M[3],
                      o.M[4] * pos.x + o.M[5] * pos.y + o.M[6] * pos.z + o.M[7],
                      o.M[8] * pos.x + o.M[9] * pos.y + o.M[10] * pos.z + o.M[11]};
        f3 n;
        if (!o.nrm.empty()) {
            auto Nn = [&](uint32_t i) { return mk3(o.nrm[3 * i], o.nrm[3 * i + 1], o.nrm[3 * i + 2]); };
            n = (1.f - bx - by) * Nn(i0) + bx * Nn(i1) + by * Nn(i2);
        } else {
            n = cross(p1 - p0, p2 - p0);
        }
        g.normal = normalize(transform_normal_raw(o.Minv, n));
        if (o.flags & PB_SHAPE_FLIP_NORMALS) g.normal *= -1.f;
        if (!o.uv.empty()) {
            auto T = [&](uint32_t i) { return mk2(o.uv[2 * i], o.uv[2 * i + 1]); };
            g.texcoord = (1.f - bx - by) * T(i0) + bx * T(i1) + by * T(i2);
            if (o.flags & PB_SHAPE_FLIP_TEX_COORDS) g.texcoord.y = 1.f - g.texcoord.y;
        }
    } else {
        g.position = ray_o + h.t * ray_d;
        f3 lp = {o.Minv[0] * g.position.x + o.Minv[1] * g.position.y + o.Minv[2] * g.position.z + o.Minv[3],
                 o.Minv[4] * g.position.x + o.Minv[5] * g.position.y + o.Minv[6] * g.position.z + o.Minv[7],
                 o.Minv[8] * g.position.x + o.Minv[9] * g.position.y + o.Minv[10] * g.position.z + o.Minv[11]};
        f3 ctr = mk3(0.f);  // render_object.cpp:23
        g.texcoord = B::sphere_texcoord(normalize(lp - ctr));
        g.normal = normalize(transform_normal_raw(o.Minv, lp - ctr));
        if (o.flags & PB_SHAPE_FLIP_NORMALS) g.normal *= -1.f;
    }
    if (dot(-ray_d, g.normal) < 0.f && m.twosided) g.normal = -g.normal;  // geometry.h:92-96
    out.geo = g;
    out.emitter_index = o.emitter_offset >= 0 ? o.emitter_offset + (int)local_prim : -1;
    out.bsdf = local_bsdf<B>(m, g.texcoord);
}

// EmitterGroup::SelectOneEmiiter (emitter.h:110-136): first i with p <= running_sum_i, found by
// binary search over the identical float prefix sums.
static const EmitterRec& select_emitter(const Ctx* c, float p) {
    size_t n = c->areas.size();
    size_t lo = 0, hi = n;
    while (lo < hi) { size_t mid = (lo + hi) / 2; if (p <= c->area_prefix[mid]) hi = mid; else lo = mid + 1; }
    if (lo < n) return c->areas[lo];
    return c->env.type != kEmNone ? c->env : c->areas[n - 1];
}

template<class B>
static f3 path_radiance(const Ctx* c, const pb_render_params& rp, uint32_t x, uint32_t y, uint32_t seed,
                        f3& albedo, f3& normal, float& test, Counters& cnt) {
    const uint32_t w = rp.width, h = rp.height;
    const uint32_t pixel_index = y * w + x;
    Rng rng;
    B::rng_init(rng, 4, pixel_index, seed);  // main.cu:53
    f3 radiance = mk3(0.f), env_radiance = mk3(0.f), throughput = mk3(1.f);
    float env_pdf = 0.f;
    bool done = false;
    const bool count_cw = (rp.flags & PB_RENDER_COUNTED) && !c->cw.empty();   // visit counts on the imported product BVH

    float jx = B::rng_next(rng), jy = B::rng_next(rng);  // main.cu:55
    f2 subpixel = {(static_cast<float>(x) + jx) / static_cast<float>(w), (static_cast<float>(y) + jy) / static_cast<float>(h)};
    f4 pof = {subpixel.x, subpixel.y, 0.f, 1.f};
    auto row = [](const float* m, int r) { return f4{m[4 * r], m[4 * r + 1], m[4 * r + 2], m[4 * r + 3]}; };
    f4 d = {dot(row(c->s2c, 0), pof), dot(row(c->s2c, 1), pof), dot(row(c->s2c, 2), pof), dot(row(c->s2c, 3), pof)};
    { float inv = 1.0f / d.w; d.x *= inv; d.y *= inv; d.z *= inv; d.w *= inv; }  // d /= d.w
    d.w = 0.f;
    d = normalize(d);
    f3 ray_direction = normalize(mk3(dot(row(c->c2w, 0), d), dot(row(c->c2w, 1), d), dot(row(c->c2w, 2), d)));
    f3 ray_origin = {c->c2w[3], c->c2w[7], c->c2w[11]};

    HitInfo hit;
    auto trace_forward = [&](f3 o, f3 dir) {  // optixTrace type 0 + CH / miss programs
        Ray r{o, 0.001f, dir, 1e16f};
        cnt.rays_closest++;
        Hit hh = bvh_trace<false>(c, r);
        if (count_cw && !same_hit(cw_trace_ctx<false>(c, r, cnt.cw_closest), hh)) cnt.cw_mismatch++;
        if (hh.prim == kMiss) {  // __miss__default, main.cu:196-212
            if (c->env.type != kEmNone) {
                f3 rd = normalize(dir);
                EmitEval ev;
                B::emitter_eval(c->env, ev, o + rd, mk3(0.f), mk2(0.f, 0.f), o);
                env_radiance = ev.radiance;
                env_pdf = ev.pdf;
            }
            done = true;
        } else {
            closest_hit_program<B>(c, hh, o, dir, hit);
        }
    };
    trace_forward(ray_origin, ray_direction);

    int depth = 0;
    HitInfo local_hit = hit;
    if (!done) {
        if (hit.emitter_index >= 0) radiance += B::emitter_radiance(c->areas[local_hit.emitter_index], local_hit.geo.texcoord);  // main.cu:88-92
        albedo = bsdf_albedo(local_hit.bsdf);
        normal = local_hit.geo.normal;
    } else {
        albedo = mk3(0.f);
        normal = mk3(0.f);
    }
    test = B::rng_next(rng);  // test[pixel] = rng.Next(), main.cu:101
    // A scene without any emitter: the reference dereferences a null emitter group in the NEE block
    // (emitter.h:110-136); normalised on both sides to "the path ends at the primary vertex" (radiance 0 either way).
    const bool no_emitters = c->areas.empty() && c->env.type == kEmNone;

    while (!done) {
        ++depth;
        if (depth >= (int)rp.max_depth || no_emitters) break;
        float rr = depth > 2 ? 0.95 : 1.0;
        if (B::rng_next(rng) > rr) break;
        throughput /= rr;

        {  // direct light sampling, main.cu:113-140
            const EmitterRec& emitter = select_emitter(c, B::rng_next(rng));
            EmitterSample es;
            float e0 = B::rng_next(rng), e1 = B::rng_next(rng);
            B::emitter_sample_direct(emitter, es, local_hit.geo.position, local_hit.geo.normal, mk2(e0, e1));
            bool occluded = true;
            if (es.valid_distance) {  // quirk 3: the reference traces with an uninitialised tmax otherwise
                Ray sr{local_hit.geo.position, 0.001f, es.wi, es.distance - 0.001f};
                cnt.rays_any++;
                occluded = bvh_trace<true>(c, sr).prim != kMiss;
                if (count_cw && (cw_trace_ctx<true>(c, sr, cnt.cw_any).prim != kMiss) != occluded) cnt.cw_mismatch++;
            }
            if (!occluded) {
                BsdfRecord er;
                er.wi = B::to_local(es.wi, local_hit.geo.normal);
                er.wo = B::to_local(-ray_direction, local_hit.geo.normal);
                B::bsdf_eval(hit.bsdf, er);
                f3 f = er.f;
                float pdf = er.pdf;
                if (!B::is_zero(f * es.pdf)) {
                    float NoL = dot(local_hit.geo.normal, es.wi);
                    if (NoL > 0.f) {
                        float mis = es.is_delta ? 1.f : B::mis_weight(es.pdf, pdf);
                        es.pdf *= emitter.select_probability;
                        radiance += throughput * es.radiance * f * NoL * mis / es.pdf;
                    }
                }
            }
        }
        {  // bsdf sampling, main.cu:141-183
            BsdfRecord br;
            br.wo = B::to_local(-ray_direction, local_hit.geo.normal);
            B::bsdf_sample(hit.bsdf, br, rng);
            if (B::is_zero(br.f * fabsf(br.wi.z)) || B::is_zero(br.pdf)) break;
            throughput *= br.f * fabsf(br.wi.z) / br.pdf;
            ray_origin = hit.geo.position;
            ray_direction = B::to_world(br.wi, local_hit.geo.normal);
            trace_forward(ray_origin, ray_direction);
            if (done) {
                float mis = B::mis_weight(br.pdf, env_pdf);
                env_radiance *= throughput * mis;
                break;
            }
            local_hit = hit;
            if (hit.emitter_index >= 0) {
                const EmitterRec& emitter = c->areas[hit.emitter_index];
                EmitEval ev;
                B::emitter_eval(emitter, ev, hit.geo.position, hit.geo.normal, hit.geo.texcoord, ray_origin);
                if (!B::is_zero(ev.pdf)) {
                    float mis = (br.sampled_type & kDelta) ? 1.f : B::mis_weight(br.pdf, ev.pdf * emitter.select_probability);
                    radiance += throughput * ev.radiance * mis;
                }
            }
        }
    }
    radiance += env_radiance;
    return radiance;
}

// Window [x0, x0+ww) x [y0, y0+wh) of the film (the whole film by default): output buffers are window-sized,
// seeds stay those of the full film (pixel_index = y * width + x), so a crop of a large film costs only its pixels.
struct Window { uint32_t x0, y0, w, h, step = 1; };   // pixel (x0 + i * step, y0 + j * step), i < w, j < h
template<class B>
static int render(Ctx* c, const pb_render_params* rp, float* sum, float* albedo, float* normal, float* test = nullptr,
                  const Window* win = nullptr) {
    const Window full{0u, 0u, rp->width, rp->height};
    const Window W = win ? *win : full;
    size_t npx = (size_t)W.w * W.h;
    if (rp->flags & PB_RENDER_CLEAR) memset(sum, 0, npx * 16);
    uint32_t stride = rp->sample_stride ? rp->sample_stride : 1;
    auto t0 = std::chrono::steady_clock::now();
    uint64_t rc = 0, ra = 0, cn = 0, ct = 0, cs = 0, an = 0, at = 0, as = 0, mm = 0;
    for (uint32_t s = 0; s < rp->sample_count; ++s) {
        uint32_t seed = rp->first_sample + s * stride;
#pragma omp parallel for schedule(dynamic, 1) reduction(+ : rc, ra, cn, ct, cs, an, at, as, mm)
        for (int yy = 0; yy < (int)W.h; ++yy) {
            Counters cnt;
            for (uint32_t xx = 0; xx < W.w; ++xx) {
                f3 a, n;
                float tv = 0.f;
                f3 L = path_radiance<B>(c, *rp, W.x0 + xx * W.step, W.y0 + (uint32_t)yy * W.step, seed, a, n, tv, cnt);
                size_t p = (size_t)yy * W.w + xx;
                sum[4 * p + 0] += L.x; sum[4 * p + 1] += L.y; sum[4 * p + 2] += L.z; sum[4 * p + 3] += 1.f;
                if (albedo) { albedo[3 * p] = a.x; albedo[3 * p + 1] = a.y; albedo[3 * p + 2] = a.z; }
                if (normal) { normal[3 * p] = n.x; normal[3 * p + 1] = n.y; normal[3 * p + 2] = n.z; }
                if (test) test[p] = tv;
            }
            rc += cnt.rays_closest; ra += cnt.rays_any;
            cn += cnt.cw_closest.nodes; ct += cnt.cw_closest.tris; cs += cnt.cw_closest.spheres;
            an += cnt.cw_any.nodes; at += cnt.cw_any.tris; as += cnt.cw_any.spheres; mm += cnt.cw_mismatch;
        }
    }
    c->stats.node_visits = cn; c->stats.prim_tests_tri = ct; c->stats.prim_tests_sphere = cs;
    c->stats.node_visits_any = an; c->stats.prim_tests_tri_any = at; c->stats.prim_tests_sphere_any = as;
    c->cw_mismatch = mm;
    if (mm) { c->err = "the ordered traversal of the imported BVH disagrees with the oracle's own BVH on " + std::to_string(mm) + " rays"; return PB_ERR_INVALID; }
    auto t1 = std::chrono::steady_clock::now();
    c->stats.rays_closest = rc; c->stats.rays_any = ra;
    c->stats.samples = (uint64_t)npx * rp->sample_count;
    c->stats.last_render_ms = std::chrono::duration<float, std::milli>(t1 - t0).count();
    return PB_OK;
}

}  // namespace orc

// =================================================================================== C ABI (orc_*)
using namespace orc;
struct orc_ctx : Ctx {};
#define ORC_API extern "C" __attribute__((visibility("default")))
#define CHECK(cond, msg) do { if (!(cond)) { c->err = msg; return PB_ERR_INVALID; } } while (0)

ORC_API int orc_create(orc_ctx** out, int) { *out = new orc_ctx(); return PB_OK; }
ORC_API void orc_destroy(orc_ctx* c) { delete c; }
ORC_API const char* orc_last_error(const orc_ctx* c) { return c->err.c_str(); }
ORC_API const char* orc_version(void) {
#ifdef ORC_WITH_REFERENCE
    return "pupil-oracle 0.1 (restated port + reference headers)";
#else
    return "pupil-oracle 0.1 (restated port)";
#endif
}
ORC_API int orc_has_reference(void) {
#ifdef ORC_WITH_REFERENCE
    return 1;
#else
    return 0;
#endif
}
// 0 = restated port, 1 = the reference's own headers (oracle/_ref builds only)
ORC_API int orc_set_backend(orc_ctx* c, int backend) {
    if (backend == 1 && !orc_has_reference()) { c->err = "reference-header backend not built into this oracle"; return PB_ERR_INVALID; }
    c->backend = backend;
    return PB_OK;
}
ORC_API int orc_max_threads(void) { return omp_get_max_threads(); }
ORC_API void orc_set_threads(int n) { omp_set_num_threads(n); }

ORC_API int orc_scene_begin(orc_ctx* c) {
    c->materials.clear(); c->objects.clear(); c->bitmaps.clear(); c->has_env = false; c->committed = false; c->in_scene = true;
    return PB_OK;
}
ORC_API int orc_add_material(orc_ctx* c, const pb_material* m, uint32_t* out_id) {
    CHECK(c->in_scene, "orc_add_material outside scene_begin/commit");
    CHECK(m->type >= 1 && m->type <= 7, "unknown material type");
    Material r;
    r.type = m->type; r.twosided = m->twosided != 0; r.int_ior = m->int_ior; r.ext_ior = m->ext_ior; r.nonlinear = m->nonlinear != 0;
    r.reflectance = make_tex(c, m->reflectance);
    r.specular_reflectance = make_tex(c, m->specular_reflectance);
    r.specular_transmittance = make_tex(c, m->specular_transmittance);
    r.alpha = make_tex(c, m->alpha);
    r.eta_tex = make_tex(c, m->eta);
    r.k_tex = make_tex(c, m->k);
    c->materials.push_back(r);
    if (out_id) *out_id = (uint32_t)c->materials.size() - 1;
    return PB_OK;
}
ORC_API int orc_add_trimesh(orc_ctx* c, const float* pos, const float* nrm, const float* uv, uint32_t n_vert,
                            const uint32_t* idx, uint32_t n_tri, const float to_world[12], uint32_t flags,
                            uint32_t material, const pb_texture* area_radiance, uint32_t* out_object) {
    CHECK(c->in_scene, "orc_add_trimesh outside scene_begin/commit");
    CHECK(material < c->materials.size(), "material id out of range");
    for (uint32_t i = 0; i < 3 * n_tri; ++i) CHECK(idx[i] < n_vert, "vertex index out of range");
    Object o;
    o.sphere = false;
    memcpy(o.M, to_world, 48);
    o.flags = flags; o.material = material;
    o.pos.assign(pos, pos + 3 * (size_t)n_vert);
    if (nrm) o.nrm.assign(nrm, nrm + 3 * (size_t)n_vert);
    if (uv) o.uv.assign(uv, uv + 2 * (size_t)n_vert);
    o.idx.assign(idx, idx + 3 * (size_t)n_tri);
    if (area_radiance) { o.is_emitter = true; o.area_radiance = make_tex(c, *area_radiance); }
    c->objects.push_back(std::move(o));
    if (out_object) *out_object = (uint32_t)c->objects.size() - 1;
    return PB_OK;
}
ORC_API int orc_add_sphere(orc_ctx* c, const float to_world[12], uint32_t flags, uint32_t material,
                           const pb_texture* area_radiance, uint32_t* out_object) {
    CHECK(c->in_scene, "orc_add_sphere outside scene_begin/commit");
    CHECK(material < c->materials.size(), "material id out of range");
    Object o;
    o.sphere = true;
    memcpy(o.M, to_world, 48);
    o.flags = flags; o.material = material;
    if (area_radiance) { o.is_emitter = true; o.area_radiance = make_tex(c, *area_radiance); }
    c->objects.push_back(std::move(o));
    if (out_object) *out_object = (uint32_t)c->objects.size() - 1;
    return PB_OK;
}
ORC_API int orc_set_env(orc_ctx* c, const pb_env* env) {
    CHECK(c->in_scene, "orc_set_env outside scene_begin/commit");
    c->env_desc = *env;
    c->has_env = env->type != PB_ENV_NONE;
    if (env->type == PB_ENV_MAP) c->env_tex = make_tex(c, env->map);
    return PB_OK;
}
ORC_API int orc_set_camera(orc_ctx* c, const float s2c[16], const float c2w[16]) {
    memcpy(c->s2c, s2c, 64); memcpy(c->c2w, c2w, 64);
    return PB_OK;
}
ORC_API int orc_scene_commit(orc_ctx* c) {
    c->cw = CwBvh{};   // an imported product BVH belongs to the scene it was exported from
    CHECK(c->in_scene, "orc_scene_commit without scene_begin");
    auto t0 = std::chrono::steady_clock::now();
    int rc;
#ifdef ORC_WITH_REFERENCE
    rc = c->backend == 1 ? commit<RefBackend>(c) : commit<PortBackend>(c);
#else
    rc = commit<PortBackend>(c);
#endif
    c->stats.build_ms = std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - t0).count();
    c->in_scene = true;  // further set_camera / update_instance allowed
    return rc;
}
ORC_API int orc_update_instance(orc_ctx* c, uint32_t object, const float to_world[12]) {
    CHECK(object < c->objects.size(), "object id out of range");
    memcpy(c->objects[object].M, to_world, 48);
    return orc_scene_commit(c);
}
static Ray to_ray(const pb_ray& r) { return Ray{{r.ox, r.oy, r.oz}, r.tmin, {r.dx, r.dy, r.dz}, r.tmax}; }
ORC_API int orc_trace_closest(orc_ctx* c, const pb_ray* rays, uint64_t n, pb_hit* hits, int, void*) {
    CHECK(c->committed, "scene not committed");
    auto t0 = std::chrono::steady_clock::now();
#pragma omp parallel for schedule(dynamic, 1024)
    for (long long i = 0; i < (long long)n; ++i) {
        Hit h = bvh_trace<false>(c, to_ray(rays[i]));
        hits[i] = h.prim == kMiss ? pb_hit{rays[i].tmax, 0.f, 0.f, PB_MISS} : pb_hit{h.t, h.u, h.v, h.prim};
    }
    c->stats.rays_closest = n;
    c->stats.last_trace_ms = std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - t0).count();
    return PB_OK;
}
ORC_API int orc_trace_any(orc_ctx* c, const pb_ray* rays, uint64_t n, uint8_t* occluded, int, void*) {
    CHECK(c->committed, "scene not committed");
    auto t0 = std::chrono::steady_clock::now();
#pragma omp parallel for schedule(dynamic, 1024)
    for (long long i = 0; i < (long long)n; ++i) occluded[i] = bvh_trace<true>(c, to_ray(rays[i])).prim != kMiss;
    c->stats.rays_any = n;
    c->stats.last_trace_ms = std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - t0).count();
    return PB_OK;
}
// ---- the product's committed BVH (pb_export_bvh) read by the oracle: import, then an ordered t-culled traversal that
// yields the visit counts SURVEY 8(d) defines B_ray on.  Import checks the leaf records against this scene's own
// world-space primitives bit for bit (same ids, same vertices / centre + extents, every primitive exactly once).
ORC_API int orc_import_cwbvh(orc_ctx* c, const void* nodes80, uint64_t n_nodes, const void* prims48, uint64_t n_prims) {
    CHECK(c->committed, "scene not committed");
    CHECK(n_prims == c->prims.size(), "primitive count differs from the oracle's scene");
    CHECK((nodes80 && prims48) || n_prims == 0, "null BVH pointers");
    CwBvh& b = c->cw;
    b.nodes.assign((const uint8_t*)nodes80, (const uint8_t*)nodes80 + 80 * n_nodes);
    b.prims.assign((const float*)prims48, (const float*)prims48 + 12 * n_prims);
    b.n_nodes = n_nodes; b.n_prims = n_prims; b.n_leaf_slots = 0;
    std::vector<uint8_t> seen(n_prims, 0);
    for (uint64_t i = 0; i < n_prims; ++i) {
        const float* r = b.prims.data() + 12 * i;
        const uint32_t gid = f2u(r[3]), obj = f2u(r[7]), kind = f2u(r[11]);
        if (gid >= n_prims || seen[gid]) { b = CwBvh{}; c->err = "exported leaf records do not hold every primitive id exactly once"; return PB_ERR_INVALID; }
        seen[gid] = 1;
        const Prim& p = c->prims[gid];
        float want[9] = {p.v0.x, p.v0.y, p.v0.z, p.v1.x, p.v1.y, p.v1.z, p.v2.x, p.v2.y, p.v2.z};
        const float got[9] = {r[0], r[1], r[2], r[4], r[5], r[6], r[8], r[9], r[10]};
        uint32_t want_kind = p.kind;
        if (kind == 2u && p.kind == 1u && obj < c->objects.size()) {
            // "uniform sphere" record of the product: centre | (extent, 1/s, 0) | translation of world->object.  Legal only
            // where this scene's own world->object matrix is (1/s) * identity + translation; then it must carry those values.
            const float* I = c->objects[obj].Minv;
            const bool uniform = I[1] == 0.f && I[2] == 0.f && I[4] == 0.f && I[6] == 0.f && I[8] == 0.f && I[9] == 0.f && I[0] == I[5] && I[0] == I[10] &&
                                 p.v1.x == p.v1.y && p.v1.x == p.v1.z;
            if (uniform) { want_kind = 2u; want[4] = I[0]; want[5] = 0.f; want[6] = I[3]; want[7] = I[7]; want[8] = I[11]; }
        }
        if (kind != want_kind || obj != p.obj || memcmp(want, got, 36) != 0) {
            b = CwBvh{}; c->err = "exported primitive record " + std::to_string(i) + " differs from the oracle's world-space primitive " + std::to_string(gid);
            return PB_ERR_INVALID;
        }
    }
    // structure: every node reachable once, child / primitive ranges inside the arrays, leaves cover all primitives
    std::vector<uint8_t> node_seen(n_nodes, 0);
    std::vector<uint32_t> todo;
    uint64_t covered = 0;
    if (n_nodes) { todo.push_back(0); node_seen[0] = 1; }
    while (!todo.empty()) {
        const uint32_t ni = todo.back(); todo.pop_back();
        const CwNode n = cw_decode(b.nodes.data() + 80ull * ni);
        uint32_t rank = 0, pcount = 0;
        for (int s = 0; s < 8; ++s) {
            const uint32_t lc = cw_leaf_count(n.pmask, s);
            if ((n.imask >> s) & 1u) {
                const uint64_t ch = (uint64_t)n.child_base + rank++;
                if (lc || ch >= n_nodes || node_seen[ch]) { b = CwBvh{}; c->err = "exported BVH is not a tree"; return PB_ERR_INVALID; }
                node_seen[ch] = 1; todo.push_back((uint32_t)ch);
            } else if (lc) { pcount += lc; ++b.n_leaf_slots; }
        }
        if ((uint64_t)n.prim_base + pcount > n_prims) { b = CwBvh{}; c->err = "exported BVH leaf range outside the primitive array"; return PB_ERR_INVALID; }
        covered += pcount;
    }
    if (covered != n_prims) { b = CwBvh{}; c->err = "exported BVH leaves do not cover every primitive"; return PB_ERR_INVALID; }
    return PB_OK;
}
ORC_API uint64_t orc_cwbvh_leaf_slots(const orc_ctx* c) { return c->cw.n_leaf_slots; }
// out: pb_hit[n] (closest) or uint8[n] (any); counts[3] = nodes fetched, triangle records tested, sphere records tested
ORC_API int orc_trace_cwbvh(orc_ctx* c, const pb_ray* rays, uint64_t n, void* out, int any, uint64_t counts[3]) {
    CHECK(c->committed && !c->cw.empty(), "no imported BVH (orc_import_cwbvh)");
    uint64_t cn = 0, ct = 0, cs = 0;
#pragma omp parallel for schedule(dynamic, 1024) reduction(+ : cn, ct, cs)
    for (long long i = 0; i < (long long)n; ++i) {
        CwCounts cnt;
        const Ray r = to_ray(rays[i]);
        if (any) {
            ((uint8_t*)out)[i] = cw_trace_ctx<true>(c, r, cnt).prim != kMiss;
        } else {
            Hit h = cw_trace_ctx<false>(c, r, cnt);
            ((pb_hit*)out)[i] = h.prim == kMiss ? pb_hit{rays[i].tmax, 0.f, 0.f, PB_MISS} : pb_hit{h.t, h.u, h.v, h.prim};
        }
        cn += cnt.nodes; ct += cnt.tris; cs += cnt.spheres;
    }
    if (counts) { counts[0] = cn; counts[1] = ct; counts[2] = cs; }
    return PB_OK;
}
// ground truth: O(N) loop over all primitives per ray
ORC_API int orc_trace_closest_brute(orc_ctx* c, const pb_ray* rays, uint64_t n, pb_hit* hits) {
    CHECK(c->committed, "scene not committed");
#pragma omp parallel for schedule(dynamic, 16)
    for (long long i = 0; i < (long long)n; ++i) {
        Hit h = brute_closest(c, to_ray(rays[i]));
        hits[i] = h.prim == kMiss ? pb_hit{rays[i].tmax, 0.f, 0.f, PB_MISS} : pb_hit{h.t, h.u, h.v, h.prim};
    }
    return PB_OK;
}
ORC_API int orc_trace_any_brute(orc_ctx* c, const pb_ray* rays, uint64_t n, uint8_t* occluded) {
    CHECK(c->committed, "scene not committed");
#pragma omp parallel for schedule(dynamic, 16)
    for (long long i = 0; i < (long long)n; ++i) {
        Ray r = to_ray(rays[i]);
        RayPre pre = ray_pre(r);
        uint8_t occ = 0;
        for (uint32_t p = 0; p < c->prims.size() && !occ; ++p) { float t, u, v; occ = prim_accept(c, r, pre, p, t, u, v); }
        occluded[i] = occ;
    }
    return PB_OK;
}
ORC_API int orc_render(orc_ctx* c, const pb_render_params* p, float* sum, float* albedo, float* normal, void*) {
    CHECK(c->committed, "scene not committed");
    CHECK(p->width && p->height, "empty film");
#ifdef ORC_WITH_REFERENCE
    if (c->backend == 1) return render<RefBackend>(c, p, sum, albedo, normal);
#endif
    return render<PortBackend>(c, p, sum, albedo, normal);
}
// pb_render_ex of the header (adds the "test" buffer, pt_pass.cpp:144-146 / main.cu:101)
ORC_API int orc_render_ex(orc_ctx* c, const pb_render_params* p, const pb_render_targets* t, void*) {
    CHECK(c->committed, "scene not committed");
    CHECK(p && t && t->accum_sum_rgba, "null render params / targets");
    CHECK(p->width && p->height, "empty film");
#ifdef ORC_WITH_REFERENCE
    if (c->backend == 1) return render<RefBackend>(c, p, t->accum_sum_rgba, t->albedo_rgb, t->normal_rgb, t->test);
#endif
    return render<PortBackend>(c, p, t->accum_sum_rgba, t->albedo_rgb, t->normal_rgb, t->test);
}
// oracle only: a window of the film (window-sized output buffers, full-film seeds)
ORC_API int orc_render_window(orc_ctx* c, const pb_render_params* p, const pb_render_targets* t, uint32_t x0, uint32_t y0,
                              uint32_t w, uint32_t h, uint32_t step) {
    CHECK(c->committed, "scene not committed");
    CHECK(p && t && t->accum_sum_rgba, "null render params / targets");
    if (step == 0) step = 1;
    CHECK(w && h && x0 + (w - 1) * step < p->width && y0 + (h - 1) * step < p->height, "window outside the film");
    const Window win{x0, y0, w, h, step};
#ifdef ORC_WITH_REFERENCE
    if (c->backend == 1) return render<RefBackend>(c, p, t->accum_sum_rgba, t->albedo_rgb, t->normal_rgb, t->test, &win);
#endif
    return render<PortBackend>(c, p, t->accum_sum_rgba, t->albedo_rgb, t->normal_rgb, t->test, &win);
}
ORC_API int orc_resolve(orc_ctx*, const float* sum, float* frame, uint32_t n, int, void*) {
    for (uint32_t i = 0; i < n; ++i) {
        float cnt = sum[4 * i + 3];
        float inv = cnt > 0.f ? 1.0f / cnt : 0.f;
        frame[4 * i] = sum[4 * i] * inv; frame[4 * i + 1] = sum[4 * i + 1] * inv; frame[4 * i + 2] = sum[4 * i + 2] * inv; frame[4 * i + 3] = 1.f;
    }
    return PB_OK;
}
ORC_API int orc_get_stats(orc_ctx* c, pb_stats* out) { *out = c->stats; return PB_OK; }

// ---- known-answer probes (used to pin the restatement against the reference headers) -----------
// RNG stream: Init(4, v0, v1) then n draws (framework/cuda/random.h)
ORC_API int orc_rng_stream(orc_ctx* c, uint32_t v0, uint32_t v1, uint32_t n, uint32_t* seed_out, float* out) {
    Rng r;
#ifdef ORC_WITH_REFERENCE
    if (c->backend == 1) { RefBackend::rng_init(r, 4, v0, v1); *seed_out = r.s; for (uint32_t i = 0; i < n; ++i) out[i] = RefBackend::rng_next(r); return PB_OK; }
#endif
    (void)c;
    r.init(4, v0, v1); *seed_out = r.s;
    for (uint32_t i = 0; i < n; ++i) out[i] = r.next();
    return PB_OK;
}
// BSDF table: for material `mat` at uv=(0.25,0.75): for each (wo, rng seed) sample → (wi, f, pdf, lobe,
// seed after) then eval at (wo, wi) → (f, pdf).  out: 16 floats per entry.
template<class B>
static void bsdf_table(Ctx* c, uint32_t mat, const float* wo, const uint32_t* seeds, uint32_t n, float* out) {
    Material m = c->materials[mat];
    prepare_material<B>(m);
    BsdfLocal b = local_bsdf<B>(m, mk2(0.25f, 0.75f));
    for (uint32_t i = 0; i < n; ++i) {
        Rng rng; rng.s = seeds[i];
        BsdfRecord r; r.wo = {wo[3 * i], wo[3 * i + 1], wo[3 * i + 2]};
        B::bsdf_sample(b, r, rng);
        BsdfRecord e; e.wo = r.wo; e.wi = r.wi;
        B::bsdf_eval(b, e);
        float* o = out + 16 * i;
        o[0] = r.wi.x; o[1] = r.wi.y; o[2] = r.wi.z; o[3] = r.f.x; o[4] = r.f.y; o[5] = r.f.z; o[6] = r.pdf;
        o[7] = u2f(r.sampled_type); o[8] = u2f(rng.s);
        o[9] = e.f.x; o[10] = e.f.y; o[11] = e.f.z; o[12] = e.pdf; o[13] = o[14] = o[15] = 0.f;
    }
}
ORC_API int orc_bsdf_table(orc_ctx* c, uint32_t mat, const float* wo, const uint32_t* seeds, uint32_t n, float* out) {
    CHECK(mat < c->materials.size(), "material id out of range");
#ifdef ORC_WITH_REFERENCE
    if (c->backend == 1) { bsdf_table<RefBackend>(c, mat, wo, seeds, n, out); return PB_OK; }
#endif
    bsdf_table<PortBackend>(c, mat, wo, seeds, n, out);
    return PB_OK;
}
// Emitter table on the committed emitters: for emitter e, hit point/normal and xi → SampleDirect
// (radiance3, wi3, pdf, distance, valid) and Eval back from the sampled point (radiance3, pdf).
template<class B>
static void emitter_table(Ctx* c, uint32_t ei, const float* hp, const float* hn, const float* xi, uint32_t n, float* out) {
    const EmitterRec& e = ei < c->areas.size() ? c->areas[ei] : c->env;
    for (uint32_t i = 0; i < n; ++i) {
        EmitterSample s;
        f3 P = {hp[3 * i], hp[3 * i + 1], hp[3 * i + 2]}, N = {hn[3 * i], hn[3 * i + 1], hn[3 * i + 2]};
        B::emitter_sample_direct(e, s, P, N, mk2(xi[2 * i], xi[2 * i + 1]));
        EmitEval ev;
        B::emitter_eval(e, ev, s.pos, s.normal, mk2(0.5f, 0.5f), P);
        float* o = out + 16 * i;
        o[0] = s.radiance.x; o[1] = s.radiance.y; o[2] = s.radiance.z; o[3] = s.wi.x; o[4] = s.wi.y; o[5] = s.wi.z;
        o[6] = s.pdf; o[7] = s.distance; o[8] = s.valid_distance ? 1.f : 0.f;
        o[9] = ev.radiance.x; o[10] = ev.radiance.y; o[11] = ev.radiance.z; o[12] = ev.pdf;
        o[13] = e.select_probability; o[14] = e.area; o[15] = e.weight;
    }
}
ORC_API int orc_emitter_table(orc_ctx* c, uint32_t ei, const float* hp, const float* hn, const float* xi, uint32_t n, float* out) {
    CHECK(c->committed, "scene not committed");
    CHECK(ei < c->areas.size() || c->env.type != kEmNone, "emitter index out of range");
#ifdef ORC_WITH_REFERENCE
    if (c->backend == 1) { emitter_table<RefBackend>(c, ei, hp, hn, xi, n, out); return PB_OK; }
#endif
    emitter_table<PortBackend>(c, ei, hp, hn, xi, n, out);
    return PB_OK;
}
