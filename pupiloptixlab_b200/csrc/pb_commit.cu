// pb_commit.cu — pb_scene_commit / pb_update_instance / ray-cast / render / stats entry points.
#include <chrono>
#include <cstring>

#include "pb_host_prep.h"
#include "pb_internal.h"

using namespace pb;
void pb_free_device_scene(pb_ctx* c);

#define PB_REQUIRE(cond, msg)                       \
    do {                                            \
        if (!(cond)) { c->err = (msg); return PB_ERR_INVALID; } \
    } while (0)

template<class T>
static int upload(pb_ctx* c, T** dst, const std::vector<T>& src) {
    *dst = nullptr;
    PB_CUDA_TRY(c, cudaMalloc(dst, (src.empty() ? 1 : src.size()) * sizeof(T)));
    if (!src.empty()) PB_CUDA_TRY(c, cudaMemcpy(*dst, src.data(), src.size() * sizeof(T), cudaMemcpyHostToDevice));
    return PB_OK;
}

// device copy of a bitmap texture's texels (CudaTextureManager::GetCudaTexture, cuda/texture.cpp:60-102, keeps a
// cudaArray + texture object per bitmap; here a plain RGBA float array read by tex_bitmap_sample)
static const float* device_texels(void* user, const pb_texture& t) {
    pb_ctx* c = static_cast<pb_ctx*>(user);
    for (size_t i = 0; i < c->bitmaps.size() && i < c->d_bitmaps.size(); ++i)
        if (c->bitmaps[i].data() == t.bitmap) return c->d_bitmaps[i];
    return nullptr;
}

// EmitterHelper::AddAreaEmitter for one object (world/emitter.cpp:169-243): world-space records of its emissive triangles
// / its sphere, appended to `emit`; returns how many
static size_t append_area_emitters(pb_ctx* c, const HostObject& o, const ObjRec& r, std::vector<EmitRec>& emit) {
    const float select_weight = prep::emitter_weight(o.area_radiance);
    if (o.sphere) {  // SetSphereAreaEmitter, world/emitter.cpp:224-243
        EmitRec e{};
        e.type = EM_SPHERE;
        const float p0[3] = {0.f, 0.f, 0.f}, p1[3] = {0.f + 1.f, 0.f, 0.f};
        float ctr[3], px[3];
        prep::transform_point(r.M, p0, ctr);
        prep::transform_point(r.M, p1, px);
        for (int k = 0; k < 3; ++k) e.p0[k] = ctr[k];
        const float dx = ctr[0] - px[0], dy = ctr[1] - px[1], dz = ctr[2] - px[2];
        const float radius = sqrtf(dx * dx + dy * dy + dz * dz);
        e.p1[0] = radius;
        e.area = 4 * 3.14159265358979323846f * radius * radius;
        e.radiance = prep::tex_rec(o.area_radiance, device_texels(c, o.area_radiance));
        e.weight = select_weight * e.area;
        emit.push_back(e);
        return 1;
    }
    // SetMeshAreaEmitter, world/emitter.cpp:169-222
    const float* P = c->h_pos.data() + 3ull * o.vert_offset;
    const float* N = c->h_nrm.data() + 3ull * o.vert_offset;
    const float* T = c->h_uv.data() + 2ull * o.vert_offset;
    const uint32_t* I = c->h_idx.data() + 3ull * o.index_offset;
    for (uint32_t f = 0; f < o.n_tri; ++f) {
        EmitRec e{};
        e.type = EM_TRI;
        float* pp[3] = {e.p0, e.p1, e.p2};
        float* nn[3] = {e.n0, e.n1, e.n2};
        float* tt[3] = {e.t0, e.t1, e.t2};
        for (int k = 0; k < 3; ++k) {
            const uint32_t vi = I[3 * f + k];
            prep::transform_point(r.M, P + 3 * vi, pp[k]);
            prep::transform_normal(r.Minv, N + 3 * vi, nn[k]);
            tt[k][0] = T[2 * vi]; tt[k][1] = T[2 * vi + 1];
        }
        const float ax = e.p1[0] - e.p0[0], ay = e.p1[1] - e.p0[1], az = e.p1[2] - e.p0[2];
        const float bx = e.p2[0] - e.p0[0], by = e.p2[1] - e.p0[1], bz = e.p2[2] - e.p0[2];
        const float cx = ay * bz - az * by, cy = az * bx - ax * bz, cz = ax * by - ay * bx;
        e.area = sqrtf(cx * cx + cy * cy + cz * cz) * 0.5f;
        e.radiance = prep::tex_rec(o.area_radiance, device_texels(c, o.area_radiance));
        e.weight = select_weight * e.area;
        emit.push_back(e);
    }
    return o.n_tri;
}

// ComputeProbability (world/emitter.cpp:321-337) over the area emitters + the environment record; leaves `emit` with the
// environment record appended (slot n_area, type EM_NONE when absent) and the running sums of emitter.h:113-121 in `prefix`
static void finish_emitter_tables(pb_ctx* c, std::vector<EmitRec>& emit, EmitRec env, std::vector<float>& prefix) {
    float area_weight_sum = 0.f;
    for (auto& e : emit) area_weight_sum += e.weight;
    if (!emit.empty())
        for (auto& e : emit) e.select_probability = e.weight / area_weight_sum * emit.size();
    // a scene without any emitter is legal here (the reference would dereference an empty group): no NEE, paths end at
    // the primary vertex (pb_wavefront.cu has_emitters)
    const size_t emitter_num = (c->has_env ? 1 : 0) + emit.size();
    if (emitter_num) {
        for (auto& e : emit) e.select_probability = e.select_probability / emitter_num;
        env.select_probability = env.weight / emitter_num;
    }
    prefix.assign(emit.size(), 0.f);
    float sum_p = 0.f;
    for (size_t i = 0; i < emit.size(); ++i) { prefix[i] = sum_p + emit[i].select_probability; sum_p += emit[i].select_probability; }
    emit.push_back(env);
    c->stats.n_emitters = emitter_num;
}

// World::LoadScene -> RenderObject / EmitterHelper / ComputeProbability, then the device BVH build
static int commit(pb_ctx* c) {
    cudaSetDevice(c->device);
    wavefront_wait(c);          // asynchronous renders may still read the old scene
    // not committed until every table below exists: a failure half way (out of memory, bad environment map, BVH
    // limit) must leave the context refusing pb_render / pb_trace_* instead of launching on freed pointers
    c->committed = false;
    ++c->scene_epoch;           // graphs captured against the old scene hold its device pointers
    pb_free_device_scene(c);
    c->stats = pb_stats{};
    auto t0 = std::chrono::steady_clock::now();

    // ---- bitmap texels, then materials
    for (auto& hb : c->bitmaps) {
        float* d = nullptr;
        PB_CUDA_TRY(c, cudaMalloc(&d, hb.size() * sizeof(float)));
        c->d_bitmaps.push_back(d);
        PB_CUDA_TRY(c, cudaMemcpy(d, hb.data(), hb.size() * sizeof(float), cudaMemcpyHostToDevice));
    }
    std::vector<MatRec> mats;
    for (auto& m : c->materials) mats.push_back(prep::mat_rec(m, device_texels, c));
    // ---- objects + emitters (emitter_index_offset: example/path_tracer/pt_pass.cpp:159-170)
    c->h_objs.clear();
    std::vector<EmitRec> emit;
    uint32_t prim = 0;
    int emitter_offset = 0;
    for (auto& o : c->objects) {
        ObjRec r{};
        memcpy(r.M, o.M, 48);
        invert_affine(r.M, r.Minv);
        r.flags = (o.sphere ? OBJ_SPHERE : 0u) | ((o.flags & PB_SHAPE_FLIP_NORMALS) ? OBJ_FLIP_NORMALS : 0u) |
                  ((o.flags & PB_SHAPE_FLIP_TEX_COORDS) ? OBJ_FLIP_TEX : 0u) | (o.has_nrm ? OBJ_HAS_NORMALS : 0u) |
                  (o.has_uv ? OBJ_HAS_UV : 0u);
        r.material = o.material;
        r.prim_offset = prim;
        r.vert_offset = o.vert_offset; r.index_offset = o.index_offset;
        r.n_prims = o.sphere ? 1u : o.n_tri;
        r.emitter_offset = -1;
        if (o.is_emitter) {
            r.emitter_offset = emitter_offset;
            emitter_offset += (int)append_area_emitters(c, o, r, emit);
        }
        prim += r.n_prims;
        c->h_objs.push_back(r);
    }
    c->n_area_emitters = (uint32_t)emit.size();

    // ---- geometry upload + world-space primitives (gives the scene box needed by the env emitter)
    if (upload(c, &c->d_pos, c->h_pos) || upload(c, &c->d_nrm, c->h_nrm) || upload(c, &c->d_uv, c->h_uv) ||
        upload(c, &c->d_idx, c->h_idx) || upload(c, &c->d_objs, c->h_objs) || upload(c, &c->d_mats, mats))
        return PB_ERR_CUDA;
    PrimRec* d_prims_in = nullptr;
    PB_CUDA_TRY(c, cudaMalloc(&d_prims_in, (prim ? prim : 1) * sizeof(PrimRec)));
    int rc = make_world_prims(c, d_prims_in, c->stream);
    if (rc == PB_OK) rc = build_bvh(c, d_prims_in, prim, c->stream);
    cudaFree(d_prims_in);
    if (rc != PB_OK) return rc;

    // ---- environment emitter (AddEmitter, world/emitter.cpp:280-313) and select probabilities
    c->has_env = false;
    EmitRec env{};
    if (c->env.type == PB_ENV_CONSTANT) {
        env.type = EM_CONST_ENV;
        for (int k = 0; k < 3; ++k) env.p0[k] = c->env.radiance[k];
        V3 ctr = (c->scene_hi + c->scene_lo) * 0.5f;
        env.p1[0] = ctr.x; env.p1[1] = ctr.y; env.p1[2] = ctr.z;
        env.weight = 1.f;
        c->has_env = true;
    } else if (c->env.type == PB_ENV_MAP) {   // AddEmitter EnvMap + BuildEnvMapCdfTable, world/emitter.cpp:107-149,293-313
        const pb_texture& mt = c->env.map;
        PB_REQUIRE(mt.type == PB_TEX_BITMAP && mt.bitmap && mt.bitmap_w >= 2 && mt.bitmap_h >= 2, "environment map needs a bitmap of at least 2 x 2 texels");
        env.type = EM_ENV_MAP;
        env.radiance = prep::tex_rec(mt, device_texels(c, mt));
        env.scale = c->env.scale;
        V3 ctr = (c->scene_hi + c->scene_lo) * 0.5f;
        env.p1[0] = ctr.x; env.p1[1] = ctr.y; env.p1[2] = ctr.z;
        env.weight = 1.f;
        const float* tw = c->env.to_world;
        float inv[12];
        invert_affine(tw, inv);
        for (int r = 0; r < 3; ++r)
            for (int k = 0; k < 3; ++k) { env.to_world[3 * r + k] = tw[4 * r + k]; env.to_local[3 * r + k] = inv[4 * r + k]; }
        const size_t w = mt.bitmap_w, h = mt.bitmap_h;
        std::vector<float> tab((h + 1) + (w + 1) * h + h, 0.f);   // row_cdf | col_cdf | row_weight
        float* row_cdf = tab.data();
        float* col_cdf = row_cdf + (h + 1);
        float* row_weight = col_cdf + (w + 1) * h;
        size_t ci = 0, ri = 0;
        float row_sum = 0.f;
        row_cdf[ri++] = 0.f;
        for (size_t y = 0; y < h; ++y) {
            float col_sum = 0.f;
            col_cdf[ci++] = 0.f;
            for (size_t x = 0; x < w; ++x) {
                const float* px = mt.bitmap + 4 * (y * w + x);
                col_sum += prep::luminance(px);
                col_cdf[ci++] = col_sum;
            }
            for (size_t x = 1; x < w; ++x) col_cdf[ci - x - 1] /= col_sum;
            col_cdf[ci - 1] = 1.f;
            const float weight = std::sin((y + 0.5f) * 3.14159265358979323846f / h);
            row_weight[y] = weight;
            row_sum += col_sum * weight;
            row_cdf[ri++] = row_sum;
        }
        for (size_t y = 1; y < h; ++y) row_cdf[ri - y - 1] /= row_sum;
        row_cdf[ri - 1] = 1.f;
        env.normalization = 1.f / (row_sum * (2.f * 3.14159265358979323846f / w) * (3.14159265358979323846f / h));
        env.map_w = (uint32_t)w; env.map_h = (uint32_t)h; env.n_row_cdf = (uint32_t)(h + 1);
        if (upload(c, &c->d_env_tables, tab)) return PB_ERR_CUDA;
        env.row_cdf = c->d_env_tables; env.col_cdf = c->d_env_tables + (h + 1); env.row_weight = c->d_env_tables + (h + 1) + (w + 1) * h;
        c->has_env = true;
    }
    c->h_env_rec = env;
    std::vector<float> prefix;
    finish_emitter_tables(c, emit, env, prefix);
    if (upload(c, &c->d_emit, emit) || upload(c, &c->d_emit_prefix, prefix)) return PB_ERR_CUDA;

    c->stats.n_prims = prim; c->stats.n_nodes = c->n_nodes;
    c->stats.n_objects = (uint32_t)c->h_objs.size(); c->stats.bvh_depth = c->bvh_levels;
    c->stats.build_ms = std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - t0).count();
    c->committed = true;
    ++c->scene_epoch;
    c->refits = 0; c->rebuilds = 0;
    return PB_OK;
}

// pb_update_instance without a rebuild: new object record, refit (pb_build.cu refit_object), emitter tables rewritten in
// place.  *done = false -> the caller falls back to commit().
static int update_by_refit(pb_ctx* c, uint32_t object, bool* done) {
    *done = false;
    if (!c->committed || !c->allow_refit || object >= c->h_objs.size()) return PB_OK;
    cudaSetDevice(c->device);
    int rc = wavefront_wait(c);          // renders in flight still traverse the old boxes
    if (rc != PB_OK) return rc;
    auto t0 = std::chrono::steady_clock::now();
    const uint32_t launches0 = c->stats.kernel_launches;
    ObjRec& r = c->h_objs[object];
    memcpy(r.M, c->objects[object].M, 48);
    invert_affine(r.M, r.Minv);
    PB_CUDA_TRY(c, cudaMemcpyAsync(c->d_objs + object, &r, sizeof(ObjRec), cudaMemcpyHostToDevice, c->stream));
    rc = refit_object(c, object, done, c->stream);
    if (rc != PB_OK || !*done) return rc;
    if (c->objects[object].is_emitter) {   // ResetAreaEmitter + ComputeProbability (world.cpp:49-52): same count, same buffers
        std::vector<EmitRec> emit;
        for (size_t oi = 0; oi < c->objects.size(); ++oi)
            if (c->objects[oi].is_emitter) append_area_emitters(c, c->objects[oi], c->h_objs[oi], emit);
        if (emit.size() != c->n_area_emitters) { *done = false; return PB_OK; }
        std::vector<float> prefix;
        finish_emitter_tables(c, emit, c->h_env_rec, prefix);
        PB_CUDA_TRY(c, cudaMemcpyAsync(c->d_emit, emit.data(), emit.size() * sizeof(EmitRec), cudaMemcpyHostToDevice, c->stream));
        if (!prefix.empty()) PB_CUDA_TRY(c, cudaMemcpyAsync(c->d_emit_prefix, prefix.data(), prefix.size() * sizeof(float), cudaMemcpyHostToDevice, c->stream));
        PB_CUDA_TRY(c, cudaStreamSynchronize(c->stream));
    }
    c->stats.build_ms = std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - t0).count();
    c->stats.kernel_launches -= launches0;      // launches of this call, like commit() reports them
    ++c->refits;
    return PB_OK;
}

extern "C" {

PB_API int pb_scene_commit(pb_ctx* c) {
    if (!c) return PB_ERR_INVALID;
    PB_REQUIRE(c->in_scene, "pb_scene_commit without pb_scene_begin");
    return commit(c);
}

PB_API int pb_update_instance(pb_ctx* c, uint32_t object, const float to_world[12]) {
    if (!c) return PB_ERR_INVALID;
    PB_REQUIRE(object < c->objects.size(), "object id out of range");
    PB_REQUIRE(to_world != nullptr, "null transform");
    memcpy(c->objects[object].M, to_world, 48);
    // IAS::Update (world/ias_manager.cpp:116-151): refit in place; a rebuild only when the move changes the scene box enough
    // to change delta (the padding every accepted hit is defined against), or before the first commit
    bool done = false;
    int rc = update_by_refit(c, object, &done);
    if (rc != PB_OK) return rc;
    if (done) return PB_OK;
    rc = commit(c);
    if (rc == PB_OK) c->rebuilds = 1;
    return rc;
}

}  // extern "C"
