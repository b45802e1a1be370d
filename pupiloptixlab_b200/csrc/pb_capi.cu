// pb_capi.cu — the C ABI of include/pupil_b200.h: context lifetime and scene ingestion.
// (commit / trace / render entry points are in pb_commit.cu.)  There is no CPU fallback: without a
// CUDA device pb_create fails with PB_ERR_NO_DEVICE.
#include <cstdlib>
#include <cstring>

#include "pb_internal.h"

using namespace pb;

#define PB_REQUIRE(cond, msg)                       \
    do {                                            \
        if (!(cond)) { c->err = (msg); return PB_ERR_INVALID; } \
    } while (0)

extern "C" {

PB_API const char* pb_version(void) { return "pupil_b200 0.1 (sm_100a)"; }

PB_API int pb_create(pb_ctx** out, int device) {
    if (!out) return PB_ERR_INVALID;
    *out = nullptr;
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess || n <= 0 || device >= n || device < 0) return PB_ERR_NO_DEVICE;
    pb_ctx* c = new pb_ctx();
    c->device = device;
    *out = c;
    PB_CUDA_TRY(c, cudaSetDevice(device));
    PB_CUDA_TRY(c, cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
    PB_CUDA_TRY(c, cudaMalloc(&c->d_counters, 16 * sizeof(unsigned long long)));
    cudaDeviceProp prop;
    PB_CUDA_TRY(c, cudaGetDeviceProperties(&prop, device));
    c->sm_count = prop.multiProcessorCount;
    if (const char* e = getenv("PB_PRIM_ITERS")) { int v = atoi(e); if (v >= 1) c->prim_iters = (uint32_t)v; }
    if (const char* e = getenv("PB_PRIM_BATCH")) { int v = atoi(e); if (v >= 1 && v <= 32) c->prim_batch = (uint32_t)v; }
    if (const char* e = getenv("PB_LEAF_MAX")) { int v = atoi(e); if (v >= 1 && v <= 3) c->leaf_max = (uint32_t)v; }
    if (const char* e = getenv("PB_BOTTOM_MAX")) { int v = atoi(e); if (v >= 0 && v <= 24) c->bottom_max = (uint32_t)v; }
    if (const char* e = getenv("PB_PRIM_ITERS_ANY")) { int v = atoi(e); if (v >= 1) c->prim_iters_any = (uint32_t)v; }
    if (const char* e = getenv("PB_PRIM_BATCH_ANY")) { int v = atoi(e); if (v >= 1 && v <= 32) c->prim_batch_any = (uint32_t)v; }
    if (const char* e = getenv("PB_REFILL_ANY")) { int v = atoi(e); if (v >= 1 && v <= 32) c->refill_any = (uint32_t)v; }
    if (getenv("PB_NO_GRAPH")) c->use_graphs = false;
    if (getenv("PB_NO_REFIT")) c->allow_refit = false;
    if (const char* e = getenv("PB_REFILL")) { int v = atoi(e); if (v >= 1 && v <= 32) c->refill = (uint32_t)v; }
    return PB_OK;
}

static void free_device_scene(pb_ctx* c) {
    cudaFree(c->d_trishade); c->d_trishade = nullptr;
    cudaFree(c->d_pos); cudaFree(c->d_nrm); cudaFree(c->d_uv); cudaFree(c->d_idx);
    cudaFree(c->d_objs); cudaFree(c->d_mats); cudaFree(c->d_emit); cudaFree(c->d_emit_prefix); cudaFree(c->d_env_tables);
    cudaFree(c->d_nodes); cudaFree(c->d_prims);
    cudaFree(c->d_node_lo); cudaFree(c->d_node_hi); cudaFree(c->d_obj_box);
    c->d_node_lo = c->d_node_hi = nullptr; c->d_obj_box = nullptr;
    c->level_first.clear(); c->h_obj_lo.clear(); c->h_obj_hi.clear();
    c->d_pos = c->d_nrm = c->d_uv = nullptr; c->d_idx = nullptr; c->d_objs = nullptr; c->d_mats = nullptr;
    c->d_emit = nullptr; c->d_emit_prefix = nullptr; c->d_env_tables = nullptr; c->d_nodes = nullptr; c->d_prims = nullptr;
    for (float* b : c->d_bitmaps) cudaFree(b);
    c->d_bitmaps.clear();
    c->n_nodes = c->n_prims = 0;
}

PB_API void pb_destroy(pb_ctx* c) {
    if (!c) return;
    cudaSetDevice(c->device);
    wavefront_free(c);
    comm_free(c);
    free_device_scene(c);
    cudaFree(c->d_counters);
    cudaFree(c->d_stage);
    if (c->stream) cudaStreamDestroy(c->stream);
    delete c;
}

PB_API const char* pb_last_error(const pb_ctx* c) { return c ? c->err.c_str() : "null context"; }

PB_API int pb_scene_begin(pb_ctx* c) {
    if (!c) return PB_ERR_INVALID;
    cudaSetDevice(c->device);
    c->materials.clear(); c->objects.clear(); c->bitmaps.clear();
    c->h_pos.clear(); c->h_nrm.clear(); c->h_uv.clear(); c->h_idx.clear();
    c->env = pb_env{};
    c->committed = false; c->in_scene = true;
    return PB_OK;
}

static void own_bitmap(pb_ctx* c, pb_texture& t) {  // keep our own copy of caller texels
    if (t.type == PB_TEX_BITMAP && t.bitmap && t.bitmap_w && t.bitmap_h) {
        c->bitmaps.emplace_back(t.bitmap, t.bitmap + (size_t)4 * t.bitmap_w * t.bitmap_h);
        t.bitmap = c->bitmaps.back().data();
    } else {
        t.bitmap = nullptr;
    }
}

PB_API int pb_add_material(pb_ctx* c, const pb_material* m, uint32_t* out_id) {
    if (!c) return PB_ERR_INVALID;
    PB_REQUIRE(c->in_scene, "pb_add_material outside pb_scene_begin/pb_scene_commit");
    PB_REQUIRE(m && m->type >= 1 && m->type <= 7, "unknown material type");
    pb_material r = *m;
    own_bitmap(c, r.reflectance); own_bitmap(c, r.specular_reflectance); own_bitmap(c, r.specular_transmittance);
    own_bitmap(c, r.alpha); own_bitmap(c, r.eta); own_bitmap(c, r.k);
    c->materials.push_back(r);
    if (out_id) *out_id = (uint32_t)c->materials.size() - 1;
    return PB_OK;
}

PB_API int pb_add_trimesh(pb_ctx* c, const float* pos, const float* nrm, const float* uv, uint32_t n_vert,
                          const uint32_t* idx, uint32_t n_tri, const float to_world[12], uint32_t flags,
                          uint32_t material, const pb_texture* area_radiance, uint32_t* out_object) {
    if (!c) return PB_ERR_INVALID;
    PB_REQUIRE(c->in_scene, "pb_add_trimesh outside pb_scene_begin/pb_scene_commit");
    PB_REQUIRE(material < c->materials.size(), "material id out of range");
    PB_REQUIRE((pos && idx) || n_tri == 0, "null geometry pointers");
    for (uint64_t i = 0; i < 3ull * n_tri; ++i) PB_REQUIRE(idx[i] < n_vert, "vertex index out of range");
    HostObject o;
    o.sphere = false;
    memcpy(o.M, to_world, 48);
    o.flags = flags; o.material = material;
    o.vert_offset = (uint32_t)(c->h_pos.size() / 3); o.n_vert = n_vert;
    o.index_offset = (uint32_t)(c->h_idx.size() / 3); o.n_tri = n_tri;
    o.has_nrm = nrm != nullptr; o.has_uv = uv != nullptr;
    c->h_pos.insert(c->h_pos.end(), pos, pos + 3ull * n_vert);
    if (nrm) c->h_nrm.insert(c->h_nrm.end(), nrm, nrm + 3ull * n_vert); else c->h_nrm.resize(c->h_nrm.size() + 3ull * n_vert, 0.f);
    if (uv) c->h_uv.insert(c->h_uv.end(), uv, uv + 2ull * n_vert); else c->h_uv.resize(c->h_uv.size() + 2ull * n_vert, 0.f);
    c->h_idx.insert(c->h_idx.end(), idx, idx + 3ull * n_tri);
    if (area_radiance) { o.is_emitter = true; o.area_radiance = *area_radiance; own_bitmap(c, o.area_radiance); }
    c->objects.push_back(o);
    if (out_object) *out_object = (uint32_t)c->objects.size() - 1;
    return PB_OK;
}

PB_API int pb_add_sphere(pb_ctx* c, const float to_world[12], uint32_t flags, uint32_t material,
                         const pb_texture* area_radiance, uint32_t* out_object) {
    if (!c) return PB_ERR_INVALID;
    PB_REQUIRE(c->in_scene, "pb_add_sphere outside pb_scene_begin/pb_scene_commit");
    PB_REQUIRE(material < c->materials.size(), "material id out of range");
    HostObject o;
    o.sphere = true;
    memcpy(o.M, to_world, 48);
    o.flags = flags; o.material = material;
    if (area_radiance) { o.is_emitter = true; o.area_radiance = *area_radiance; own_bitmap(c, o.area_radiance); }
    c->objects.push_back(o);
    if (out_object) *out_object = (uint32_t)c->objects.size() - 1;
    return PB_OK;
}

PB_API int pb_set_env(pb_ctx* c, const pb_env* env) {
    if (!c) return PB_ERR_INVALID;
    PB_REQUIRE(c->in_scene, "pb_set_env outside pb_scene_begin/pb_scene_commit");
    c->env = *env;
    if (c->env.type == PB_ENV_MAP) own_bitmap(c, c->env.map);
    return PB_OK;
}

PB_API int pb_set_camera(pb_ctx* c, const float s2c[16], const float c2w[16]) {
    if (!c) return PB_ERR_INVALID;
    memcpy(c->cam.s2c, s2c, 64);
    memcpy(c->cam.c2w, c2w, 64);
    ++c->scene_epoch;
    return PB_OK;
}

}  // extern "C"

// shared with pb_commit.cu
void pb_free_device_scene(pb_ctx* c) { free_device_scene(c); }
