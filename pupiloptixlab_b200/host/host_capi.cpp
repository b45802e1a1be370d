// host_capi.cpp — the C entry points of libpupil_host.so (include/pupil_host.h).
#include <cstring>

#include "b200_pt_pass.h"
#include <cstdlib>
#include <cstring>

#include "image_io.h"
#include "scene_xml.h"

using namespace pupil::host;

struct ph_scene { Scene scene; };
static thread_local std::string g_err;

static int sink_begin(void* c) { return pb_scene_begin(static_cast<pb_ctx*>(c)); }
static int sink_mat(void* c, const pb_material* m, uint32_t* id) { return pb_add_material(static_cast<pb_ctx*>(c), m, id); }
static int sink_mesh(void* c, const float* p, const float* n, const float* uv, uint32_t nv, const uint32_t* idx, uint32_t nt,
                     const float tw[12], uint32_t fl, uint32_t mat, const pb_texture* rad, uint32_t* out) {
    return pb_add_trimesh(static_cast<pb_ctx*>(c), p, n, uv, nv, idx, nt, tw, fl, mat, rad, out);
}
static int sink_sphere(void* c, const float tw[12], uint32_t fl, uint32_t mat, const pb_texture* rad, uint32_t* out) {
    return pb_add_sphere(static_cast<pb_ctx*>(c), tw, fl, mat, rad, out);
}
static int sink_env(void* c, const pb_env* e) { return pb_set_env(static_cast<pb_ctx*>(c), e); }
static int sink_cam(void* c, const float a[16], const float b[16]) { return pb_set_camera(static_cast<pb_ctx*>(c), a, b); }
static int sink_commit(void* c) { return pb_scene_commit(static_cast<pb_ctx*>(c)); }
static const ph_sink g_default_sink = {sink_begin, sink_mat, sink_mesh, sink_sphere, sink_env, sink_cam, sink_commit};

extern "C" {

PH_API const ph_sink* ph_default_sink(void) { return &g_default_sink; }
PH_API const char* ph_last_error(void) { return g_err.c_str(); }

PH_API int ph_scene_load_xml(const char* path, ph_scene** out) {
    if (!path || !out) return PB_ERR_INVALID;
    auto s = std::make_unique<ph_scene>();
    if (!load_scene_xml_file(path, s->scene, g_err)) { *out = nullptr; return PB_ERR_INVALID; }
    *out = s.release();
    return PB_OK;
}
PH_API int ph_scene_load_xml_string(const char* xml, const char* root_dir, ph_scene** out) {
    if (!xml || !out) return PB_ERR_INVALID;
    auto s = std::make_unique<ph_scene>();
    if (!load_scene_xml_text(xml, root_dir ? root_dir : "", s->scene, g_err)) { *out = nullptr; return PB_ERR_INVALID; }
    *out = s.release();
    return PB_OK;
}
PH_API void ph_scene_free(ph_scene* s) { delete s; }

PH_API int ph_scene_info_get(const ph_scene* s, ph_scene_info* o) {
    if (!s || !o) return PB_ERR_INVALID;
    const Scene& sc = s->scene;
    *o = ph_scene_info{};
    o->max_depth = sc.max_depth; o->film_w = sc.film_w; o->film_h = sc.film_h;
    o->fov_y = sc.fov; o->near_clip = sc.near_clip; o->far_clip = sc.far_clip;
    o->n_shapes = static_cast<uint32_t>(sc.shape_instances.size());
    for (auto& i : sc.shape_instances) {
        if (i.is_emitter) ++o->n_emissive_shapes;
        if (i.type == ShapeType::Sphere) ++o->n_spheres; else o->n_triangles += static_cast<uint32_t>(i.mesh->indices.size() / 3);
    }
    o->has_env = sc.has_env ? sc.env.type : PB_ENV_NONE;
    o->n_warnings = static_cast<uint32_t>(sc.warnings.size());
    return PB_OK;
}
PH_API const char* ph_scene_warning(const ph_scene* s, uint32_t i) { return s && i < s->scene.warnings.size() ? s->scene.warnings[i].c_str() : ""; }

PH_API int ph_scene_camera(const ph_scene* s, float s2c[16], float c2w[16]) {
    if (!s) return PB_ERR_INVALID;
    const Camera cam = s->scene.camera();
    std::memcpy(s2c, cam.GetSampleToCameraMatrix().e, 64);
    std::memcpy(c2w, cam.GetToWorldMatrix().e, 64);
    return PB_OK;
}
PH_API int ph_scene_shape(const ph_scene* s, uint32_t i, uint32_t* type, float tw[12], pb_material* mat, uint32_t* flags,
                          uint32_t* is_emitter, pb_texture* radiance) {
    if (!s || i >= s->scene.shape_instances.size()) return PB_ERR_INVALID;
    const ShapeInstance& ins = s->scene.shape_instances[i];
    if (type) *type = static_cast<uint32_t>(ins.type);
    if (tw) std::memcpy(tw, ins.transform.matrix.e, 48);
    if (mat) *mat = ins.mat;
    if (flags) *flags = (ins.flip_normals ? PB_SHAPE_FLIP_NORMALS : 0u) | (ins.flip_tex_coords ? PB_SHAPE_FLIP_TEX_COORDS : 0u);
    if (is_emitter) *is_emitter = ins.is_emitter ? 1u : 0u;
    if (radiance) *radiance = ins.radiance;
    return PB_OK;
}
PH_API int ph_scene_upload(const ph_scene* s, const ph_sink* sink, void* ctx) {
    if (!s || !ctx) return PB_ERR_INVALID;
    g_err.clear();
    return upload_scene(s->scene, sink ? *sink : g_default_sink, ctx, g_err);
}

PH_API int ph_image_load(const char* path, float** rgba, uint32_t* w, uint32_t* h) {
    g_err.clear();
    if (!path || !rgba || !w || !h) { g_err = "null argument"; return PB_ERR_INVALID; }
    pupil::host::Image img;
    if (!pupil::host::load_image(path, img, g_err)) return PB_ERR_INVALID;
    float* p = static_cast<float*>(malloc(img.rgba.size() * sizeof(float)));
    if (!p) { g_err = "out of memory"; return PB_ERR_INVALID; }
    memcpy(p, img.rgba.data(), img.rgba.size() * sizeof(float));
    *rgba = p; *w = (uint32_t)img.w; *h = (uint32_t)img.h;
    return PB_OK;
}
PH_API void ph_image_free(float* rgba) { free(rgba); }
PH_API int ph_image_save(const char* path, const float* rgba, uint32_t w, uint32_t h) {
    g_err.clear();
    if (!path) { g_err = "null path"; return PB_ERR_INVALID; }
    return pupil::host::save_image(path, rgba, (int)w, (int)h, g_err) ? PB_OK : PB_ERR_INVALID;
}
PH_API int ph_image_save_png8(const char* path, const uint8_t* rgba8, uint32_t w, uint32_t h) {
    g_err.clear();
    if (!path || !rgba8 || !w || !h) { g_err = "null argument"; return PB_ERR_INVALID; }
    return pupil::host::save_png8(path, rgba8, (int)w, (int)h, g_err) ? PB_OK : PB_ERR_INVALID;
}

struct ph_session {
    Pupil::System system;
    std::unique_ptr<Pupil::pt::B200PTPass> pass;
};

PH_API int ph_session_create(const char* xml_path, int device, ph_session** out) {
    return ph_session_create_multi(xml_path, &device, 1, out);
}
PH_API int ph_session_set_frames_per_run(ph_session* s, uint32_t frames) {
    if (!s || !frames) return PB_ERR_INVALID;
    s->pass->frames_per_run = frames;
    return PB_OK;
}
PH_API int ph_session_create_multi(const char* xml_path, const int* devices, int n, ph_session** out) {
    using namespace Pupil;
    g_err.clear();
    if (!xml_path || !out || !devices || n < 1) return PB_ERR_INVALID;
    *out = nullptr;
    auto s = std::make_unique<ph_session>();
    s->system.Init(false);
    s->pass = std::make_unique<pt::B200PTPass>(&s->system, std::vector<int>(devices, devices + n));
    if (!s->pass->ok()) { g_err = s->pass->error(); return PB_ERR_NO_DEVICE; }
    s->system.AddPass(s->pass.get());
    if (!s->system.SetScene(xml_path) || !s->pass->ok()) {
        g_err = s->system.last_error.empty() ? s->pass->error() : s->system.last_error;
        return PB_ERR_INVALID;
    }
    *out = s.release();
    return PB_OK;
}
PH_API void ph_session_destroy(ph_session* s) {
    if (!s) return;
    s->pass.reset();          // the pass's buffers belong to the system's BufferManager: drop the pass first
    s->system.Destroy();
    delete s;
}
PH_API int ph_session_run(ph_session* s, uint32_t frames) {
    if (!s) return PB_ERR_INVALID;
    s->system.RunFrames(frames);
    if (!s->pass->ok()) { g_err = s->pass->error(); return PB_ERR_CUDA; }
    return PB_OK;
}
PH_API int ph_session_move_object(ph_session* s, uint32_t index, const float to_world[16]) {
    if (!s || !to_world) return PB_ERR_INVALID;
    auto ros = s->system.world.GetRenderobjects();
    if (index >= ros.size()) { g_err = "render object index out of range"; return PB_ERR_INVALID; }
    pupil::host::Transform t;
    std::memcpy(t.matrix.e, to_world, 64);
    ros[index]->UpdateTransform(t);
    return PB_OK;
}
PH_API int ph_session_read(ph_session* s, const char* name, void* dst, uint64_t bytes) {
    if (!s || !name || !dst) return PB_ERR_INVALID;
    Pupil::Buffer* b = s->system.buffers.GetBuffer(name);
    if (!b) { g_err = std::string("no buffer named '") + name + "'"; return PB_ERR_INVALID; }
    if (bytes != b->size_in_byte()) { g_err = "size mismatch for buffer '" + std::string(name) + "'"; return PB_ERR_INVALID; }
    if (cudaMemcpy(dst, b->cuda_ptr, bytes, cudaMemcpyDeviceToHost) != cudaSuccess) { g_err = "cudaMemcpy failed"; return PB_ERR_CUDA; }
    return PB_OK;
}
PH_API int ph_session_film(const ph_session* s, uint32_t* w, uint32_t* h) {
    if (!s || !s->system.world.scene) return PB_ERR_INVALID;
    if (w) *w = static_cast<uint32_t>(s->system.world.scene->film_w);
    if (h) *h = static_cast<uint32_t>(s->system.world.scene->film_h);
    return PB_OK;
}
PH_API uint32_t ph_session_sample_count(const ph_session* s) { return s ? s->pass->sample_cnt() : 0u; }

PH_API int ph_run_path_tracer(const char* xml_path, int device, uint32_t frames, float* final_rgba, float* albedo_rgb,
                              float* normal_rgb, uint32_t* out_w, uint32_t* out_h) {
    using namespace Pupil;
    g_err.clear();
    System system;
    system.Init(false);
    pt::B200PTPass pass(&system, device);
    if (!pass.ok()) { g_err = pass.error(); return PB_ERR_NO_DEVICE; }
    system.AddPass(&pass);
    if (!system.SetScene(xml_path) || !pass.ok()) { g_err = system.last_error.empty() ? pass.error() : system.last_error; return PB_ERR_INVALID; }
    system.RunFrames(frames);
    if (!pass.ok()) { g_err = pass.error(); return PB_ERR_CUDA; }
    const uint32_t w = static_cast<uint32_t>(system.world.scene->film_w), h = static_cast<uint32_t>(system.world.scene->film_h);
    if (out_w) *out_w = w;
    if (out_h) *out_h = h;
    auto fetch = [&](const char* name, float* dst, size_t bytes) {
        if (!dst) return true;
        Buffer* b = system.buffers.GetBuffer(name);
        return b && cudaMemcpy(dst, b->cuda_ptr, bytes, cudaMemcpyDeviceToHost) == cudaSuccess;
    };
    const size_t n = static_cast<size_t>(w) * h;
    if (!fetch("final result", final_rgba, n * 16) || !fetch("albedo", albedo_rgb, n * 12) || !fetch("normal", normal_rgb, n * 12)) {
        g_err = "reading back the output buffers failed";
        return PB_ERR_CUDA;
    }
    system.Destroy();
    return PB_OK;
}

}  // extern "C"
