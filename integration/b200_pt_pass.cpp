// integration/b200_pt_pass.cpp — see b200_pt_pass.h.  Links -lpupil_b200 -lcudart; nothing of framework/optix is used.
#include "b200_pt_pass.h"

#include <cstring>

namespace Pupil::pt {

namespace {
// util::Texture -> pb_texture: the same three variants; rows 0 and 1 of the uv transform are what cuda::Texture::Sample
// multiplies by (framework/cuda/texture.h:34-36); bitmap texels are the RGBA floats BitmapTexture::Load produced
// (framework/util/texture.cpp:87-148) and are copied by pb_add_material / pb_add_* / pb_set_env during the call.
pb_texture ToPb(const util::Texture& t) noexcept {
    pb_texture r{};
    std::memcpy(r.uv_r0, t.transform.matrix.e + 0, sizeof r.uv_r0);
    std::memcpy(r.uv_r1, t.transform.matrix.e + 4, sizeof r.uv_r1);
    switch (t.type) {
        case util::ETextureType::RGB:
            r.type = PB_TEX_RGB;
            r.a[0] = t.rgb.color.x; r.a[1] = t.rgb.color.y; r.a[2] = t.rgb.color.z;
            break;
        case util::ETextureType::Checkerboard:
            r.type = PB_TEX_CHECKERBOARD;
            r.a[0] = t.checkerboard.patch1.x; r.a[1] = t.checkerboard.patch1.y; r.a[2] = t.checkerboard.patch1.z;
            r.b[0] = t.checkerboard.patch2.x; r.b[1] = t.checkerboard.patch2.y; r.b[2] = t.checkerboard.patch2.z;
            break;
        case util::ETextureType::Bitmap:
            r.type = PB_TEX_BITMAP;
            r.bitmap = t.bitmap.data;
            r.bitmap_w = static_cast<uint32_t>(t.bitmap.w); r.bitmap_h = static_cast<uint32_t>(t.bitmap.h);
            r.address_mode = static_cast<uint32_t>(t.bitmap.address_mode);
            r.filter_mode = static_cast<uint32_t>(t.bitmap.filter_mode);
            break;
    }
    return r;
}

// resource::Material (a tagged union, material.h:66-84) -> pb_material (one flat struct with the union of the fields).
// EMatType's enumerators are numbered by decl/material_decl.inl:4-11 exactly as PB_MAT_* (Diffuse = 1 ... RoughPlastic = 7).
pb_material ToPb(const resource::Material& m) noexcept {
    pb_material r{};
    r.type = static_cast<uint32_t>(m.type);
    r.twosided = m.twosided ? 1u : 0u;
    switch (m.type) {
        case EMatType::Diffuse:
            r.reflectance = ToPb(m.diffuse.reflectance);
            break;
        case EMatType::Dielectric:
            r.int_ior = m.dielectric.int_ior; r.ext_ior = m.dielectric.ext_ior;
            r.specular_reflectance = ToPb(m.dielectric.specular_reflectance);
            r.specular_transmittance = ToPb(m.dielectric.specular_transmittance);
            break;
        case EMatType::RoughDielectric:
            r.int_ior = m.rough_dielectric.int_ior; r.ext_ior = m.rough_dielectric.ext_ior;
            r.alpha = ToPb(m.rough_dielectric.alpha);
            r.specular_reflectance = ToPb(m.rough_dielectric.specular_reflectance);
            r.specular_transmittance = ToPb(m.rough_dielectric.specular_transmittance);
            break;
        case EMatType::Conductor:
            r.eta = ToPb(m.conductor.eta); r.k = ToPb(m.conductor.k);
            r.specular_reflectance = ToPb(m.conductor.specular_reflectance);
            break;
        case EMatType::RoughConductor:
            r.alpha = ToPb(m.rough_conductor.alpha);
            r.eta = ToPb(m.rough_conductor.eta); r.k = ToPb(m.rough_conductor.k);
            r.specular_reflectance = ToPb(m.rough_conductor.specular_reflectance);
            break;
        case EMatType::Plastic:
            r.int_ior = m.plastic.int_ior; r.ext_ior = m.plastic.ext_ior; r.nonlinear = m.plastic.nonlinear ? 1u : 0u;
            r.reflectance = ToPb(m.plastic.diffuse_reflectance);
            r.specular_reflectance = ToPb(m.plastic.specular_reflectance);
            break;
        case EMatType::RoughPlastic:
            r.int_ior = m.rough_plastic.int_ior; r.ext_ior = m.rough_plastic.ext_ior; r.nonlinear = m.rough_plastic.nonlinear ? 1u : 0u;
            r.alpha = ToPb(m.rough_plastic.alpha);
            r.reflectance = ToPb(m.rough_plastic.diffuse_reflectance);
            r.specular_reflectance = ToPb(m.rough_plastic.specular_reflectance);
            break;
        default: break;
    }
    return r;
}
}  // namespace

B200PTPass::B200PTPass(std::string_view name, int device) noexcept : Pass(name) {
    // cuda::Context::Init + optix::Context::Init + the PTPass constructor (pt_pass.cpp:29-37) in one call; there is no CPU
    // fallback: without a CUDA device this fails and the pass stays inert
    if (pb_create(&m_ctx, device) != PB_OK) { m_error = "pb_create failed: no CUDA device"; m_ctx = nullptr; return; }
    cudaStreamCreateWithFlags(&m_stream, cudaStreamNonBlocking);
    BindingEventCallback();
}

B200PTPass::~B200PTPass() noexcept {
    if (m_stream) cudaStreamDestroy(m_stream);
    if (m_ctx) pb_destroy(m_ctx);
}

void B200PTPass::Fail(const char* what) noexcept {
    if (m_error.empty()) m_error = std::string(what) + ": " + (m_ctx ? pb_last_error(m_ctx) : "");
}

void B200PTPass::BindingEventCallback() noexcept {
    EventBinder<EWorldEvent::CameraChange>([this](void*) { m_dirty = true; });
    // the payload is the RenderObject whose transform changed (render_object.cpp:46-54 -> world.cpp:45-58); PTPass only sets
    // its dirty flag and re-fetches the IAS handle in OnRun (pt_pass.cpp:46,216-218) — here the object and its matrix are
    // noted and handed to pb_update_instance before the next render
    EventBinder<EWorldEvent::RenderInstanceUpdate>([this](void* p) {
        if (auto* ro = static_cast<world::RenderObject*>(p)) m_moved.emplace_back(ro, ro->transform.matrix);
        m_dirty = true;
    });
    EventBinder<ESystemEvent::SceneLoad>([this](void* p) { SetScene(static_cast<world::World*>(p)); });
}

void B200PTPass::SetScene(world::World* world) noexcept {
    if (!m_ctx || !world || !world->scene) return;
    m_world = world;
    resource::Scene& scene = *world->scene;

    // ---- World::LoadScene (world.cpp:101-139): one object per usable shape instance, in scene order — the order in which
    // World creates its RenderObjects, so object id == position in world->GetRenderobjects()
    if (pb_scene_begin(m_ctx) != PB_OK) return Fail("pb_scene_begin");
    for (const resource::ShapeInstance& ins : scene.shape_instances) {
        if (!ins.shape || ins.shape->type == resource::EShapeType::_unknown) continue;      // world.cpp:122
        const pb_material mat = ToPb(ins.mat);
        uint32_t mat_id = 0;
        if (pb_add_material(m_ctx, &mat, &mat_id) != PB_OK) return Fail("pb_add_material");
        uint32_t flags = 0u;
        if (ins.flip_normals) flags |= PB_SHAPE_FLIP_NORMALS;
        if (ins.flip_tex_coords) flags |= PB_SHAPE_FLIP_TEX_COORDS;
        pb_texture radiance{};
        if (ins.is_emitter) radiance = ToPb(ins.emitter.area.radiance);                      // resource::AreaEmitter, emitter.h:16-18
        const pb_texture* rad = ins.is_emitter ? &radiance : nullptr;
        const float* to_world = ins.transform.matrix.e;                                      // rows 0..2 of the row-major Mat4
        int rc;
        if (ins.shape->type == resource::EShapeType::_sphere) {
            // ShapeLoader<_sphere> already folded centre and radius into ins.transform (shape.cpp:106-125): unit sphere
            rc = pb_add_sphere(m_ctx, to_world, flags, mat_id, rad, nullptr);
        } else {
            const resource::Mesh& mesh = ins.shape->mesh;                                    // host arrays kept by ShapeManager
            rc = pb_add_trimesh(m_ctx, mesh.positions, mesh.normals, mesh.texcoords, mesh.vertex_num, mesh.indices, mesh.face_num,
                                to_world, flags, mat_id, rad, nullptr);
        }
        if (rc != PB_OK) return Fail("pb_add_sphere / pb_add_trimesh");
    }
    // ---- environment emitters (world.cpp:133-135 -> EmitterHelper::AddEmitter, world/emitter.cpp:280-313)
    for (const resource::Emitter& e : scene.emitters) {
        pb_env env{};
        if (e.type == resource::EEmitterType::ConstEnv) {
            env.type = PB_ENV_CONSTANT;
            env.radiance[0] = e.const_env.radiance.x; env.radiance[1] = e.const_env.radiance.y; env.radiance[2] = e.const_env.radiance.z;
        } else if (e.type == resource::EEmitterType::EnvMap) {
            env.type = PB_ENV_MAP;
            env.scale = e.env_map.scale;
            std::memcpy(env.to_world, e.env_map.transform.matrix.e, sizeof env.to_world);
            env.map = ToPb(e.env_map.radiance);
        } else {
            continue;       // point / directional emitters are not on the path tracer's path (emitter.h:110-136 scans them, C1-C5 have none)
        }
        if (pb_set_env(m_ctx, &env) != PB_OK) return Fail("pb_set_env");
    }
    // ---- camera (CameraHelper::GetCudaMemory, world/camera.cpp:72-92) and commit (emitter tables + device BVH build)
    const util::Mat4 s2c = world->camera->GetSampleToCameraMatrix(), c2w = world->camera->GetToWorldMatrix();
    if (pb_set_camera(m_ctx, s2c.e, c2w.e) != PB_OK) return Fail("pb_set_camera");
    if (pb_scene_commit(m_ctx) != PB_OK) return Fail("pb_scene_commit");

    // ---- launch config and output buffers (pt_pass.cpp:110-147)
    m_w = static_cast<uint32_t>(scene.sensor.film.w);
    m_h = static_cast<uint32_t>(scene.sensor.film.h);
    m_max_depth = scene.integrator.max_depth;
    auto buf_mngr = util::Singleton<BufferManager>::instance();
    m_frame = buf_mngr->GetBuffer(buf_mngr->DEFAULT_FINAL_RESULT_BUFFER_NAME);
    BufferDesc desc{ "pt accum buffer", EBufferFlag::None, m_w, m_h, static_cast<uint32_t>(sizeof(float) * 4) };
    m_accum = buf_mngr->AllocBuffer(desc);            // holds SUMS in .xyz and the frame count in .w (pb_resolve divides)
    desc.name = "albedo"; desc.flag = EBufferFlag::AllowDisplay; desc.stride_in_byte = sizeof(float) * 3;
    m_albedo = buf_mngr->AllocBuffer(desc);
    desc.name = "normal";
    m_normal = buf_mngr->AllocBuffer(desc);
    desc.name = "test"; desc.stride_in_byte = sizeof(float);
    m_test = buf_mngr->AllocBuffer(desc);
    m_moved.clear();
    m_dirty = true;
}

void B200PTPass::OnRun() noexcept {
    if (!m_ctx || !m_world || !m_error.empty()) return;
    if (m_dirty) {
        // moved instances -> IAS::Update analogue (world/ias_manager.cpp:116-151,187-205); object id = position of the
        // RenderObject in World's list
        if (!m_moved.empty()) {
            const std::vector<world::RenderObject*> ros = m_world->GetRenderobjects();
            for (auto& mv : m_moved)
                for (size_t i = 0; i < ros.size(); ++i)
                    if (ros[i] == mv.first && pb_update_instance(m_ctx, static_cast<uint32_t>(i), mv.second.e) != PB_OK) return Fail("pb_update_instance");
            m_moved.clear();
        }
        const util::Mat4 s2c = m_world->camera->GetSampleToCameraMatrix(), c2w = m_world->camera->GetToWorldMatrix();
        if (pb_set_camera(m_ctx, s2c.e, c2w.e) != PB_OK) return Fail("pb_set_camera");
        m_sample_cnt = 0;                       // pt_pass.cpp:43-44
        m_random_seed = 0;
        m_dirty = false;
    }
    pb_render_params p{};
    p.width = m_w; p.height = m_h;
    p.max_depth = static_cast<uint32_t>(m_max_depth);
    p.first_sample = m_random_seed;             // frame f of pixel p is seeded Init(4, p, f), main.cu:53
    p.sample_count = 1;                         // one frame per OnRun like optixLaunch(w, h, 1); larger values batch frames
    p.sample_stride = 1;
    p.flags = 0u;
    if (m_sample_cnt == 0 || !m_accumulated_flag) p.flags |= PB_RENDER_CLEAR;
    pb_render_targets targets{};
    targets.accum_sum_rgba = reinterpret_cast<float*>(m_accum->cuda_ptr);     // Buffer::cuda_ptr is a CUdeviceptr (buffer.h:36)
    targets.albedo_rgb = reinterpret_cast<float*>(m_albedo->cuda_ptr);
    targets.normal_rgb = reinterpret_cast<float*>(m_normal->cuda_ptr);
    targets.test = reinterpret_cast<float*>(m_test->cuda_ptr);
    if (pb_render_ex(m_ctx, &p, &targets, m_stream) != PB_OK) return Fail("pb_render_ex");
    if (pb_resolve(m_ctx, targets.accum_sum_rgba, reinterpret_cast<float*>(m_frame->cuda_ptr), m_w * m_h, 0, m_stream) != PB_OK) return Fail("pb_resolve");
    m_sample_cnt += m_accumulated_flag ? 1u : 0u;     // pt_pass.cpp:55
    ++m_random_seed;                                  // pt_pass.cpp:56
}

}  // namespace Pupil::pt
