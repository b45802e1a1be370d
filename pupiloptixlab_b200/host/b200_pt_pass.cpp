// b200_pt_pass.cpp — see b200_pt_pass.h.
#include "b200_pt_pass.h"

namespace Pupil::pt {

B200PTPass::B200PTPass(System* system, int device, std::string_view name) noexcept : Pass(name), m_system(system) {
    init({device});
}
B200PTPass::B200PTPass(System* system, const std::vector<int>& devices, std::string_view name) noexcept : Pass(name), m_system(system) {
    init(devices.empty() ? std::vector<int>{0} : devices);
}

// cuda::Context::Init + optix::Context::Init + PTPass ctor (pt_pass.cpp:29-37): one call per GPU
void B200PTPass::init(const std::vector<int>& devices) noexcept {
    m_devices = devices;
    m_ctxs.assign(devices.size(), nullptr);
    int rc;
    if (devices.size() == 1) rc = pb_create(&m_ctxs[0], devices[0]);
    else rc = pb_create_multi(m_ctxs.data(), devices.data(), static_cast<int>(devices.size()));
    if (rc != PB_OK) {
        m_error = rc == PB_ERR_NCCL ? "pb_create_multi: NCCL is not available or failed to initialise"
                                    : std::string("pb_create: ") + (m_ctxs[0] ? pb_last_error(m_ctxs[0]) : "no CUDA device (there is no CPU fallback)");
        for (auto& c : m_ctxs) { if (c) pb_destroy(c); c = nullptr; }
        m_ctxs.clear();
        return;
    }
    m_ctx = m_ctxs[0];
    for (size_t i = 0; i < devices.size(); ++i) {
        cudaSetDevice(devices[i]);
        cudaStream_t st = nullptr;
        if (cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking) != cudaSuccess) m_error = "cudaStreamCreate failed";
        m_streams.push_back(st);
    }
    cudaSetDevice(devices[0]);
    m_stream = m_streams.empty() ? nullptr : m_streams[0];
    BindingEventCallback();
}

B200PTPass::~B200PTPass() noexcept {
    for (size_t i = 0; i < m_streams.size(); ++i) { cudaSetDevice(m_devices[i]); if (m_streams[i]) cudaStreamDestroy(m_streams[i]); }
    for (pb_ctx* c : m_ctxs) if (c) pb_destroy(c);
    if (!m_devices.empty()) cudaSetDevice(m_devices[0]);
}

void B200PTPass::fail(const char* what) noexcept {
    if (m_error.empty()) m_error = std::string(what) + ": " + (m_ctx ? pb_last_error(m_ctx) : "");
    m_system->last_error = m_error;
}

// pt_pass.cpp:211-223
void B200PTPass::BindingEventCallback() noexcept {
    m_system->Bind(EWorldEvent::CameraChange, [this](void*) { m_dirty = true; });
    // pt_pass.cpp:216-218 only sets the dirty flag and OnRun re-fetches the IAS handle with the new instance
    // transforms (GetIASHandle(2, true), pt_pass.cpp:46 -> IASManager::UpdateInstance, ias_manager.cpp:187-205); here the
    // moved object's id and matrix are recorded and handed to pb_update_instance before the next pb_render
    m_system->Bind(EWorldEvent::RenderInstanceUpdate, [this](void* p) {
        if (auto* ro = static_cast<world::RenderObject*>(p)) {
            bool found = false;
            for (auto& m : m_moved) if (m.first == ro->index) { m.second = ro->transform.matrix; found = true; }
            if (!found) m_moved.emplace_back(ro->index, ro->transform.matrix);
        }
        m_dirty = true;
    });
    m_system->Bind(ESystemEvent::SceneLoad, [this](void* p) { SetScene(static_cast<world::World*>(p)); });
}

// pt_pass.cpp:107-209: launch config, output buffers, acceleration structure, "SBT" (= object / material /
// emitter tables inside pb_scene_commit)
void B200PTPass::SetScene(world::World* world) noexcept {
    if (!m_ctx || !world || !world->scene) return;
    m_world = world;
    const pupil::host::Scene& sc = *world->scene;
    std::string err;
    const ph_sink* sink = ph_default_sink();
    for (pb_ctx* c : m_ctxs)      // replicated: every GPU ingests the same host data and builds the identical tree
        if (pupil::host::upload_scene(sc, *sink, c, err) != PB_OK) { m_error = err + ": " + pb_last_error(c); m_system->last_error = m_error; return; }
    cudaSetDevice(m_devices[0]);  // the BufferManager's buffers live on the first device
    m_w = static_cast<uint32_t>(sc.film_w); m_h = static_cast<uint32_t>(sc.film_h);
    max_depth = sc.max_depth;
    BufferManager& bm = m_system->buffers;
    m_frame = bm.GetBuffer(BufferManager::DEFAULT_FINAL_RESULT_BUFFER_NAME);
    m_accum = bm.AllocBuffer({"pt accum buffer", EBufferFlag::None, m_w, m_h, sizeof(float) * 4});   // sums.xyz + frame count.w
    m_albedo = bm.AllocBuffer({"albedo", EBufferFlag::AllowDisplay, m_w, m_h, sizeof(float) * 3});
    m_normal = bm.AllocBuffer({"normal", EBufferFlag::AllowDisplay, m_w, m_h, sizeof(float) * 3});
    m_test = bm.AllocBuffer({"test", EBufferFlag::AllowDisplay, m_w, m_h, sizeof(float)});                // pt_pass.cpp:144-146
    m_moved.clear();
    if (!m_frame || !m_accum || !m_albedo || !m_normal || !m_test) { m_error = "output buffer allocation failed"; m_system->last_error = m_error; return; }
    m_dirty = true;
}

// pt_pass.cpp:39-57
void B200PTPass::OnRun() noexcept {
    if (!m_ctx || !m_world || !m_error.empty()) return;
    cudaSetDevice(m_devices[0]);
    if (m_dirty) {
        const pupil::host::Mat4 s2c = m_world->camera.GetSampleToCameraMatrix(), c2w = m_world->camera.GetToWorldMatrix();
        for (pb_ctx* c : m_ctxs) {
            for (auto& mv : m_moved)
                if (pb_update_instance(c, mv.first, mv.second.e) != PB_OK) { m_error = std::string("pb_update_instance: ") + pb_last_error(c); m_system->last_error = m_error; return; }
            if (pb_set_camera(c, s2c.e, c2w.e) != PB_OK) { m_error = std::string("pb_set_camera: ") + pb_last_error(c); m_system->last_error = m_error; return; }
        }
        m_moved.clear();
        m_sample_cnt = 0;
        m_random_seed = 0;
        m_dirty = false;
    }
    pb_render_params p{};
    p.width = m_w; p.height = m_h;
    p.max_depth = static_cast<uint32_t>(max_depth);
    p.first_sample = m_random_seed;                      // frame f of pixel p is seeded Init(4, p, f), main.cu:53
    p.sample_count = frames_per_run;
    p.sample_stride = 1;
    p.flags = (m_sample_cnt == 0 || !accumulated_flag) ? PB_RENDER_CLEAR : 0u;
    const pb_render_targets tg{static_cast<float*>(m_accum->cuda_ptr), static_cast<float*>(m_albedo->cuda_ptr),
                               static_cast<float*>(m_normal->cuda_ptr), static_cast<float*>(m_test->cuda_ptr)};
    if (m_ctxs.size() == 1) {
        if (pb_render_ex(m_ctx, &p, &tg, m_stream) != PB_OK) return fail("pb_render_ex");
    } else {   // the frames of this run dealt round-robin to the GPUs, one NCCL reduce into m_accum (rank 0)
        std::vector<void*> streams(m_streams.begin(), m_streams.end());
        if (pb_render_multi(m_ctxs.data(), static_cast<int>(m_ctxs.size()), &p, &tg, streams.data()) != PB_OK) return fail("pb_render_multi");
    }
    if (pb_resolve(m_ctx, static_cast<const float*>(m_accum->cuda_ptr), static_cast<float*>(m_frame->cuda_ptr), m_w * m_h, 0, m_stream) != PB_OK)
        return fail("pb_resolve");
    if (accumulated_flag) m_sample_cnt += frames_per_run;   // sample_cnt += accumulated_flag (pt_pass.cpp:55)
    m_random_seed += frames_per_run;                         // ++random_seed (pt_pass.cpp:56)
}

}  // namespace Pupil::pt
