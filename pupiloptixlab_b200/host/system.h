// system.h — headless mirror of the reference's pass-plugin runtime: Pass, BufferManager, World,
// System.  Same names, argument meaning and threading contract as
//   Pupil::Pass              framework/system/pass.h:22-43, pass.cpp:6-13
//   Pupil::BufferManager     framework/system/buffer.h:34-59, buffer.cpp:39-45 (cudaMalloc branch; the
//                            DX12-shared branch :46-85 is graphics interop: unavailable on this box)
//   Pupil::world::World      framework/world/world.h:26-54 (scene + camera hand-off only)
//   Pupil::System            framework/system/system.h:22-46, system.cpp:33-165 (Init(false) headless
//                            path :55-59; Run's frame loop :93-101 as RunFrames(n))
// so that B200PTPass reads like the reference's PTPass.  No window, no GUI pass, no render thread:
// RunFrames() executes the per-frame pass list synchronously on the caller's thread.
#pragma once
#include <cuda_runtime.h>

#include <chrono>
#include <functional>
#include <memory>
#include <string>
#include <string_view>
#include <unordered_map>
#include <vector>

#include "scene_xml.h"

namespace Pupil {

enum class EPassTag : uint32_t { None = 0, Pre = 1 << 0, Post = 1 << 1, Asyn = 1 << 2 };
inline bool operator&(EPassTag a, EPassTag b) noexcept { return static_cast<uint32_t>(a) & static_cast<uint32_t>(b); }

class Pass {
protected:
    double m_last_exec_time = 0.;
    bool m_enable = true;
public:
    const std::string name;
    const EPassTag tag;
    Pass(std::string_view name, EPassTag tag = EPassTag::None) noexcept : name(name), tag(tag) {}
    virtual ~Pass() = default;
    virtual void Run() noexcept {                       // pass.cpp:6-13: timer around OnRun when enabled
        if (!m_enable) return;
        const auto t0 = std::chrono::steady_clock::now();
        OnRun();
        m_last_exec_time = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    }
    virtual void Inspector() noexcept {}
    virtual void OnRun() noexcept = 0;
    void Toggle() noexcept { m_enable ^= true; }
    void SetEnablility(bool enable) noexcept { m_enable = enable; }
    bool IsEnabled() const noexcept { return m_enable; }
    double LastExecTimeMs() const noexcept { return m_last_exec_time; }
};

enum class EBufferFlag : unsigned int { None = 0, SharedWithDX12 = 1, AllowDisplay = 1 << 1 };
struct BufferDesc {
    const char* name = nullptr;
    EBufferFlag flag = EBufferFlag::None;
    uint32_t width = 1, height = 1, stride_in_byte = 1;
};
struct Buffer {
    BufferDesc desc{};
    void* cuda_ptr = nullptr;                           // CUdeviceptr in the reference
    Buffer() noexcept = default;
    explicit Buffer(const BufferDesc& d) noexcept : desc(d) {}
    ~Buffer() noexcept { if (cuda_ptr) cudaFree(cuda_ptr); }
    size_t size_in_byte() const noexcept { return static_cast<size_t>(desc.width) * desc.height * desc.stride_in_byte; }
};

class BufferManager {
public:
    constexpr static std::string_view DEFAULT_FINAL_RESULT_BUFFER_NAME = "final result";
    Buffer* GetBuffer(std::string_view id) noexcept {
        auto it = m_buffers.find(std::string(id));
        return it == m_buffers.end() ? nullptr : it->second.get();
    }
    // same name => the old buffer is replaced (the reference overwrites the map entry, buffer.cpp:39-92)
    Buffer* AllocBuffer(const BufferDesc& desc) noexcept {
        auto buf = std::make_unique<Buffer>(desc);
        if (cudaMalloc(&buf->cuda_ptr, buf->size_in_byte()) != cudaSuccess) { buf->cuda_ptr = nullptr; return nullptr; }
        cudaMemset(buf->cuda_ptr, 0, buf->size_in_byte());
        const std::string id = desc.name ? desc.name : "";
        if (!m_buffers.count(id)) m_buffer_names.push_back(id);
        Buffer* raw = buf.get();
        m_buffers[id] = std::move(buf);
        return raw;
    }
    const std::vector<std::string>& GetBufferNameList() const noexcept { return m_buffer_names; }
    void Destroy() noexcept { m_buffers.clear(); m_buffer_names.clear(); }
private:
    std::unordered_map<std::string, std::unique_ptr<Buffer>> m_buffers;
    std::vector<std::string> m_buffer_names;
};

enum class ESystemEvent { Quit, Precompute, StartRendering, StopRendering, SceneLoad, FrameFinished };
enum class EWorldEvent { CameraChange, CameraMove, CameraFovChange, CameraViewChange, RenderInstanceTransform, RenderInstanceUpdate, RenderInstanceRemove };

namespace world {
class World;
// world::RenderObject (framework/world/render_object.h:11-31) reduced to what the event contract needs: identity, the
// current transform and UpdateTransform / ApplyTransform, which raise EWorldEvent::RenderInstanceTransform with `this`
// as the payload exactly as render_object.cpp:46-54 does.  `index` = position in scene->shape_instances = object id
// behind the C ABI (World keeps the same mapping in m_ro_in_scene_index, world.cpp:46-47).
struct RenderObject {
    std::string name;
    uint32_t index = 0;
    pupil::host::Transform transform;
    World* world = nullptr;
    void UpdateTransform(const pupil::host::Transform& new_transform) noexcept;
    void ApplyTransform(const pupil::host::Transform& new_transform) noexcept;
};

// what the pass pulls from World (pt_pass.cpp:107-150): the loaded scene, the camera helper, the render objects
class World {
public:
    std::unique_ptr<pupil::host::Scene> scene;
    pupil::host::Camera camera;
    std::function<void(EWorldEvent, void*)> dispatch;   // set by System: the reference's EventDispatcher is a global
    bool LoadScene(const std::string& path, std::string& err) noexcept {
        scene = std::make_unique<pupil::host::Scene>();
        if (!pupil::host::load_scene_xml_file(path, *scene, err)) { scene.reset(); return false; }
        camera = scene->camera();
        m_ros.clear();
        for (size_t i = 0; i < scene->shape_instances.size(); ++i) {   // world.cpp:115-127: one RenderObject per instance
            auto ro = std::make_unique<RenderObject>();
            ro->name = scene->shape_instances[i].name; ro->index = static_cast<uint32_t>(i);
            ro->transform = scene->shape_instances[i].transform; ro->world = this;
            m_ros.push_back(std::move(ro));
        }
        return true;
    }
    std::vector<RenderObject*> GetRenderobjects() noexcept {   // world.cpp:141-148
        std::vector<RenderObject*> v;
        for (auto& r : m_ros) v.push_back(r.get());
        return v;
    }
    // the RenderInstanceTransform handler of world.cpp:45-54: the scene's instance takes the new transform, then
    // RenderInstanceUpdate goes out (emitter tables are rebuilt by pb_update_instance on the pass side)
    void OnInstanceTransform(RenderObject* ro) noexcept {
        if (!scene || !ro || ro->index >= scene->shape_instances.size()) return;
        scene->shape_instances[ro->index].transform = ro->transform;
        if (dispatch) dispatch(EWorldEvent::RenderInstanceUpdate, ro);
    }
private:
    std::vector<std::unique_ptr<RenderObject>> m_ros;
};
inline void RenderObject::UpdateTransform(const pupil::host::Transform& new_transform) noexcept {
    transform = new_transform;
    if (world && world->dispatch) world->dispatch(EWorldEvent::RenderInstanceTransform, this);
}
inline void RenderObject::ApplyTransform(const pupil::host::Transform& new_transform) noexcept {
    transform.matrix = pupil::host::mul(new_transform.matrix, transform.matrix);
    if (world && world->dispatch) world->dispatch(EWorldEvent::RenderInstanceTransform, this);
}
}  // namespace world

class System {
public:
    bool render_flag = true;
    bool quit_flag = false;
    BufferManager buffers;
    world::World world;
    std::string last_error;

    System() {
        world.dispatch = [this](EWorldEvent e, void* p) { Dispatch(e, p); };
        Bind(EWorldEvent::RenderInstanceTransform, [this](void* p) { world.OnInstanceTransform(static_cast<world::RenderObject*>(p)); });
    }
    System(const System&) = delete;
    System& operator=(const System&) = delete;
    void Init(bool has_window = false) noexcept { (void)has_window; /* headless only: system.cpp:55-59 */ }
    void AddPass(Pass* pass) noexcept {                 // system.cpp:129-134: not owned
        if (pass->tag & EPassTag::Pre) m_pre_passes.push_back(pass);
        else m_passes.push_back(pass);
    }
    // events (util/event.h): passes bind callbacks; dispatch is synchronous
    void Bind(ESystemEvent e, std::function<void(void*)> fn) { m_sys_events[static_cast<int>(e)].push_back(std::move(fn)); }
    void Bind(EWorldEvent e, std::function<void(void*)> fn) { m_world_events[static_cast<int>(e)].push_back(std::move(fn)); }
    void Dispatch(ESystemEvent e, void* p = nullptr) { for (auto& f : m_sys_events[static_cast<int>(e)]) f(p); }
    void Dispatch(EWorldEvent e, void* p = nullptr) { for (auto& f : m_world_events[static_cast<int>(e)]) f(p); }

    // system.cpp:136-165: load, allocate the default frame buffer, tell the passes, run the pre passes
    bool SetScene(const std::string& path) noexcept {
        if (!world.LoadScene(path, last_error)) return false;
        BufferDesc d{BufferManager::DEFAULT_FINAL_RESULT_BUFFER_NAME.data(), EBufferFlag::AllowDisplay,
                     static_cast<uint32_t>(world.scene->film_w), static_cast<uint32_t>(world.scene->film_h), sizeof(float) * 4};
        if (!buffers.AllocBuffer(d)) { last_error = "cudaMalloc of the final result buffer failed"; return false; }
        Dispatch(ESystemEvent::SceneLoad, &world);
        for (Pass* p : m_pre_passes) p->Run();           // ESystemEvent::Precompute, system.cpp:50-53
        return last_error.empty();
    }
    // the body of the render thread's loop (system.cpp:96-101), n times
    void RunFrames(uint32_t n) noexcept {
        for (uint32_t i = 0; i < n && !quit_flag; ++i) {
            if (!render_flag) continue;
            for (Pass* p : m_passes) p->Run();
            Dispatch(ESystemEvent::FrameFinished);
        }
    }
    void Destroy() noexcept { m_passes.clear(); m_pre_passes.clear(); buffers.Destroy(); }
private:
    std::vector<Pass*> m_passes, m_pre_passes;
    std::unordered_map<int, std::vector<std::function<void(void*)>>> m_sys_events, m_world_events;
};

}  // namespace Pupil
