// b200_pt_pass.h — the path-tracing pass behind the reference's plug-in interface, driving
// libpupil_b200.so through its C ABI.  Replaces Pupil::pt::PTPass
// (example/path_tracer/pt_pass.h:21-41, pt_pass.cpp:29-237): same constructor / OnRun / SetScene
// contract, same named output buffers ("final result", "pt accum buffer", "albedo", "normal", "test"), same
// dirty handling (camera or instance change => sample count and seed restart, pt_pass.cpp:40-49).
#pragma once
#include "../../include/pupil_b200.h"
#include "system.h"

#include <utility>

namespace Pupil::pt {

class B200PTPass : public Pass {
public:
    explicit B200PTPass(System* system, int device = 0, std::string_view name = "B200 Path Tracing") noexcept;
    // several GPUs of this box: scene and BVH replicated on each, the frames of one OnRun dealt round-robin to them
    // (frame f of pixel p is seeded Init(4, p, f) wherever it runs), one NCCL reduce into devices[0]'s "pt accum buffer"
    B200PTPass(System* system, const std::vector<int>& devices, std::string_view name = "B200 Path Tracing") noexcept;
    ~B200PTPass() noexcept override;
    void OnRun() noexcept override;
    void SetScene(world::World* world) noexcept;
    void BindingEventCallback() noexcept;

    bool ok() const noexcept { return m_error.empty(); }
    const std::string& error() const noexcept { return m_error; }
    uint32_t sample_cnt() const noexcept { return m_sample_cnt; }
    int max_depth = 1;            // config.max_depth, editable like the reference's ImGui slider (pt_pass.cpp:225-237)
    bool accumulated_flag = true; // config.accumulated_flag
    uint32_t frames_per_run = 1;  // >1 renders several frames per OnRun (the B200 path needs no launch per spp); with n GPUs
                                  // an OnRun keeps all of them busy once this is >= n
    size_t device_count() const noexcept { return m_ctxs.size(); }

private:
    void fail(const char* what) noexcept;
    System* m_system;
    void init(const std::vector<int>& devices) noexcept;
    pb_ctx* m_ctx = nullptr;              // rank 0 (= m_ctxs[0]): owns the output buffers' device
    cudaStream_t m_stream = nullptr;
    std::vector<pb_ctx*> m_ctxs;
    std::vector<cudaStream_t> m_streams;
    std::vector<int> m_devices;
    world::World* m_world = nullptr;
    Buffer *m_frame = nullptr, *m_accum = nullptr, *m_albedo = nullptr, *m_normal = nullptr, *m_test = nullptr;
    std::vector<std::pair<uint32_t, pupil::host::Mat4>> m_moved;   // instances whose transform changed since the last OnRun
    uint32_t m_w = 0, m_h = 0, m_sample_cnt = 0, m_random_seed = 0;
    bool m_dirty = true;
    std::string m_error;
};

}  // namespace Pupil::pt
