// integration/b200_pt_pass.h — THE REFERENCE-SIDE BINDING: the translation unit a PupilOptixLab maintainer adds to
// example/path_tracer (next to, or instead of, pt_pass.{h,cpp}) to run the path tracer on libpupil_b200.so.
//
// Written against the reference's own types, member for member:
//   Pupil::Pass                       framework/system/pass.h:22-43
//   Pupil::BufferManager / BufferDesc / Buffer::cuda_ptr      framework/system/buffer.h:25-59
//   Pupil::world::World  (scene, camera, GetRenderobjects)    framework/world/world.h:26-54
//   Pupil::world::CameraHelper::Get{SampleToCamera,ToWorld}Matrix   framework/world/camera.h:41-43
//   Pupil::world::RenderObject::transform                      framework/world/render_object.h:11-31
//   Pupil::resource::Scene  (integrator, sensor.film, shape_instances, emitters)   framework/resource/scene.h:16-56
//   Pupil::resource::ShapeInstance / Shape / Mesh              framework/resource/shape.h:35-74
//   Pupil::resource::Material and its per-type structs         framework/resource/material.h:16-84
//   Pupil::resource::Emitter (ConstEnv / EnvMap)               framework/resource/emitter.h:7-54
//   Pupil::util::Texture (rgb / bitmap / checkerboard)         framework/util/texture.h:9-66
//   EventBinder<...>                                           framework/util/event.h:34-38
// It cannot be compiled against the real tree here (MSVC / DX12 / OptiX), so tests/test_integration_shim.py
// compiles it (-fsyntax-only, and for real) against integration/ref_stub/, a header set that declares exactly the
// reference members used here, each with the reference file:line it restates, and — where /root/reference exists —
// checks every such citation against the reference's text.  The file this repository ships and runs on the GPU is the
// same code over its headless mirror of those interfaces: pupiloptixlab_b200/host/b200_pt_pass.{h,cpp}.
#pragma once
#include <cuda_runtime.h>

#include <string>
#include <string_view>
#include <utility>
#include <vector>

#include "pupil_b200.h"

#include "resource/scene.h"
#include "system/buffer.h"
#include "system/pass.h"
#include "system/system.h"
#include "util/event.h"
#include "util/util.h"
#include "world/world.h"

namespace Pupil::pt {

class B200PTPass : public Pass {
public:
    B200PTPass(std::string_view name = "B200 Path Tracing", int device = 0) noexcept;   // PTPass::PTPass, pt_pass.cpp:29-37
    ~B200PTPass() noexcept;
    void OnRun() noexcept override;                                                     // PTPass::OnRun, pt_pass.cpp:39-57
    void SetScene(world::World* world) noexcept;                                        // PTPass::SetScene, pt_pass.cpp:107-209

    const std::string& error() const noexcept { return m_error; }

private:
    void BindingEventCallback() noexcept;                                               // pt_pass.cpp:211-223
    void Fail(const char* what) noexcept;

    pb_ctx* m_ctx = nullptr;
    cudaStream_t m_stream = nullptr;
    world::World* m_world = nullptr;
    Buffer *m_frame = nullptr, *m_accum = nullptr, *m_albedo = nullptr, *m_normal = nullptr, *m_test = nullptr;
    uint32_t m_w = 0, m_h = 0;
    int m_max_depth = 1;
    bool m_accumulated_flag = true;
    uint32_t m_sample_cnt = 0, m_random_seed = 0;      // OptixLaunchParams::{sample_cnt, random_seed}, type.h:22-23
    bool m_dirty = true;
    std::vector<std::pair<world::RenderObject*, util::Mat4>> m_moved;   // instances whose transform changed since the last OnRun
    std::string m_error;
};

}  // namespace Pupil::pt
