#pragma once
#include "util/type.h"
namespace Pupil::world {
class CameraHelper {
public:
    util::Mat4 GetSampleToCameraMatrix() noexcept;   // REF framework/world/camera.h:41 | util::Mat4 GetSampleToCameraMatrix() noexcept { return m_camera.GetSampleToCameraMatrix(); }
    util::Mat4 GetToWorldMatrix() noexcept;          // REF framework/world/camera.h:43 | util::Mat4 GetToWorldMatrix() noexcept { return m_camera.GetToWorldMatrix(); }
};
}  // namespace Pupil::world
