#pragma once
#include <memory>
#include <vector>
#include "camera.h"
#include "render_object.h"
#include "resource/scene.h"
namespace Pupil {
enum class EWorldEvent {
    CameraChange,                            // REF framework/world/world.h:16 | CameraChange,
    CameraMove, CameraFovChange, CameraViewChange, RenderInstanceTransform,
    RenderInstanceUpdate,                    // REF framework/world/world.h:21 | RenderInstanceUpdate,
    RenderInstanceRemove
};
namespace world {
class World {
public:
    std::unique_ptr<resource::Scene> scene = nullptr;    // REF framework/world/world.h:28 | std::unique_ptr<resource::Scene> scene = nullptr;
    std::unique_ptr<CameraHelper> camera = nullptr;      // REF framework/world/world.h:29 | std::unique_ptr<CameraHelper> camera = nullptr;
    std::vector<RenderObject*> GetRenderobjects() noexcept;   // REF framework/world/world.h:48 | std::vector<RenderObject *> GetRenderobjects() noexcept;
};
}  // namespace world
}  // namespace Pupil
