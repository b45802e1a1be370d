#pragma once
#include <string>
#include "util/transform.h"
namespace Pupil::world {
struct RenderObject {
    std::string name;                        // REF framework/world/render_object.h:13 | std::string name;
    util::Transform transform{};             // REF framework/world/render_object.h:17 | util::Transform transform{};
};
}  // namespace Pupil::world
