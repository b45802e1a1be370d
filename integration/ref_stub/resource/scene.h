#pragma once
#include <vector>
#include "emitter.h"
#include "shape.h"
namespace Pupil::resource {
struct Integrator {
    int max_depth = 1;                       // REF framework/resource/scene.h:18 | int max_depth = 1;
};
struct Film {
    int w = 768;                             // REF framework/resource/scene.h:23 | int w = 768;
    int h = 576;                             // REF framework/resource/scene.h:24 | int h = 576;
};
struct Sensor {
    Film film{};                             // REF framework/resource/scene.h:33 | Film film{};
};
class Scene {
public:
    Integrator integrator;                   // REF framework/resource/scene.h:44 | Integrator integrator;
    Sensor sensor;                           // REF framework/resource/scene.h:45 | Sensor sensor;
    std::vector<ShapeInstance> shape_instances;   // REF framework/resource/scene.h:46 | std::vector<ShapeInstance> shape_instances;
    std::vector<Emitter> emitters;           // REF framework/resource/scene.h:47 | std::vector<Emitter> emitters;
};
}  // namespace Pupil::resource
