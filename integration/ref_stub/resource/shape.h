#pragma once
#include <cstdint>
#include <string>
#include "emitter.h"
#include "material.h"
#include "util/transform.h"
namespace Pupil::resource {
enum class EShapeType : unsigned int { _unknown = 0, _obj, _sphere, _cube, _rectangle };   // REF framework/resource/shape.h:29 | obj, sphere, cube, rectangle
struct Mesh {
    uint32_t vertex_num;                     // REF framework/resource/shape.h:35 | uint32_t vertex_num;
    uint32_t face_num;                       // REF framework/resource/shape.h:36 | uint32_t face_num;
    float* positions;                        // REF framework/resource/shape.h:38 | float *positions;
    float* normals;                          // REF framework/resource/shape.h:39 | float *normals;
    float* texcoords;                        // REF framework/resource/shape.h:40 | float *texcoords;
    uint32_t* indices;                       // REF framework/resource/shape.h:41 | uint32_t *indices;
};
struct Shape {
    EShapeType type = EShapeType::_unknown;  // REF framework/resource/shape.h:53 | EShapeType type = EShapeType::_unknown;
    union {
        Mesh mesh;                           // REF framework/resource/shape.h:55 | Mesh mesh;
    };
};
struct ShapeInstance {
    std::string name = "";
    Shape* shape = nullptr;                  // REF framework/resource/shape.h:64 | Shape *shape = nullptr;
    bool flip_tex_coords = false;            // REF framework/resource/shape.h:66 | bool flip_tex_coords = false;
    bool flip_normals = false;               // REF framework/resource/shape.h:67 | bool flip_normals = false;
    Material mat{};                          // REF framework/resource/shape.h:69 | Material mat{};
    bool is_emitter = false;                 // REF framework/resource/shape.h:70 | bool is_emitter = false;
    Emitter emitter{};                       // REF framework/resource/shape.h:71 | Emitter emitter{};
    util::Transform transform{};             // REF framework/resource/shape.h:73 | util::Transform transform{};
};
}  // namespace Pupil::resource
