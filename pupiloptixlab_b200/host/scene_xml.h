// scene_xml.h — mitsuba3-style scene XML -> host scene description.
//
// Mirrors Pupil::resource::Scene::LoadFromXML and the loaders behind it
// (framework/resource/scene.cpp:27-224, shape.cpp:20-218, material.cpp:27-190,
//  xml/visitor.h:23-209, xml/object.cpp:9-63, xml/util_loader.cpp:8-210): same tags, same
// defaults, same order of operations, same quirks:
//   * <default name value> defines $name substitutions for every LATER attribute; <ref id> re-attaches a
//     previously declared object; unknown tags (<sampler>, <rfilter>, ...) are skipped with their subtree
//   * integrator: only max_depth (default 1); film: hdrfilm width/height (768 x 576); sample counts are
//     never read
//   * transforms: `matrix` (16 or 9 values) OR `lookat` OR scale -> rotate -> translate in that fixed order
//     whatever the XML order; to_uv supports only scale
//   * a <lookat> sensor is flipped twice (once in the transform loader, once in the sensor case)
//   * `distribution` is ignored (always GGX); face_normals is parsed and unused
// The result is plain data using the C ABI's POD structs (include/pupil_b200.h); upload() feeds it to a
// sink with the pb_* signatures in the order World::LoadScene does (framework/world/world.cpp:101-139).
#pragma once
#include <memory>
#include <string>
#include <unordered_map>
#include <vector>

#include "../../include/pupil_b200.h"
#include "../../include/pupil_host.h"
#include "mathx.h"
#include "xml_lite.h"

namespace pupil::host {

// ---- generic object tree (xml::Object, framework/resource/xml/object.h:20-45) --------------------
struct XObject {
    std::string obj_name, var_name, id, type;
    std::vector<std::pair<std::string, std::string>> properties;
    std::vector<XObject*> sub_object;   // owned by the pool; <ref> shares pointers
    std::string GetProperty(const std::string& n) const {
        for (auto& p : properties) if (p.first == n) return p.second;
        return "";
    }
    XObject* GetUniqueSubObject(const std::string& n) const {
        for (auto* s : sub_object) if (s->obj_name == n) return s;
        return nullptr;
    }
    // a parameter is either a sub object named `n` (e.g. <texture name="reflectance">) or a property
    std::pair<XObject*, std::string> GetParameter(const std::string& n) const {
        for (auto* s : sub_object) if (s->var_name == n) return {s, ""};
        return {nullptr, GetProperty(n)};
    }
};

struct XTree {
    std::vector<std::unique_ptr<XObject>> pool;   // pool[0] is the scene root
    std::unordered_map<std::string, std::string> defaults;
    std::unordered_map<std::string, XObject*> refs;
    std::vector<std::string> skipped;             // tags skipped with their subtree (diagnostics)
    XObject* root() const { return pool.empty() ? nullptr : pool[0].get(); }
};
bool build_tree(const xml::Node& doc, XTree& out, std::string& err);

// ---- resource::Scene (framework/resource/scene.h:14-56) -------------------------------------------
enum class ShapeType { Obj, Sphere, Cube, Rectangle };

struct MeshData {                       // ShapeManager::MeshData (shape.h:113-121)
    std::vector<float> positions, normals, texcoords;   // normals / texcoords may be empty
    std::vector<uint32_t> indices;
};

struct ShapeInstance {                  // resource::ShapeInstance (shape.h:60-72)
    std::string name;
    ShapeType type = ShapeType::Cube;
    std::shared_ptr<MeshData> mesh;     // cube / rectangle / obj
    bool face_normals = false, flip_tex_coords = false, flip_normals = false;
    pb_material mat{};
    bool is_emitter = false;
    pb_texture radiance{};
    Transform transform;
};

struct Scene {
    int max_depth = 1;
    int film_w = 768, film_h = 576;
    float fov = 90.f, near_clip = 0.01f, far_clip = 10000.f;   // fov already converted to the vertical fov
    Transform sensor_transform;
    std::vector<ShapeInstance> shape_instances;
    bool has_env = false;
    pb_env env{};
    std::vector<std::unique_ptr<std::vector<float>>> bitmaps;   // texel storage referenced by pb_texture::bitmap
    std::vector<std::string> warnings;
    std::string root_path;

    Camera camera() const {   // World::LoadScene -> CameraHelper::Reset (world.cpp:105-113)
        Camera c;
        c.fov_y = fov; c.aspect = static_cast<float>(film_w) / film_h;
        c.near_clip = near_clip; c.far_clip = far_clip; c.to_world = sensor_transform;
        return c;
    }
};

bool load_scene_xml_text(const std::string& text, const std::string& root_dir, Scene& out, std::string& err);
bool load_scene_xml_file(const std::string& path, Scene& out, std::string& err);
// World::LoadScene: every shape instance (insertion order), then the environment emitter, camera, commit
int upload_scene(const Scene& s, const ph_sink& sink, void* ctx, std::string& err);

// built-in meshes (the reference's literal tables, framework/resource/shape.cpp:20-66)
std::shared_ptr<MeshData> builtin_rectangle();
std::shared_ptr<MeshData> builtin_cube();
bool load_obj_file(const std::string& path, MeshData& out, std::string& err);

}  // namespace pupil::host
