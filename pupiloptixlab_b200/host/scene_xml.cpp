// scene_xml.cpp — see scene_xml.h.  Part 1: XML DOM -> object tree (the reference's visitors).
#include "scene_xml.h"

#include <algorithm>
#include <cstdlib>
#include <fstream>
#include <sstream>

#include <map>

#include "image_io.h"
#include "ior_tables.h"

namespace pupil::host {

namespace {

// ---- $default substitution (GlobalManager::ReplaceDefaultValue, xml/object.cpp:9-24).  The reference
// iterates an unordered_map; longest name first makes "$ab" win over "$a" deterministically.
std::string substitute(const XTree& t, std::string v) {
    if (v.find('$') == std::string::npos) return v;
    std::vector<std::pair<std::string, std::string>> params(t.defaults.begin(), t.defaults.end());
    std::sort(params.begin(), params.end(), [](auto& a, auto& b) { return a.first.size() != b.first.size() ? a.first.size() > b.first.size() : a.first < b.first; });
    for (auto& [name, value] : params) {
        const std::string key = "$" + name;
        size_t pos = 0;
        while ((pos = v.find(key, pos)) != std::string::npos) { v.replace(pos, key.size(), value); pos += value.size(); }
    }
    return v;
}

struct Builder {
    XTree& t;
    XObject* current = nullptr;

    std::string attr(const xml::Node& n, const char* name) const { return substitute(t, n.attr_or(name)); }

    XObject* make(const std::string& obj_name, const std::string& type) {
        t.pool.emplace_back(new XObject());
        XObject* o = t.pool.back().get();
        o->obj_name = obj_name; o->type = type;
        return o;
    }
    void property(const xml::Node& n) {                 // PropertyVisitor, visitor.h:23-33
        std::string name = attr(n, "name");
        if (name.empty()) name = n.name;
        if (current) current->properties.emplace_back(name, attr(n, "value"));
    }
    void xyz_property(const xml::Node& n, const char* dx, const char* dy, const char* dz) {   // visitor.h:66-88
        std::string name = attr(n, "name");
        if (name.empty()) name = n.name;
        std::string value = attr(n, "value");
        if (value.empty()) {
            std::string x = attr(n, "x"), y = attr(n, "y"), z = attr(n, "z");
            if (x.empty()) x = dx;
            if (y.empty()) y = dy;
            if (z.empty()) z = dz;
            value = x + "," + y + "," + z;
        }
        if (current) current->properties.emplace_back(name, value);
    }
    void object(const xml::Node& n) {                   // ObjectVisitor, visitor.h:35-55
        XObject* o = make(n.name, attr(n, "type"));
        if (n.has("id")) { o->id = attr(n, "id"); t.refs[o->id] = o; }
        if (n.has("name")) o->var_name = attr(n, "name");
        if (current) current->sub_object.push_back(o);
        current = o;
    }

    // returns false when the subtree is skipped
    bool visit(const xml::Node& n) {
        const std::string& tag = n.name;
        if (tag == "scene") { current = make(n.name, n.attr_or("version")); return true; }
        if (tag == "default") { t.defaults[attr(n, "name")] = attr(n, "value"); return true; }
        if (tag == "ref") {
            auto it = t.refs.find(attr(n, "id"));
            if (n.has("id") && it != t.refs.end() && current) current->sub_object.push_back(it->second);
            return true;
        }
        if (tag == "lookat") {                          // visitor.h:126-142: a sub object, never "current"
            XObject* o = make(n.name, "");
            o->properties.emplace_back("origin", attr(n, "origin"));
            o->properties.emplace_back("target", attr(n, "target"));
            o->properties.emplace_back("up", attr(n, "up"));
            if (current) current->sub_object.push_back(o);
            return true;
        }
        if (tag == "rotate") {                          // visitor.h:146-168
            XObject* o = make(n.name, "");
            std::string axis;
            if (n.has("value")) axis = attr(n, "value");
            else if (n.has("x")) axis = "1, 0, 0";
            else if (n.has("y")) axis = "0, 1, 0";
            else if (n.has("z")) axis = "0, 0, 1";
            o->properties.emplace_back("axis", axis);
            o->properties.emplace_back("angle", attr(n, "angle"));
            if (current) current->sub_object.push_back(o);
            return true;
        }
        if (tag == "scale") { xyz_property(n, "1", "1", "1"); return true; }
        if (tag == "point" || tag == "translate") { xyz_property(n, "0", "0", "0"); return true; }
        if (tag == "integer" || tag == "string" || tag == "float" || tag == "rgb" || tag == "boolean" || tag == "matrix") {
            property(n);
            return true;
        }
        if (tag == "bsdf" || tag == "emitter" || tag == "film" || tag == "integrator" || tag == "sensor" || tag == "shape" ||
            tag == "texture" || tag == "transform") {
            object(n);
            return true;
        }
        t.skipped.push_back(tag);                       // Visitor<_unknown>: "XML Node [..] skip"
        return false;
    }
    void dfs(const xml::Node& n) {                      // DfsParse, xml/parser.cpp:22-38
        XObject* parent = current;
        if (!visit(n)) return;
        for (auto& c : n.children) dfs(*c);
        if (n.name != "scene") current = parent;
    }
};

}  // namespace

bool build_tree(const xml::Node& doc, XTree& out, std::string& err) {
    if (doc.name != "scene") { err = "document element is <" + doc.name + ">, expected <scene>"; return false; }
    Builder b{out};
    b.dfs(doc);
    return out.root() != nullptr;
}

// ---- value parsing (xml/util_loader.cpp:8-125, util/util.cpp:4-29) ---------------------------------
namespace {

std::vector<std::string> split(const std::string& s, const std::string& d) {
    std::vector<std::string> r;
    size_t a = 0, e;
    while ((e = s.find(d, a)) != std::string::npos) { r.push_back(s.substr(a, e - a)); a = e + d.size(); }
    r.push_back(s.substr(a));
    return r;
}
float to_f(const std::string& s) { return std::strtof(s.c_str(), nullptr); }   // std::stof: leading blanks skipped

bool load_int(const XObject* o, const char* name, int& v, int def) {
    const std::string s = o->GetProperty(name);
    if (s.empty()) { v = def; return false; }
    v = (int)std::strtol(s.c_str(), nullptr, 10);
    return true;
}
bool load_float(const XObject* o, const char* name, float& v, float def = 0.f) {
    const std::string s = o->GetProperty(name);
    if (s.empty()) { v = def; return false; }
    v = to_f(s);
    return true;
}
// LoadFloat3: 3 values or 1 replicated; Load3Float: exactly 3
bool parse_float3(const std::string& s, float v[3], const float def[3], bool allow_one) {
    if (s.empty()) { v[0] = def[0]; v[1] = def[1]; v[2] = def[2]; return false; }
    auto p = split(s, ",");
    if (p.size() == 3) { v[0] = to_f(p[0]); v[1] = to_f(p[1]); v[2] = to_f(p[2]); return true; }
    if (p.size() == 1 && allow_one) { v[0] = v[1] = v[2] = to_f(p[0]); return true; }
    return false;
}
bool load_bool(const XObject* o, const char* name, bool& v, bool def) {
    const std::string s = o->GetProperty(name);
    if (s == "true") { v = true; return true; }
    if (s == "false") { v = false; return true; }
    v = def;
    return false;
}

pb_texture color_texture(float r, float g, float b) {   // TextureManager::GetColorTexture
    pb_texture t{};
    t.type = PB_TEX_RGB;
    t.a[0] = r; t.a[1] = g; t.a[2] = b;
    t.uv_r0[0] = 1.f; t.uv_r1[1] = 1.f;                  // identity uv transform
    return t;
}
void set_uv_transform(pb_texture& t, const Transform& tr) {
    for (int k = 0; k < 4; ++k) { t.uv_r0[k] = tr.matrix.at(0, k); t.uv_r1[k] = tr.matrix.at(1, k); }
}

}  // namespace
}  // namespace pupil::host

// ---- Part 2: object tree -> Scene (framework/resource/scene.cpp:75-224 and friends) ----------------
namespace pupil::host {
namespace {

struct Loader {
    Scene& sc;
    std::string& err;
    bool failed = false;
    struct BitmapRef { const float* texels; uint32_t w, h; };
    std::map<std::string, BitmapRef> bitmap_cache;

    void warn(const std::string& m) { sc.warnings.push_back(m); }
    void fail(const std::string& m) { if (!failed) err = m; failed = true; }

    // LoadTransform3D, util_loader.cpp:128-192
    void transform3d(const XObject* o, Transform& tr) {
        const std::string m = o->GetProperty("matrix");
        if (!m.empty()) {
            std::vector<std::string> e;
            for (auto& s : split(m, " ")) if (!s.empty()) e.push_back(s);   // the reference splits on single blanks
            if (e.size() == 16) {
                for (int i = 0; i < 16; ++i) tr.matrix.e[i] = to_f(e[i]);
            } else if (e.size() == 9) {
                for (int i = 0, j = 0; j < 9; ++j) { tr.matrix.e[i] = to_f(e[j]); ++i; if ((j + 1) % 3 == 0) ++i; }
            } else {
                warn("transform matrix size is " + std::to_string(e.size()) + " (must be 9 or 16)");
                for (size_t i = 0; i < e.size() && i < 16; ++i) tr.matrix.e[i] = to_f(e[i]);
            }
            return;
        }
        if (const XObject* la = o->GetUniqueSubObject("lookat")) {
            const float d_o[3] = {1, 0, 0}, d_t[3] = {0, 0, 0}, d_u[3] = {0, 1, 0};
            float org[3], tgt[3], up[3];
            parse_float3(la->GetProperty("origin"), org, d_o, false);
            parse_float3(la->GetProperty("target"), tgt, d_t, false);
            parse_float3(la->GetProperty("up"), up, d_u, false);
            tr.LookAt({org[0], org[1], org[2]}, {tgt[0], tgt[1], tgt[2]}, {up[0], up[1], up[2]});
            tr.FlipXZ();
            if (!o->GetProperty("scale").empty() || o->GetUniqueSubObject("rotate") || !o->GetProperty("translate").empty())
                warn("transform scale/rotate/translate is ignored as look_at exists");
            return;
        }
        const float zero[3] = {0, 0, 0};
        float v[3];
        if (parse_float3(o->GetProperty("scale"), v, zero, true)) tr.Scale(v[0], v[1], v[2]);
        if (const XObject* r = o->GetUniqueSubObject("rotate")) {
            float axis[3], angle;
            if (parse_float3(r->GetProperty("axis"), axis, zero, false) && load_float(r, "angle", angle)) tr.Rotate(axis[0], axis[1], axis[2], angle);
        }
        if (parse_float3(o->GetProperty("translate"), v, zero, false)) tr.Translate(v[0], v[1], v[2]);
    }
    // LoadTransform, util_loader.cpp:194-210
    void transform(const XObject* o, Transform& tr) {
        if (!o) return;
        if (o->var_name == "to_world") transform3d(o, tr);
        else if (o->var_name == "to_uv") {
            const float zero[3] = {0, 0, 0};
            float v[3];
            if (parse_float3(o->GetProperty("scale"), v, zero, true)) tr.Scale(v[0], v[1], v[2]);
        } else warn("transform [" + o->var_name + "] UNKNOWN");
    }
    // TextureManager::GetTexture (resource/texture.cpp): one load per path, texels kept by the scene
    pb_texture bitmap_texture(const std::string& file) {
        pb_texture t = color_texture(0.f, 0.f, 0.f);
        const std::string path = sc.root_path.empty() || (!file.empty() && file[0] == '/') ? file : sc.root_path + "/" + file;
        auto it = bitmap_cache.find(path);
        if (it == bitmap_cache.end()) {
            Image img;
            std::string e;
            if (!load_image(path, img, e)) { fail(e); return t; }
            sc.bitmaps.emplace_back(new std::vector<float>(std::move(img.rgba)));
            it = bitmap_cache.emplace(path, BitmapRef{sc.bitmaps.back()->data(), (uint32_t)img.w, (uint32_t)img.h}).first;
        }
        t.type = PB_TEX_BITMAP;
        t.bitmap = it->second.texels; t.bitmap_w = it->second.w; t.bitmap_h = it->second.h;
        return t;
    }
    // Scene::LoadXmlObj case _texture, scene.cpp:141-180
    pb_texture texture(const XObject* o) {
        pb_texture t = color_texture(0.f, 0.f, 0.f);
        if (o->type == "bitmap") {                       // scene.cpp:144-165
            t = bitmap_texture(o->GetProperty("filename"));
            t.filter_mode = o->GetProperty("filter_type") == "bilinear" ? 1u : 0u;
            const std::string wrap = o->GetProperty("wrap_mode");
            t.address_mode = wrap == "mirror" ? 2u : (wrap == "clamp" ? 1u : 0u);   // default is repeat
        } else if (o->type == "checkerboard") {
            const float d0[3] = {0.4f, 0.4f, 0.4f}, d1[3] = {0.2f, 0.2f, 0.2f};
            t.type = PB_TEX_CHECKERBOARD;
            parse_float3(o->GetProperty("color0"), t.a, d0, true);
            parse_float3(o->GetProperty("color1"), t.b, d1, true);
        } else {
            warn("unknown texture type [" + o->type + "]");
        }
        Transform uv;
        transform(o->GetUniqueSubObject("transform"), uv);
        set_uv_transform(t, uv);
        return t;
    }
    // LoadTextureOrRGB, util_loader.cpp:92-109
    pb_texture texture_or_rgb(const XObject* o, const char* name, float dr, float dg, float db) {
        auto [tex, rgb] = o->GetParameter(name);
        if (tex) return texture(tex);
        const float def[3] = {dr, dg, db};
        float c[3];
        parse_float3(rgb, c, def, true);
        return color_texture(c[0], c[1], c[2]);
    }
    pb_texture texture_or_rgb(const XObject* o, const char* name, float d) { return texture_or_rgb(o, name, d, d, d); }

    // LoadMaterialFromXml, material.cpp:27-190
    pb_material material(const XObject* o) {
        pb_material m{};
        m.reflectance = m.specular_reflectance = m.specular_transmittance = m.alpha = m.eta = m.k = color_texture(0, 0, 0);
        if (!o) return m;
        const std::string& ty = o->type;
        auto iors = [&](float def_int) {
            m.int_ior = dielectric_ior(o->GetProperty("int_ior"), def_int);
            m.ext_ior = dielectric_ior(o->GetProperty("ext_ior"), 1.000277f);
        };
        auto conductor = [&]() {
            float eta[3] = {0, 0, 0}, k[3] = {1, 1, 1};
            if (!conductor_ior(o->GetProperty("material"), eta, k)) { eta[0] = eta[1] = eta[2] = 0.f; k[0] = k[1] = k[2] = 1.f; }
            m.eta = texture_or_rgb(o, "eta", eta[0], eta[1], eta[2]);
            m.k = texture_or_rgb(o, "k", k[0], k[1], k[2]);
            m.specular_reflectance = texture_or_rgb(o, "specular_reflectance", 1.f);
        };
        if (ty == "diffuse") {
            m.type = PB_MAT_DIFFUSE;
            m.reflectance = texture_or_rgb(o, "reflectance", 0.5f);
        } else if (ty == "dielectric" || ty == "roughdielectric") {
            m.type = ty == "dielectric" ? PB_MAT_DIELECTRIC : PB_MAT_ROUGH_DIELECTRIC;
            iors(1.5046f);
            if (ty == "roughdielectric") m.alpha = texture_or_rgb(o, "alpha", 0.1f);
            m.specular_reflectance = texture_or_rgb(o, "specular_reflectance", 1.f);
            m.specular_transmittance = texture_or_rgb(o, "specular_transmittance", 1.f);
        } else if (ty == "conductor") {
            m.type = PB_MAT_CONDUCTOR;
            conductor();
        } else if (ty == "roughconductor") {
            m.type = PB_MAT_ROUGH_CONDUCTOR;
            m.alpha = texture_or_rgb(o, "alpha", 0.1f);
            conductor();
        } else if (ty == "plastic" || ty == "roughplastic") {
            m.type = ty == "plastic" ? PB_MAT_PLASTIC : PB_MAT_ROUGH_PLASTIC;
            iors(1.49f);
            m.nonlinear = o->GetProperty("nonlinear") == "true" ? 1u : 0u;
            if (ty == "roughplastic") m.alpha = texture_or_rgb(o, "alpha", 0.1f);
            m.reflectance = texture_or_rgb(o, "diffuse_reflectance", 0.5f);
            m.specular_reflectance = texture_or_rgb(o, "specular_reflectance", 1.f);
        } else if (ty == "twosided") {
            m = material(o->GetUniqueSubObject("bsdf"));
            m.twosided = 1u;
        } else {
            warn("unknown bsdf [" + ty + "]");
        }
        return m;
    }

    // LoadShapeInstanceFromXml, shape.cpp:181-218 with the ShapeLoader<> specialisations :86-146
    bool shape(const XObject* o, ShapeInstance& ins) {
        ins.name = o->id;
        Transform own;   // sphere: Scale(radius) then Translate(center) (shape.cpp:118-121)
        if (o->type == "cube") { ins.type = ShapeType::Cube; ins.mesh = builtin_cube(); }
        else if (o->type == "rectangle") { ins.type = ShapeType::Rectangle; ins.mesh = builtin_rectangle(); }
        else if (o->type == "sphere") {
            ins.type = ShapeType::Sphere;
            const float zero[3] = {0, 0, 0};
            float c[3], radius;
            parse_float3(o->GetProperty("center"), c, zero, false);
            load_float(o, "radius", radius, 1.f);
            own.Scale(radius, radius, radius);
            own.Translate(c[0], c[1], c[2]);
        } else if (o->type == "obj") {
            ins.type = ShapeType::Obj;
            ins.mesh = std::make_shared<MeshData>();
            std::string e;
            const std::string file = o->GetProperty("filename");
            const std::string path = sc.root_path.empty() ? file : sc.root_path + "/" + file;
            if (!load_obj_file(path, *ins.mesh, e)) { warn("Mesh load from " + path + " failed: " + e); return false; }
            load_bool(o, "face_normals", ins.face_normals, false);
            load_bool(o, "flip_tex_coords", ins.flip_tex_coords, true);
        } else {
            warn("Unknown shape type [" + o->type + "]");
            return false;
        }
        load_bool(o, "flip_normals", ins.flip_normals, false);
        ins.mat = material(o->GetUniqueSubObject("bsdf"));
        Transform tr;
        transform(o->GetUniqueSubObject("transform"), tr);
        if (ins.type == ShapeType::Sphere) ins.transform.matrix = mul(tr.matrix, own.matrix);
        else ins.transform = tr;
        ins.is_emitter = false;
        if (const XObject* e = o->GetUniqueSubObject("emitter")) {
            if (e->type == "area") { ins.is_emitter = true; ins.radiance = texture_or_rgb(e, "radiance", 0.f); }
            else warn("shape emitter not support");
        }
        return true;
    }

    // Scene::LoadXmlObj case _sensor, scene.cpp:97-140
    void sensor(const XObject* o) {
        if (o->type != "perspective") { warn("sensor only support perspective"); return; }
        load_float(o, "fov", sc.fov, 90.f);
        load_float(o, "near_clip", sc.near_clip, 0.01f);
        load_float(o, "far_clip", sc.far_clip, 10000.f);
        if (const XObject* film = o->GetUniqueSubObject("film")) {
            if (film->type != "hdrfilm") warn("film only support hdrfilm");
            else { load_int(film, "width", sc.film_w, 768); load_int(film, "height", sc.film_h, 576); }
        }
        char axis = 'x';
        const std::string a = o->GetProperty("fov_axis");
        if (a == "y" || a == "Y") axis = 'y';
        else if (!a.empty() && a != "x" && a != "X") warn("sensor fov_axis must be x or y");
        if (axis == 'x') {
            const float aspect = static_cast<float>(sc.film_h) / static_cast<float>(sc.film_w);
            const float radian = sc.fov * 3.14159265358979323846f / 180.f * 0.5f;
            const float t = std::tan(radian) * aspect;
            sc.fov = 2.f * std::atan(t) * 180.f / 3.14159265358979323846f;
        }
        transform(o->GetUniqueSubObject("transform"), sc.sensor_transform);
        sc.sensor_transform.FlipXZ();
    }

    // top-level emitters, scene.cpp:44-49,189-224 (area emitters only exist attached to shapes;
    // "point" is parsed by the reference and never used, world/emitter.cpp:314-316)
    void emitter(const XObject* o) {
        if (o->type == "constant") {
            const float zero[3] = {0, 0, 0};
            sc.env = pb_env{};
            sc.env.type = PB_ENV_CONSTANT;
            parse_float3(o->GetProperty("radiance"), sc.env.radiance, zero, true);
            sc.has_env = true;
        } else if (o->type == "envmap") {                // scene.cpp:208-220
            sc.env = pb_env{};
            sc.env.type = PB_ENV_MAP;
            load_float(o, "scale", sc.env.scale, 1.f);
            sc.env.map = bitmap_texture(o->GetProperty("filename"));
            sc.env.map.filter_mode = 1u;                  // Linear
            sc.env.map.address_mode = 0u;                 // Wrap
            Transform tr;
            transform(o->GetUniqueSubObject("transform"), tr);
            for (int r = 0; r < 3; ++r)
                for (int k = 0; k < 4; ++k) sc.env.to_world[4 * r + k] = tr.matrix.at(r, k);
            sc.has_env = true;
        } else if (o->type == "point" || o->type == "area") {
            if (o->type == "point") warn("point emitters are parsed and ignored (as in the reference)");
        } else {
            warn("unknown emitter type [" + o->type + "]");
        }
    }

    void scene(const XObject* root) {                    // Scene::LoadFromXML, scene.cpp:27-54
        for (const XObject* o : root->sub_object) {
            if (o->obj_name == "integrator") load_int(o, "max_depth", sc.max_depth, 1);
            else if (o->obj_name == "sensor") sensor(o);
            else if (o->obj_name == "shape") { ShapeInstance ins; if (shape(o, ins)) sc.shape_instances.push_back(std::move(ins)); }
            else if (o->obj_name == "emitter") emitter(o);
        }
    }
};

}  // namespace

bool load_scene_xml_text(const std::string& text, const std::string& root_dir, Scene& out, std::string& err) {
    out = Scene{};
    out.root_path = root_dir;
    xml::Reader rd(text);
    auto doc = rd.parse();
    if (!doc) { err = "XML parse error: " + rd.error(); return false; }
    XTree tree;
    if (!build_tree(*doc, tree, err)) return false;
    for (auto& s : tree.skipped) out.warnings.push_back("XML Node [" + s + "] skip");
    Loader ld{out, err};
    ld.scene(tree.root());
    return !ld.failed;
}

bool load_scene_xml_file(const std::string& path, Scene& out, std::string& err) {
    std::ifstream f(path, std::ios::binary);
    if (!f) { err = "cannot open " + path; return false; }
    std::stringstream ss;
    ss << f.rdbuf();
    const size_t slash = path.find_last_of('/');
    return load_scene_xml_text(ss.str(), slash == std::string::npos ? "" : path.substr(0, slash), out, err);
}

}  // namespace pupil::host

// ---- Part 3: built-in meshes, OBJ reader, World::LoadScene hand-off --------------------------------
namespace pupil::host {

// rectangle: XY square [-1,1]^2, +Z normal, 2 triangles (shape.cpp:20-32)
std::shared_ptr<MeshData> builtin_rectangle() {
    static std::shared_ptr<MeshData> m = [] {
        auto r = std::make_shared<MeshData>();
        r->positions = {-1, -1, 0, 1, -1, 0, 1, 1, 0, -1, 1, 0};
        r->normals = {0, 0, 1, 0, 0, 1, 0, 0, 1, 0, 0, 1};
        r->texcoords = {0, 0, 1, 0, 1, 1, 0, 1};
        r->indices = {0, 1, 2, 0, 2, 3};
        return r;
    }();
    return m;
}
// cube: 6 faces x 4 vertices, per-face normals and [0,1]^2 uvs, 12 triangles (shape.cpp:35-66)
std::shared_ptr<MeshData> builtin_cube() {
    static std::shared_ptr<MeshData> m = [] {
        auto r = std::make_shared<MeshData>();
        struct Face { float n[3]; float c[4][3]; };
        static const Face faces[6] = {
            {{-1, 0, 0}, {{-1, -1, -1}, {-1, -1, 1}, {-1, 1, 1}, {-1, 1, -1}}},
            {{0, 0, -1}, {{1, -1, -1}, {-1, -1, -1}, {-1, 1, -1}, {1, 1, -1}}},
            {{1, 0, 0}, {{1, -1, 1}, {1, -1, -1}, {1, 1, -1}, {1, 1, 1}}},
            {{0, 0, 1}, {{-1, -1, 1}, {1, -1, 1}, {1, 1, 1}, {-1, 1, 1}}},
            {{0, 1, 0}, {{-1, 1, 1}, {1, 1, 1}, {1, 1, -1}, {-1, 1, -1}}},
            {{0, -1, 0}, {{-1, -1, -1}, {1, -1, -1}, {1, -1, 1}, {-1, -1, 1}}},
        };
        static const float uv[8] = {0, 0, 1, 0, 1, 1, 0, 1};
        for (uint32_t f = 0; f < 6; ++f) {
            for (int v = 0; v < 4; ++v) {
                r->positions.insert(r->positions.end(), faces[f].c[v], faces[f].c[v] + 3);
                r->normals.insert(r->normals.end(), faces[f].n, faces[f].n + 3);
                r->texcoords.push_back(uv[2 * v]); r->texcoords.push_back(uv[2 * v + 1]);
            }
            const uint32_t b = 4 * f;
            const uint32_t idx[6] = {b, b + 1, b + 2, b, b + 2, b + 3};
            r->indices.insert(r->indices.end(), idx, idx + 6);
        }
        return r;
    }();
    return m;
}

// Wavefront OBJ: v / vt / vn / f with fan triangulation, one mesh per file, vertices unified per
// (v, vt, vn) triple — what the reference gets from assimp's aiProcess_Triangulate import of a
// single-mesh file (shape.cpp:219-278).  Normals / texcoords are dropped unless every corner has one.
bool load_obj_file(const std::string& path, MeshData& out, std::string& err) {
    std::ifstream f(path);
    if (!f) { err = "cannot open file"; return false; }
    std::vector<float> v, vt, vn;
    struct Key { int a, b, c; bool operator==(const Key& o) const { return a == o.a && b == o.b && c == o.c; } };
    struct KeyHash { size_t operator()(const Key& k) const { return ((size_t)k.a * 73856093u) ^ ((size_t)k.b * 19349663u) ^ ((size_t)k.c * 83492791u); } };
    std::unordered_map<Key, uint32_t, KeyHash> remap;
    bool all_vt = true, all_vn = true;
    std::string line;
    while (std::getline(f, line)) {
        std::istringstream ls(line);
        std::string tag;
        ls >> tag;
        if (tag == "v") { float x, y, z; ls >> x >> y >> z; v.insert(v.end(), {x, y, z}); }
        else if (tag == "vt") { float a = 0, b = 0; ls >> a >> b; vt.insert(vt.end(), {a, b}); }
        else if (tag == "vn") { float x, y, z; ls >> x >> y >> z; vn.insert(vn.end(), {x, y, z}); }
        else if (tag == "f") {
            std::vector<uint32_t> corner;
            std::string tok;
            while (ls >> tok) {
                int i[3] = {0, 0, 0};
                int k = 0;
                size_t a = 0;
                while (k < 3) {
                    size_t e = tok.find('/', a);
                    const std::string part = tok.substr(a, e == std::string::npos ? std::string::npos : e - a);
                    if (!part.empty()) i[k] = std::atoi(part.c_str());
                    ++k;
                    if (e == std::string::npos) break;
                    a = e + 1;
                }
                auto fix = [](int idx, size_t n) { return idx < 0 ? (int)n + idx + 1 : idx; };   // negative = relative
                Key key{fix(i[0], v.size() / 3), fix(i[1], vt.size() / 2), fix(i[2], vn.size() / 3)};
                if (key.a <= 0 || (size_t)key.a > v.size() / 3) { err = "face index out of range"; return false; }
                if (key.b <= 0 || (size_t)key.b > vt.size() / 2) { key.b = 0; all_vt = false; }
                if (key.c <= 0 || (size_t)key.c > vn.size() / 3) { key.c = 0; all_vn = false; }
                auto it = remap.find(key);
                if (it == remap.end()) {
                    it = remap.emplace(key, (uint32_t)(out.positions.size() / 3)).first;
                    out.positions.insert(out.positions.end(), v.begin() + 3 * (key.a - 1), v.begin() + 3 * key.a);
                    if (key.b) out.texcoords.insert(out.texcoords.end(), vt.begin() + 2 * (key.b - 1), vt.begin() + 2 * key.b);
                    else out.texcoords.insert(out.texcoords.end(), {0.f, 0.f});
                    if (key.c) out.normals.insert(out.normals.end(), vn.begin() + 3 * (key.c - 1), vn.begin() + 3 * key.c);
                    else out.normals.insert(out.normals.end(), {0.f, 0.f, 0.f});
                }
                corner.push_back(it->second);
            }
            for (size_t k = 2; k < corner.size(); ++k) out.indices.insert(out.indices.end(), {corner[0], corner[k - 1], corner[k]});
        }
    }
    if (out.indices.empty()) { err = "no faces"; return false; }
    if (!all_vt) out.texcoords.clear();
    if (!all_vn) out.normals.clear();
    return true;
}

// World::LoadScene (world.cpp:101-139) + RenderObject ctor (render_object.cpp:10-39): one render object
// per shape instance in scene order, then top-level emitters, then the camera; commit builds the
// emitter tables and the BVH (EmitterHelper::ComputeProbability + IASManager::SetInstance).
int upload_scene(const Scene& s, const ph_sink& sink, void* ctx, std::string& err) {
    auto ck = [&](int rc, const char* what) { if (rc != PB_OK && err.empty()) err = std::string(what) + " failed (" + std::to_string(rc) + ")"; return rc; };
    int rc;
    if ((rc = ck(sink.scene_begin(ctx), "scene_begin"))) return rc;
    for (const ShapeInstance& ins : s.shape_instances) {
        uint32_t mat = 0;
        if ((rc = ck(sink.add_material(ctx, &ins.mat, &mat), ("add_material for shape '" + ins.name + "'").c_str()))) return rc;
        const uint32_t flags = (ins.flip_normals ? PB_SHAPE_FLIP_NORMALS : 0u) | (ins.flip_tex_coords ? PB_SHAPE_FLIP_TEX_COORDS : 0u);
        const pb_texture* rad = ins.is_emitter ? &ins.radiance : nullptr;
        if (ins.type == ShapeType::Sphere) {
            rc = sink.add_sphere(ctx, ins.transform.matrix.e, flags, mat, rad, nullptr);
        } else {
            const MeshData& m = *ins.mesh;
            rc = sink.add_trimesh(ctx, m.positions.data(), m.normals.empty() ? nullptr : m.normals.data(),
                                  m.texcoords.empty() ? nullptr : m.texcoords.data(), (uint32_t)(m.positions.size() / 3),
                                  m.indices.data(), (uint32_t)(m.indices.size() / 3), ins.transform.matrix.e, flags, mat, rad, nullptr);
        }
        if (ck(rc, ("add shape '" + ins.name + "'").c_str())) return rc;
    }
    if (s.has_env && (rc = ck(sink.set_env(ctx, &s.env), "set_env"))) return rc;
    const Camera cam = s.camera();
    const Mat4 s2c = cam.GetSampleToCameraMatrix(), c2w = cam.GetToWorldMatrix();
    if ((rc = ck(sink.set_camera(ctx, s2c.e, c2w.e), "set_camera"))) return rc;
    return ck(sink.scene_commit(ctx), "scene_commit");
}

}  // namespace pupil::host
