// json_loader.cpp — scene JSON -> rtb::Scene / rtb::Camera.
//
// Host surface kept identical to the reference loader (raytracer/src/json_loader.cpp): same
// keys, same defaults, same validation and the same error texts, so a scene file (and a
// failing scene file) behaves the same.  Reading is done with json_min instead of nlohmann.
#include <fstream>
#include <sstream>
#include <stdexcept>

#include "host_scene.hpp"
#include "json_min.hpp"

using json_min::Value;
using namespace rtb;

namespace {

struct Fail : std::runtime_error { using std::runtime_error::runtime_error; };

Vec3d vec3_of(const Value& v) {                           // json_loader.cpp:56-59
    if (!v.is_array() || v.size() != 3) throw Fail("Expected array[3]");
    return Vec3d{v[0].as_double(), v[1].as_double(), v[2].as_double()};
}
ColorD rgb_of(const Value& v) {                           // json_loader.cpp:66-69
    if (!v.is_array() || v.size() != 3) throw Fail("Expected color array[3]");
    return ColorD{v[0].as_double(), v[1].as_double(), v[2].as_double()};
}
double mean3(const ColorD& c) { return (c.r + c.g + c.b) / 3.0; }

// "color" block (json_loader.cpp:83-123): defaults are the black material; specular /
// reflected / refracted RGB triples collapse to their mean; kd/ks keys are ignored.
Material material_of(const Value& block) {
    if (!block.is_object()) throw Fail("color must be an object");
    Material m;
    m.albedo = ColorD{0, 0, 0};
    m.shininess = 1.0;
    if (block.contains("diffuse"))   m.albedo = rgb_of(block.at("diffuse"));
    if (block.contains("ambient"))   m.ambient = rgb_of(block.at("ambient"));
    if (block.contains("specular"))  m.ks = mean3(rgb_of(block.at("specular")));
    if (block.contains("reflected")) m.kr = mean3(rgb_of(block.at("reflected")));
    if (block.contains("refracted")) m.kt = mean3(rgb_of(block.at("refracted")));
    if (block.contains("shininess")) m.shininess = block.at("shininess").as_double();
    return m;
}

void require(const Value& j, std::initializer_list<const char*> keys, const char* message) {
    for (const char* k : keys) if (!j.contains(k)) throw Fail(message);
}

std::shared_ptr<Node> node_of(const Value& jnode);

std::shared_ptr<Node> fold_left(const Value& arr, CsgOp op, size_t at_least, const char* message) {
    // variadic union / intersection / difference: ((n0 op n1) op n2) ... (json_loader.cpp:375-401)
    if (!arr.is_array() || arr.size() < at_least) throw Fail(message);
    std::shared_ptr<Node> acc = node_of(arr[0]);
    for (size_t i = 1; i < arr.size(); ++i) {
        auto parent = std::make_shared<Node>();
        parent->kind = NodeKind::Csg; parent->op = op;
        parent->a = acc; parent->b = node_of(arr[i]);
        acc = parent;
    }
    return acc;
}

std::shared_ptr<Node> node_of(const Value& jnode) {        // json_loader.cpp:409-434
    if (!jnode.is_object() || jnode.size() != 1) throw Fail("Each object node must be a one-entry object");
    const std::string& kind = jnode.obj.front().first;
    const Value& j = jnode.obj.front().second;
    auto n = std::make_shared<Node>();

    if (kind == "sphere") {                                // json_loader.cpp:211-227
        require(j, {"position", "radius", "color"}, "sphere requires 'position', 'radius', 'color'");
        n->kind = NodeKind::Sphere;
        n->center = vec3_of(j.at("position"));
        n->radius = j.at("radius").as_double();
        n->mat = material_of(j.at("color"));
        if (j.contains("index")) n->mat.refractive_index = j.at("index").as_double();
        return n;
    }
    if (kind == "halfSpace") {                             // json_loader.cpp:234-248 + geometry.h:124-134
        require(j, {"position", "normal", "color"}, "halfSpace requires 'position', 'normal', 'color'");
        n->kind = NodeKind::HalfSpace;
        n->p0 = vec3_of(j.at("position"));
        Vec3d N = vec3_of(j.at("normal"));
        n->mat = material_of(j.at("color"));
        if (j.contains("index")) n->mat.refractive_index = j.at("index").as_double();
        const double len2 = N.x * N.x + N.y * N.y + N.z * N.z;
        if (len2 > 0.0) { const double inv = 1.0 / std::sqrt(len2); n->normal = Vec3d{N.x * inv, N.y * inv, N.z * inv}; }
        else n->normal = Vec3d{0, 1, 0};
        return n;
    }
    if (kind == "pokeball") {                              // json_loader.cpp:256-297
        require(j, {"position", "radius"}, "pokeball requires 'position' and 'radius'.");
        n->kind = NodeKind::Pokeball;
        n->center = vec3_of(j.at("position"));
        n->radius = j.at("radius").as_double();
        n->top.albedo = ColorD{0.88, 0.12, 0.20};    n->top.ks = 0.15;    n->top.shininess = 64;
        n->bottom.albedo = ColorD{0.95, 0.95, 0.98}; n->bottom.ks = 0.08; n->bottom.shininess = 32;
        n->belt.albedo = ColorD{0.12, 0.12, 0.15};
        n->ring.albedo = ColorD{0.35, 0.35, 0.40};
        n->button.albedo = ColorD{0.96, 0.96, 0.99}; n->button.ks = 0.25; n->button.shininess = 64;
        if (j.contains("colors")) {
            const Value& c = j.at("colors");
            if (c.contains("top"))    n->top = material_of(c.at("top"));
            if (c.contains("bottom")) n->bottom = material_of(c.at("bottom"));
            if (c.contains("belt"))   n->belt = material_of(c.at("belt"));
            if (c.contains("ring"))   n->ring = material_of(c.at("ring"));
            if (c.contains("button")) n->button = material_of(c.at("button"));
        }
        if (j.contains("belt_half"))    n->belt_half = j.at("belt_half").as_double();
        if (j.contains("button_outer")) n->button_outer = j.at("button_outer").as_double();
        if (j.contains("ring_width"))   n->ring_width = j.at("ring_width").as_double();
        Vec3d d{1, 0, 0};
        if (j.contains("button_dir")) d = vec3_of(j.at("button_dir"));
        const double len = std::sqrt(d.x * d.x + d.y * d.y + d.z * d.z);   // Dir3::normalized, core.h:95-101
        n->button_dir = len > 1e-6 ? Vec3d{d.x / len, d.y / len, d.z / len} : Vec3d{0, 1, 0};
        return n;
    }
    if (kind == "scaling" || kind == "translation") {      // json_loader.cpp:304-323
        if (!j.contains("factors") || !j.contains("subject"))
            throw Fail(kind + " requires 'factors' and 'subject'");
        n->kind = kind == "scaling" ? NodeKind::Scaling : NodeKind::Translation;
        n->factors = vec3_of(j.at("factors"));
        n->a = node_of(j.at("subject"));
        return n;
    }
    if (kind == "rotation") {                              // json_loader.cpp:330-346
        require(j, {"angle", "direction", "subject"}, "rotation requires 'angle', 'direction', and 'subject'");
        const double deg = j.at("angle").as_double();
        const int axis = j.at("direction").as_int();
        if (axis < 0 || axis > 2) throw Fail("rotation direction must be 0 (X), 1 (Y), or 2 (Z)");
        n->kind = NodeKind::Rotation; n->axis = axis; n->angle_rad = deg * M_PI / 180.0;
        n->a = node_of(j.at("subject"));
        return n;
    }
    if (kind == "csg") {                                   // json_loader.cpp:355-367
        require(j, {"operator", "left", "right"}, "csg requires 'operator', 'left', 'right'");
        const std::string& op = j.at("operator").as_string();
        if (op == "union") n->op = CsgOp::Union;
        else if (op == "intersection") n->op = CsgOp::Intersection;
        else if (op == "difference") n->op = CsgOp::Difference;
        else throw Fail("csg.operator must be union/intersection/difference");
        n->kind = NodeKind::Csg;
        n->a = node_of(j.at("left"));
        n->b = node_of(j.at("right"));
        return n;
    }
    if (kind == "union")        return fold_left(j, CsgOp::Union, 1, "CSG array must be a non-empty array");
    if (kind == "intersection") return fold_left(j, CsgOp::Intersection, 1, "CSG array must be a non-empty array");
    if (kind == "difference")   return fold_left(j, CsgOp::Difference, 2, "difference array must have at least 2 elements");
    throw Fail("unknown object kind: " + kind);
}

void read_screen(const Value& root, Camera& cam) {         // json_loader.cpp:133-164
    if (!root.contains("screen")) return;
    const Value& j = root.at("screen");
    const int dpi = j.contains("dpi") ? j.at("dpi").as_int() : 72;
    double Lx = 1.0, Ly = 1.0;
    if (j.contains("dimensions")) {
        const Value& d = j.at("dimensions");
        if (!d.is_array() || d.size() != 2) throw Fail("screen.dimensions must be [Lx, Ly]");
        Lx = d[0].as_double(); Ly = d[1].as_double();
    }
    if (!j.contains("position")) throw Fail("screen.position is required");
    const Vec3d P = vec3_of(j.at("position"));
    if (!j.contains("observer")) throw Fail("screen.observer is required");
    const Vec3d eye = vec3_of(j.at("observer"));
    cam.screen.P = P; cam.screen.Lx = Lx; cam.screen.Ly = Ly; cam.screen.dpi = dpi; cam.eye = eye;
}

}  // namespace

namespace jsonio {

Material make_material_from_color_block(const Value& color_block) { return material_of(color_block); }

bool load_scene_from_json_text(const std::string& json_text, Scene& scene, Camera& cam) {   // json_loader.cpp:458-490
    Value root;
    try {
        root = json_min::parse(json_text);
    } catch (const json_min::parse_error& e) {
        throw std::runtime_error(std::string("JSON parse error: ") + e.what());
    }
    try {
        read_screen(root, cam);
        if (root.contains("medium")) {                     // json_loader.cpp:171-184
            const Value& m = root.at("medium");
            if (m.contains("ambient"))   scene.ambient = rgb_of(m.at("ambient"));
            if (m.contains("index"))     scene.medium_index = m.at("index").as_double();
            if (m.contains("recursion")) scene.recursion_limit = m.at("recursion").as_int();
        }
        if (root.contains("sources")) {                    // json_loader.cpp:191-204
            const Value& arr = root.at("sources");
            if (!arr.is_array()) throw Fail("'sources' must be an array");
            for (const Value& s : arr.arr) {
                if (!s.contains("position") || !s.contains("intensity"))
                    throw Fail("each source needs 'position' and 'intensity'");
                scene.point_lights.push_back(PointLight{vec3_of(s.at("position")), rgb_of(s.at("intensity"))});
            }
        }
        if (root.contains("background")) scene.background = rgb_of(root.at("background"));
        scene.objects.clear();
        if (root.contains("objects")) {
            const Value& arr = root.at("objects");
            if (!arr.is_array()) throw Fail("'objects' must be an array");
            for (const Value& node : arr.arr) scene.objects.push_back(node_of(node));
        }
        return true;
    } catch (const std::exception& e) {
        throw std::runtime_error(std::string("JSON processing error: ") + e.what());
    }
}

bool load_scene_from_json(const std::string& filename, Scene& scene, Camera& cam) {          // json_loader.cpp:499-504
    std::ifstream in(filename);
    if (!in) throw std::runtime_error("Cannot open JSON file: " + filename);
    std::ostringstream text; text << in.rdbuf();
    return load_scene_from_json_text(text.str(), scene, cam);
}

}  // namespace jsonio
