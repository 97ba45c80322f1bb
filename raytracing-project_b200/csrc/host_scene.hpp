// host_scene.hpp — host-side scene graph: what the JSON loader produces and the scene
// compiler consumes.  Mirrors the data the reference keeps in Scene / Camera / Primitive
// subclasses (raytracer/src/scene.h:32-47, camera.h:8-36, geometry.h, csg.h, transform.h),
// but as plain data: no virtual intersect(), because all ray queries run on the device.
#pragma once
#include <algorithm>
#include <cmath>
#include <memory>
#include <string>
#include <vector>

namespace rtb {

struct Vec3d { double x{0}, y{0}, z{0}; };
struct ColorD { double r{0}, g{0}, b{0}; };

// Material (geometry.h:5-22)
struct Material {
    ColorD albedo{1, 1, 1};
    ColorD ambient{0, 0, 0};
    double kd{1.0}, ks{0.0}, kr{0.0}, kt{0.0};
    double shininess{32.0};
    double refractive_index{1.0};
};

enum class NodeKind { Sphere, HalfSpace, Pokeball, Translation, Scaling, Rotation, Csg };
enum class CsgOp { Union, Intersection, Difference };     // csg.h:6

struct Node {
    NodeKind kind{NodeKind::Sphere};
    // Sphere / Pokeball (geometry.h:79-108, 161-242)
    Vec3d center; double radius{1.0};
    Material mat;                                   // sphere / half-space material
    // HalfSpace (geometry.h:111-134): normal is unit (falls back to +Y)
    Vec3d p0, normal{0, 1, 0};
    // Pokeball regions and controls
    Material top, bottom, belt, ring, button;
    double belt_half{0.06}, button_outer{0.28}, ring_width{0.06};
    Vec3d button_dir{1, 0, 0};                      // unit
    // Transforms (transform.h:83-185): factors = translation vector or scale factors
    Vec3d factors; int axis{0}; double angle_rad{0.0};
    // CSG (csg.h:14-49) children; transforms use `a` as their subject
    CsgOp op{CsgOp::Union};
    std::shared_ptr<Node> a, b;
};

// ScreenSpec / Camera (camera.h:8-36)
struct ScreenSpec {
    Vec3d P{0, 0, 0}; double Lx{1.0}, Ly{1.0}; int dpi{72};
    int nx() const { return std::max(1, int(std::round(Lx * dpi))); }
    int ny() const { return std::max(1, int(std::round(Ly * dpi))); }
};
struct Camera { Vec3d eye{0, 0, 1}; ScreenSpec screen; };

struct PointLight { Vec3d pos; ColorD intensity{10, 10, 10}; };   // scene.h:22-27
struct DirectionalLight { Vec3d dir{0, -1, 0}; ColorD radiance{1, 1, 1}; };   // scene.h:10-19 (dir: from the light toward the scene)

// Scene (scene.h:32-47).  Owns its nodes (the reference keeps them in file-static pools,
// json_loader.cpp:27-35, so only one scene can be alive per process; here each Scene is independent).
struct Scene {
    std::vector<std::shared_ptr<Node>> objects;
    std::vector<DirectionalLight> dir_lights;      // never filled by the JSON loader (json_loader.cpp:191-204)
    std::vector<PointLight> point_lights;
    ColorD background{0, 0, 0};
    ColorD ambient{0, 0, 0};
    double medium_index{1.0};
    int recursion_limit{5};
};

}  // namespace rtb

// Same entry points as the reference's loader (json_loader.h:27-43).
namespace json_min { class Value; }
namespace jsonio {
bool load_scene_from_json(const std::string& filename, rtb::Scene& scene, rtb::Camera& cam);
bool load_scene_from_json_text(const std::string& json_text, rtb::Scene& scene, rtb::Camera& cam);
rtb::Material make_material_from_color_block(const json_min::Value& color_block);
}  // namespace jsonio
