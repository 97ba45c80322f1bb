// scene_compiler.cpp — host scene graph -> flat, device-ready rt_scene_desc.
//
// The reference walks a pointer graph of virtual Primitive objects for every ray
// (scene.cpp:15-21).  Here the graph is flattened once on the host:
//   * top-level objects stay in JSON order (closest-hit tie rule and Scene::occluded order,
//     scene.cpp:10-42) as 48-byte records: sphere / half-space / pokeball / CSG program;
//   * Translation / Rotation / uniform Scaling nodes (transform.cpp) are BAKED into the leaves
//     they enclose (centres, radii, plane equations, pokeball axes) in double precision, so the
//     device never touches a matrix for them; only a sphere/pokeball under a non-uniform
//     scaling keeps a world->local affine map (RT_FLAG_AFFINE);
//   * every CSG subtree becomes a postfix program over an accumulator + small stack
//     (csg.cpp:61-163 evaluated bottom-up), left spines need no stack at all;
//   * what the transforms do besides moving geometry — re-deriving front_face, collapsing
//     non-finite interval ends to t=0, the [0,inf] child query followed by an exclusive
//     (tmin,tmax) test — is kept as flags (RT_FLAG_XFORMED / RT_FLAG_WRAPPED) and RT_OP_XFIX ops.
#include <cmath>
#include <cstring>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/rt_b200.h"
#include "host_scene.hpp"

using namespace rtb;

namespace {

// world -> local affine map x_l = A x + b, tracked in double while descending the graph
struct Affine {
    double A[3][3]{{1, 0, 0}, {0, 1, 0}, {0, 0, 1}};
    double b[3]{0, 0, 0};
    bool identity{true};
};

Affine then_apply(const Affine& first, const double B[3][3], const double c[3]) {
    // result(x) = B * first(x) + c
    Affine r; r.identity = false;
    for (int i = 0; i < 3; ++i) {
        for (int j = 0; j < 3; ++j) {
            double s = 0; for (int k = 0; k < 3; ++k) s += B[i][k] * first.A[k][j];
            r.A[i][j] = s;
        }
        double s = c[i]; for (int k = 0; k < 3; ++k) s += B[i][k] * first.b[k];
        r.b[i] = s;
    }
    return r;
}

bool invert3(const double M[3][3], double out[3][3]) {
    const double det = M[0][0] * (M[1][1] * M[2][2] - M[1][2] * M[2][1])
                     - M[0][1] * (M[1][0] * M[2][2] - M[1][2] * M[2][0])
                     + M[0][2] * (M[1][0] * M[2][1] - M[1][1] * M[2][0]);
    if (std::abs(det) < 1e-300) return false;
    const double id = 1.0 / det;
    out[0][0] = (M[1][1] * M[2][2] - M[1][2] * M[2][1]) * id;
    out[0][1] = (M[0][2] * M[2][1] - M[0][1] * M[2][2]) * id;
    out[0][2] = (M[0][1] * M[1][2] - M[0][2] * M[1][1]) * id;
    out[1][0] = (M[1][2] * M[2][0] - M[1][0] * M[2][2]) * id;
    out[1][1] = (M[0][0] * M[2][2] - M[0][2] * M[2][0]) * id;
    out[1][2] = (M[0][2] * M[1][0] - M[0][0] * M[1][2]) * id;
    out[2][0] = (M[1][0] * M[2][1] - M[1][1] * M[2][0]) * id;
    out[2][1] = (M[0][1] * M[2][0] - M[0][0] * M[2][1]) * id;
    out[2][2] = (M[0][0] * M[1][1] - M[0][1] * M[1][0]) * id;
    return true;
}

bool is_transform(const Node& n) {
    return n.kind == NodeKind::Translation || n.kind == NodeKind::Scaling || n.kind == NodeKind::Rotation;
}

struct Ctx { bool xformed{false}; bool never{false}; };

class Compiler {
public:
    std::vector<rt_prim> objects, leaves;
    std::vector<rt_pokeball> pokeballs;
    std::vector<rt_affine> affines;
    std::vector<uint32_t> ops;
    std::vector<rt_material> materials;
    std::vector<rt_light> lights;
    std::vector<float> fallbacks;
    const Scene* scene{nullptr};

    // Descend through one transform node: world->local of the child = inverse(node) o current.
    Affine enter(const Node& t, const Affine& cur, Ctx& ctx) const {
        double B[3][3]{{1, 0, 0}, {0, 1, 0}, {0, 0, 1}}, c[3]{0, 0, 0};
        if (t.kind == NodeKind::Translation) {                 // transform.cpp:24-27: o - t
            c[0] = -t.factors.x; c[1] = -t.factors.y; c[2] = -t.factors.z;
        } else if (t.kind == NodeKind::Scaling) {              // transform.cpp:98-108: o / s, d / s
            const double s[3] = {t.factors.x, t.factors.y, t.factors.z};
            for (int i = 0; i < 3; ++i) {
                if (std::abs(s[i]) < 1e-6) { ctx.never = true; B[i][i] = 1.0; }   // transform.cpp:103
                else B[i][i] = 1.0 / s[i];
            }
        } else {                                               // Rotation: inverse rotation (core.h:200-234,241-263)
            const double cs = std::cos(t.angle_rad), sn = std::sin(t.angle_rad);
            // forward R; inverse = transpose
            double R[3][3]{{1, 0, 0}, {0, 1, 0}, {0, 0, 1}};
            if (t.axis == 0)      { R[1][1] = cs; R[1][2] = -sn; R[2][1] = sn;  R[2][2] = cs; }
            else if (t.axis == 1) { R[0][0] = cs; R[0][2] = sn;  R[2][0] = -sn; R[2][2] = cs; }
            else                  { R[0][0] = cs; R[0][1] = -sn; R[1][0] = sn;  R[1][1] = cs; }
            for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) B[i][j] = R[j][i];
        }
        ctx.xformed = true;
        return then_apply(cur, B, c);
    }

    int add_material(const Material& m) {
        rt_material r{};
        r.albedo[0] = (float)m.albedo.r; r.albedo[1] = (float)m.albedo.g; r.albedo[2] = (float)m.albedo.b;
        r.ambient[0] = (float)(m.ambient.r * scene->ambient.r);            // shading.cpp:39
        r.ambient[1] = (float)(m.ambient.g * scene->ambient.g);
        r.ambient[2] = (float)(m.ambient.b * scene->ambient.b);
        r.ks = (float)m.ks; r.kr = (float)m.kr; r.kt = (float)m.kt; r.kd = (float)m.kd;
        r.shininess = (float)m.shininess; r.ior = (float)m.refractive_index;
        materials.push_back(r);
        return (int)materials.size() - 1;
    }

    // A^T A == sigma^2 I ?  (rotations, translations, uniform scalings and their products)
    static bool similarity(const Affine& M, double& sigma) {
        double g[3][3];
        for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) {
            g[i][j] = 0; for (int k = 0; k < 3; ++k) g[i][j] += M.A[k][i] * M.A[k][j];
        }
        const double s2 = (g[0][0] + g[1][1] + g[2][2]) / 3.0;
        if (!(s2 > 0)) return false;
        for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) {
            const double want = i == j ? s2 : 0.0;
            if (std::abs(g[i][j] - want) > 1e-12 * s2) return false;
        }
        sigma = std::sqrt(s2);
        return true;
    }

    rt_prim bake_leaf(const Node& n, const Affine& M, const Ctx& ctx) {
        rt_prim p{}; p.xf = -1; p.aux = -1; p.flags = 0;
        if (ctx.xformed) p.flags |= RT_FLAG_XFORMED;
        if (ctx.never) p.flags |= RT_FLAG_NEVER;
        if (n.kind == NodeKind::HalfSpace) {
            // plane n_l.(A x + b - p0) >= 0  ->  (A^T n_l).x >= n_l.(p0 - b)
            p.kind = RT_PRIM_HALFSPACE;
            double g[3];
            const double nl[3] = {n.normal.x, n.normal.y, n.normal.z};
            const double p0[3] = {n.p0.x, n.p0.y, n.p0.z};
            for (int j = 0; j < 3; ++j) g[j] = M.A[0][j] * nl[0] + M.A[1][j] * nl[1] + M.A[2][j] * nl[2];
            double L = std::sqrt(g[0] * g[0] + g[1] * g[1] + g[2] * g[2]);
            if (!(L > 0)) { L = 1; g[0] = 0; g[1] = 1; g[2] = 0; }
            const double off = nl[0] * (p0[0] - M.b[0]) + nl[1] * (p0[1] - M.b[1]) + nl[2] * (p0[2] - M.b[2]);
            p.geom[0] = (float)(g[0] / L); p.geom[1] = (float)(g[1] / L); p.geom[2] = (float)(g[2] / L);
            p.geom[3] = (float)(off / L);
            p.mat = add_material(n.mat);
            return p;
        }
        // sphere / pokeball
        p.kind = n.kind == NodeKind::Pokeball ? RT_PRIM_POKEBALL : RT_PRIM_SPHERE;
        double sigma = 1.0;
        const bool sim = M.identity || similarity(M, sigma);
        double Q[3][3]{{1, 0, 0}, {0, 1, 0}, {0, 0, 1}};      // local = sigma * Q * world (+b)
        if (M.identity || ctx.never) {
            p.geom[0] = (float)n.center.x; p.geom[1] = (float)n.center.y; p.geom[2] = (float)n.center.z;
            p.geom[3] = (float)n.radius;
        } else if (sim) {
            double Ai[3][3];
            invert3(M.A, Ai);
            const double cl[3] = {n.center.x - M.b[0], n.center.y - M.b[1], n.center.z - M.b[2]};
            for (int i = 0; i < 3; ++i) p.geom[i] = (float)(Ai[i][0] * cl[0] + Ai[i][1] * cl[1] + Ai[i][2] * cl[2]);
            p.geom[3] = (float)(n.radius / sigma);
            for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) Q[i][j] = M.A[i][j] / sigma;
        } else {
            p.flags |= RT_FLAG_AFFINE;
            rt_affine a{};
            for (int i = 0; i < 3; ++i) { for (int j = 0; j < 3; ++j) a.m[i][j] = (float)M.A[i][j]; a.m[i][3] = (float)M.b[i]; }
            affines.push_back(a);
            p.xf = (int)affines.size() - 1;
            p.geom[0] = (float)n.center.x; p.geom[1] = (float)n.center.y; p.geom[2] = (float)n.center.z;
            p.geom[3] = (float)n.radius;
        }
        if (n.kind == NodeKind::Sphere) { p.mat = add_material(n.mat); return p; }
        // pokeball: materials in region order top,bottom,belt,ring,button; axes carried to world by Q^T
        p.mat = add_material(n.top); add_material(n.bottom); add_material(n.belt); add_material(n.ring); add_material(n.button);
        rt_pokeball pb{};
        const double bd[3] = {n.button_dir.x, n.button_dir.y, n.button_dir.z};
        for (int j = 0; j < 3; ++j) {
            pb.btn_dir[j] = (float)(Q[0][j] * bd[0] + Q[1][j] * bd[1] + Q[2][j] * bd[2]);
            pb.up[j] = (float)Q[1][j];
        }
        // geometry.cpp:169-174 without acos: ang in [0,pi]
        const double outer = n.button_outer, inner = std::max(0.0, n.button_outer - n.ring_width);
        pb.cos_outer = outer < 0.0 ? 2.0f : (outer >= M_PI ? -2.0f : (float)std::cos(outer));
        pb.cos_inner = inner > M_PI ? -2.0f : (float)std::cos(inner);
        pb.belt_half = (float)n.belt_half;
        pokeballs.push_back(pb);
        p.aux = (int)pokeballs.size() - 1;
        return p;
    }

    static const Node* strip(const Node* n, std::vector<const Node*>* chain = nullptr) {
        while (is_transform(*n)) { if (chain) chain->push_back(n); n = n->a.get(); }
        return n;
    }

    int add_leaf(const rt_prim& p) {
        leaves.push_back(p);
        if (leaves.size() >= (1u << 20)) throw std::runtime_error("too many CSG leaves");
        return (int)leaves.size() - 1;
    }

    // Emit the interval program of `n` (in the frame described by M); result lands in the accumulator.
    void emit(const Node& n, const Affine& M, Ctx ctx, int depth, int& max_depth) {
        if (is_transform(n)) {
            const Affine outside = M;
            Ctx inner = ctx;
            Affine Mi = enter(n, M, inner);
            emit(*n.a, Mi, inner, depth, max_depth);
            if (strip(&n)->kind == NodeKind::Csg) {
                // transform over a CSG result: collapse non-finite ends, orient; Scaling/Rotation also
                // replace a zero normal by normalized() 's fallback (0,1,0) of the frame outside this node.
                uint32_t arg = 0xFFFFF;
                if (n.kind != NodeKind::Translation) {
                    double g[3] = {outside.A[1][0], outside.A[1][1], outside.A[1][2]};   // A^T e_y
                    double L = std::sqrt(g[0] * g[0] + g[1] * g[1] + g[2] * g[2]);
                    if (!(L > 0)) { g[0] = 0; g[1] = 1; g[2] = 0; L = 1; }
                    fallbacks.push_back((float)(g[0] / L)); fallbacks.push_back((float)(g[1] / L));
                    fallbacks.push_back((float)(g[2] / L)); fallbacks.push_back(0.f);
                    arg = (uint32_t)(fallbacks.size() / 4 - 1);
                    if (arg > 63) throw std::runtime_error("more than 64 rotated/scaled CSG nodes nested inside CSG nodes");
                }
                ops.push_back(RT_OP_WORD(RT_OP_XFIX, 0, arg));
            }
            return;
        }
        if (n.kind != NodeKind::Csg) {
            ops.push_back(RT_OP_WORD(RT_OP_LEAF, 0, add_leaf(bake_leaf(n, M, ctx))));
            return;
        }
        const int op = n.op == CsgOp::Union ? RT_CSG_UNION : n.op == CsgOp::Intersection ? RT_CSG_INTERSECTION : RT_CSG_DIFFERENCE;
        emit(*n.a, M, ctx, depth, max_depth);
        // right operand: a (possibly transformed) leaf combines straight into the accumulator
        Affine Mb = M; Ctx cb = ctx;
        const Node* rb = n.b.get();
        while (is_transform(*rb)) { Mb = enter(*rb, Mb, cb); rb = rb->a.get(); }
        if (rb->kind != NodeKind::Csg) {
            ops.push_back(RT_OP_WORD(RT_OP_COMBINE_LEAF, op, add_leaf(bake_leaf(*rb, Mb, cb))));
        } else {
            if (depth + 1 > RT_MAX_CSG_STACK)
                throw std::runtime_error("CSG right-nesting deeper than " + std::to_string(RT_MAX_CSG_STACK));
            max_depth = std::max(max_depth, depth + 1);
            ops.push_back(RT_OP_WORD(RT_OP_PUSH, 0, 0));
            emit(*n.b, M, ctx, depth + 1, max_depth);
            ops.push_back(RT_OP_WORD(RT_OP_COMBINE_POP, op, 0));
        }
    }

    // Peephole over one object's program: consecutive COMBINE_LEAF ops with the same csg op over consecutive
    // plain-sphere leaves become one RT_OP_COMBINE_SPHERES (the device runs them as a tight loop).
    void fuse_sphere_runs(size_t begin) {
        std::vector<uint32_t> out(ops.begin(), ops.begin() + (long)begin);
        auto plain_sphere = [&](uint32_t leaf) {
            const rt_prim& l = leaves[leaf];
            return l.kind == RT_PRIM_SPHERE && (l.flags & (RT_FLAG_AFFINE | RT_FLAG_NEVER)) == 0;
        };
        for (size_t i = begin; i < ops.size();) {
            const uint32_t w = ops[i];
            const uint32_t opcode = w & 15u, csg = (w >> 4) & 3u, leaf = w >> 8;
            if (opcode == RT_OP_COMBINE_LEAF && plain_sphere(leaf)) {
                size_t j = i + 1;
                while (j < ops.size() && (ops[j] & 15u) == RT_OP_COMBINE_LEAF && ((ops[j] >> 4) & 3u) == csg &&
                       (ops[j] >> 8) == leaf + (uint32_t)(j - i) && plain_sphere(ops[j] >> 8) && (j - i) < 1023) ++j;
                if (j - i >= 2 && leaf < 16384u) {
                    out.push_back(RT_OP_WORD(RT_OP_COMBINE_SPHERES, csg, leaf | ((uint32_t)(j - i) << 14)));
                    i = j; continue;
                }
            }
            out.push_back(w); ++i;
        }
        ops.swap(out);
    }

    // World-space bounding sphere of a sphere-like leaf.  Baked leaves are world spheres already; a leaf under a
    // non-uniform scaling is the ellipsoid { x : |A x + b - c| <= r }, bounded by centre A^-1 (c - b) and radius
    // r * ||A^-1||_F (>= r * sigma_max(A^-1)).
    bool leaf_world_sphere(const rt_prim& l, double c[3], double& R) const {
        if (l.kind == RT_PRIM_HALFSPACE) return false;
        if (!(l.flags & RT_FLAG_AFFINE)) {
            c[0] = l.geom[0]; c[1] = l.geom[1]; c[2] = l.geom[2]; R = std::abs((double)l.geom[3]);
            return true;
        }
        const rt_affine& a = affines[(size_t)l.xf];
        double A[3][3], Ai[3][3];
        for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) A[i][j] = a.m[i][j];
        if (!invert3(A, Ai)) return false;
        const double v[3] = {l.geom[0] - a.m[0][3], l.geom[1] - a.m[1][3], l.geom[2] - a.m[2][3]};
        double f2 = 0;
        for (int i = 0; i < 3; ++i) {
            c[i] = Ai[i][0] * v[0] + Ai[i][1] * v[1] + Ai[i][2] * v[2];
            for (int j = 0; j < 3; ++j) f2 += Ai[i][j] * Ai[i][j];
        }
        R = std::abs((double)l.geom[3]) * std::sqrt(f2);
        return std::isfinite(R);
    }

    // Conservative bounding sphere of a program's leaves, inflated well beyond the FP32 rounding of the device's
    // discriminant, so that "misses the bound" implies "misses every leaf" on the device.  None when a leaf is a
    // half-space (unbounded).
    void bound_program(rt_prim& p, size_t leaf0) const {
        p.geom[0] = p.geom[1] = p.geom[2] = 0.f; p.geom[3] = -1.f;
        double lo[3] = {1e300, 1e300, 1e300}, hi[3] = {-1e300, -1e300, -1e300};
        int n = 0;
        for (size_t i = leaf0; i < leaves.size(); ++i) {
            const rt_prim& l = leaves[i];
            if (l.flags & RT_FLAG_NEVER) continue;
            double c[3], R;
            if (!leaf_world_sphere(l, c, R)) return;
            for (int k = 0; k < 3; ++k) { lo[k] = std::min(lo[k], c[k] - R); hi[k] = std::max(hi[k], c[k] + R); }
            ++n;
        }
        if (n == 0) return;
        const double ctr[3] = {0.5 * (lo[0] + hi[0]), 0.5 * (lo[1] + hi[1]), 0.5 * (lo[2] + hi[2])};
        double Rmax = 0;
        for (size_t i = leaf0; i < leaves.size(); ++i) {
            const rt_prim& l = leaves[i];
            if (l.flags & RT_FLAG_NEVER) continue;
            double c[3] = {0, 0, 0}, R = 0;
            leaf_world_sphere(l, c, R);
            const double dx = c[0] - ctr[0], dy = c[1] - ctr[1], dz = c[2] - ctr[2];
            Rmax = std::max(Rmax, std::sqrt(dx * dx + dy * dy + dz * dz) + R);
        }
        if (!std::isfinite(Rmax)) return;
        p.geom[0] = (float)ctr[0]; p.geom[1] = (float)ctr[1]; p.geom[2] = (float)ctr[2];
        p.geom[3] = (float)(Rmax * 1.002 + 1e-3);
    }

    void add_object(const Node& top) {
        Affine M; Ctx ctx;
        const Node* core = &top;
        while (is_transform(*core)) { M = enter(*core, M, ctx); core = core->a.get(); }
        const bool wrapped = core != &top;
        if (core->kind != NodeKind::Csg) {
            rt_prim p = bake_leaf(*core, M, ctx);
            if (wrapped) p.flags |= RT_FLAG_WRAPPED;
            objects.push_back(p);
            return;
        }
        rt_prim p{}; p.kind = RT_PRIM_PROGRAM; p.mat = -1;
        p.flags = (wrapped ? (RT_FLAG_WRAPPED | RT_FLAG_XFORMED) : 0) | (ctx.never ? RT_FLAG_NEVER : 0);
        p.aux = (int)ops.size();
        int max_depth = 0;
        const size_t leaf0 = leaves.size();
        emit(*core, M, ctx, 0, max_depth);
        fuse_sphere_runs((size_t)p.aux);
        p.xf = (int)ops.size() - p.aux;
        bound_program(p, leaf0);
        objects.push_back(p);
    }
};

}  // namespace

struct rt_compiled {
    Compiler c;
    std::vector<rt_dir_light> dir_lights;
    rt_scene_desc desc{};
};

namespace {
// shading.cpp:47: wi = (-light.dir).normalized(); the Ray constructor normalises once more (core.h:278)
rt_dir_light make_dir_light(const rtb::DirectionalLight& L) {
    auto unit = [](double& x, double& y, double& z) {
        const double len = std::sqrt(x * x + y * y + z * z);
        if (len > 1e-6) { x /= len; y /= len; z /= len; } else { x = 0; y = 1; z = 0; }       // core.h:95-101
    };
    double x = -L.dir.x, y = -L.dir.y, z = -L.dir.z;
    unit(x, y, z); unit(x, y, z);
    rt_dir_light l{};
    l.to_light[0] = (float)x; l.to_light[1] = (float)y; l.to_light[2] = (float)z;
    l.radiance[0] = (float)L.radiance.r; l.radiance[1] = (float)L.radiance.g; l.radiance[2] = (float)L.radiance.b;
    return l;
}
}  // namespace

namespace rtb {

// Build the flat description.  Throws std::runtime_error on a limit violation.
rt_compiled* compile_scene(const Scene& scene, const Camera& cam) {
    if (scene.recursion_limit > RT_MAX_RECURSION)
        throw std::runtime_error("medium.recursion " + std::to_string(scene.recursion_limit) + " exceeds the device limit " +
                                 std::to_string(RT_MAX_RECURSION));
    auto out = std::make_unique<rt_compiled>();
    Compiler& c = out->c;
    c.scene = &scene;
    for (const auto& obj : scene.objects) if (obj) c.add_object(*obj);
    for (const auto& L : scene.point_lights) {
        rt_light l{};
        l.pos[0] = (float)L.pos.x; l.pos[1] = (float)L.pos.y; l.pos[2] = (float)L.pos.z;
        l.intensity[0] = (float)L.intensity.r; l.intensity[1] = (float)L.intensity.g; l.intensity[2] = (float)L.intensity.b;
        c.lights.push_back(l);
    }
    for (const auto& L : scene.dir_lights) out->dir_lights.push_back(make_dir_light(L));
    rt_scene_desc& d = out->desc;
    d.n_dir_lights = (int)out->dir_lights.size(); d.dir_lights = out->dir_lights.data();
    d.eye[0] = cam.eye.x; d.eye[1] = cam.eye.y; d.eye[2] = cam.eye.z;
    d.screen_p[0] = cam.screen.P.x; d.screen_p[1] = cam.screen.P.y; d.screen_p[2] = cam.screen.P.z;
    d.lx = cam.screen.Lx; d.ly = cam.screen.Ly;
    d.width = cam.screen.nx(); d.height = cam.screen.ny();
    d.background[0] = (float)scene.background.r; d.background[1] = (float)scene.background.g; d.background[2] = (float)scene.background.b;
    d.medium_index = (float)scene.medium_index;
    d.recursion = scene.recursion_limit;
    d.n_objects = (int)c.objects.size();     d.objects = c.objects.data();
    d.n_leaves = (int)c.leaves.size();       d.leaves = c.leaves.data();
    d.n_pokeballs = (int)c.pokeballs.size(); d.pokeballs = c.pokeballs.data();
    d.n_affines = (int)c.affines.size();     d.affines = c.affines.data();
    d.n_ops = (int)c.ops.size();             d.ops = c.ops.data();
    d.n_materials = (int)c.materials.size(); d.materials = c.materials.data();
    d.n_lights = (int)c.lights.size();       d.lights = c.lights.data();
    d.n_fallbacks = (int)c.fallbacks.size() / 4; d.fallbacks = c.fallbacks.data();
    c.scene = nullptr;
    return out.release();
}

}  // namespace rtb

namespace rtb {
void destroy_compiled(rt_compiled* c) { delete c; }
const rt_scene_desc* desc_of(const rt_compiled* c) { return &c->desc; }
void add_dir_light(rt_compiled* c, const DirectionalLight& L) {
    c->dir_lights.push_back(make_dir_light(L));
    c->desc.n_dir_lights = (int)c->dir_lights.size(); c->desc.dir_lights = c->dir_lights.data();
}
}  // namespace rtb
