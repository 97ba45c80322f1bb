// rt_api.cu — C ABI (include/rt_b200.h) over the sm_100a kernels.
//
// Every CUDA device the process renders on has its own Context (stream, counters, jitter checkpoints, frame buffers, scene
// arenas); the constant scene header is per device anyway.  rt_init selects the process's render device (one process per
// GPU under torchrun); rt_render_multi drives several devices from ONE process (the CLI with RAY_GPUS).  Multi-GPU frames
// need no collective and no peer mapping: every device stores its tiles as packed RGB8 straight into one pinned, mapped
// host frame (rt_host_frame_*), over its own PCIe link, while the other tiles are still being traced.
// There is no CPU fallback: every compute entry point fails with RT_ERR_CUDA without a device.
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <algorithm>
#include <atomic>
#include <cerrno>
#include <chrono>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <map>
#include <memory>
#include <mutex>
#include <string>
#include <vector>

#include "rt_kernels.cuh"
#include "rt_internal.hpp"

struct rt_scene;

namespace {

int fail(int code, const std::string& msg) { rt_set_error(msg); return code; }
#define CU(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) \
    return fail(RT_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(e_)); } while (0)

struct Arena { unsigned char* ptr; size_t bytes; };

struct ContextState {
    int device{-1}, sms{0};
    cudaStream_t stream{nullptr};
    cudaEvent_t ev0{nullptr}, ev1{nullptr};
    // One frame in flight per device: the tile counter, the counters, the wavefront / G-buffer scratch and the constant scene
    // header are shared by all frames of a device.  Every launch records ev_last; a launch on ANOTHER stream waits for it
    // first, so callers may use any stream without overlapping two frames.
    cudaEvent_t ev_last{nullptr}; cudaStream_t last_stream{nullptr}; bool in_flight{false};
    // jitter-stream checkpoints (scene independent) of the tiles rank, rank + world, ... : entry k = tile rank + k * world
    uint32_t* mt_table{nullptr}; long long mt_tiles{0}, mt_capacity{0}; int mt_rank{0}, mt_world{1};
    uint32_t* mt_resume{nullptr}; long long mt_resume_tiles{0};    // sequential walk (test hook): state after mt_resume_tiles tiles
    uint32_t* d_jump_poly{nullptr};                                // mt::kJumpPoly on the device
    unsigned long long* counters{nullptr}; unsigned* tile_counter{nullptr};
    const void* const_owner{nullptr};     // which scene currently occupies constant memory
    // frame buffers of the host-output path (rt_render): owned by the context and grown on demand, so that a
    // fresh scene (the CLI, the bench's end-to-end leg) does not pay cudaMalloc/cudaFree of ~100 MB per frame
    uchar4* d_rgba8{nullptr}; float4* d_rgbf{nullptr}; int* d_ids{nullptr}; unsigned char* d_gbuf{nullptr}; unsigned char* d_pack{nullptr};
    size_t rgba8_cap{0}, rgbf_cap{0}, ids_cap{0}, gbuf_cap{0}, pack_cap{0};     // bytes
    float4* d_wave{nullptr}; size_t wave_cap{0};                                  // wavefront kernel: per-CTA sample state
};
// Context objects are never deleted (scenes keep a pointer to theirs): rt_shutdown releases the device resources and
// resets the state; scenes that were still alive are marked dead and only their host side remains to be freed.
struct Context : ContextState {
    std::vector<Arena> arenas;            // device memory of freed scenes, reused by the next upload (no cudaMalloc per frame)
    std::vector<rt_scene*> live;          // scenes resident on this device
    std::mutex mu;
};
constexpr int kMaxDevices = 64;
Context* g_ctxs[kMaxDevices] = {};        // created on demand (rt_init, rt_init_devices, rt_render_multi), destroyed by rt_shutdown
Context* g_cur = nullptr;                 // the process's render device (rt_init)
std::mutex g_mu;                          // guards g_ctxs / g_cur / the host-frame registry; taken BEFORE a context's mutex, never after
// Test hooks (rt_debug_set_option): -1 = the library's own choice, 0 / 1 = forced.  Never read from the environment.
struct DebugOptions { std::atomic<int> flat{-1}, cull{-1}, wavefront{-1}, global_scene{-1}, mt_sequential{-1}, grid{-1}, grid_cells{-1}; };
DebugOptions g_opts;                      // (atomics: read under a context's mutex, written under g_mu)
constexpr size_t kSceneSmemBudget = 80 * 1024;   // + 25 KB of MT19937 buffers: two CTAs per SM still fit in 227 KB

}  // namespace

struct rt_scene {
    rtd::SceneHeader hdr{};
    Context* cx{nullptr};     // the device the scene lives on
    bool alive{true};         // false once rt_shutdown released the device under it
    bool fits_const{false};   // scanned arrays fit the shared-memory budget (staged per CTA)
    bool flat{false};         // no ray can spawn a child (no reflective / refractive material, or recursion == 1)
    bool bounces{false};      // most materials reflect or refract and medium.recursion >= 2: deep, uneven ray trees
    bool clustered{false};    // a long run of plain spheres / pokeballs was reordered into spatial clusters (kernel variant K)
    size_t smem_bytes{0};
    std::vector<float4> obj_geom, leaf_geom, pokeballs, affines, materials, lights, fallbacks, clusters, masks, cull_geom;
    std::vector<uint32_t> cull_plane_bits;
    int cull_words{0}, cull_leaf_word{0};
    rtd::GridHeader grid{};                // uniform grid over the plain spheres / pokeballs (rt_grid.cuh); words == 0: none
    std::vector<uint32_t> grid_masks;
    float mask_scale{0.f};
    std::vector<int4> obj_meta, leaf_meta, runs;
    std::vector<int> obj_orig;
    std::vector<uint32_t> ops;
    Arena arena{nullptr, 0};  // ONE device allocation holds every array below (one H2D copy per upload)
    int width{0}, height{0};
};

namespace {

int4 meta_of(const rt_prim& p) { return make_int4((p.kind & 0xff) | (p.flags << 8), p.mat, p.aux, p.xf); }

// ---- clustered runs ------------------------------------------------------------------------------------------
// A long run of plain spheres / pokeballs is reordered spatially (recursive median split along the longest axis,
// leaves of exactly kClusterSize primitives) and every kClusterSize consecutive primitives get a conservative
// bounding sphere.  The scans test the bounds first (same candidate-mask loop), so a query costs
// n/16 + 16 x (clusters whose bound it crosses) tests instead of n.  The reference has no acceleration structure
// (scene.cpp:10-42 is a linear scan); results are identical because a sphere's candidate root does not depend on
// the scan order and exact ties are resolved by JSON index on the device (obj_orig).
constexpr int kClusterSize = 16;      // the device code shifts by 4
constexpr int kClusterMinRun = 48;

void split_run(std::vector<int>& idx, int lo, int hi, const rt_prim* objs) {
    if (hi - lo <= kClusterSize) return;
    float mn[3] = {1e30f, 1e30f, 1e30f}, mx[3] = {-1e30f, -1e30f, -1e30f};
    for (int k = lo; k < hi; ++k) for (int a = 0; a < 3; ++a) { mn[a] = std::min(mn[a], objs[idx[k]].geom[a]); mx[a] = std::max(mx[a], objs[idx[k]].geom[a]); }
    int axis = 0; for (int a = 1; a < 3; ++a) if (mx[a] - mn[a] > mx[axis] - mn[axis]) axis = a;
    // split at a multiple of the cluster size so that only the very last cluster of the run can be partial
    const int n_cl = (hi - lo + kClusterSize - 1) / kClusterSize;
    const int mid = lo + (n_cl / 2) * kClusterSize;
    std::nth_element(idx.begin() + lo, idx.begin() + mid, idx.begin() + hi, [&](int a, int b) {
        const float ga = objs[a].geom[axis], gb = objs[b].geom[axis];
        return ga < gb || (ga == gb && a < b);
    });
    split_run(idx, lo, mid, objs);
    split_run(idx, mid, hi, objs);
}

float4 bound_of(const rt_prim* objs, const int* idx, int n) {
    double lo[3] = {1e300, 1e300, 1e300}, hi[3] = {-1e300, -1e300, -1e300};
    for (int k = 0; k < n; ++k) for (int a = 0; a < 3; ++a) {
        const rt_prim& p = objs[idx[k]];
        lo[a] = std::min(lo[a], (double)p.geom[a] - std::abs((double)p.geom[3]));
        hi[a] = std::max(hi[a], (double)p.geom[a] + std::abs((double)p.geom[3]));
    }
    const double c[3] = {0.5 * (lo[0] + hi[0]), 0.5 * (lo[1] + hi[1]), 0.5 * (lo[2] + hi[2])};
    double R = 0;
    for (int k = 0; k < n; ++k) {
        const rt_prim& p = objs[idx[k]];
        const double dx = p.geom[0] - c[0], dy = p.geom[1] - c[1], dz = p.geom[2] - c[2];
        R = std::max(R, std::sqrt(dx * dx + dy * dy + dz * dz) + std::abs((double)p.geom[3]));
    }
    // inflated far beyond the FP32 rounding of the device discriminant (same margin as CSG program bounds)
    return make_float4((float)c[0], (float)c[1], (float)c[2], (float)(R * 1.002 + 1e-3));
}

void build_host_arrays(const rt_scene_desc& d, rt_scene& s, bool want_grid = false, double grid_cells_per_sphere = 2.0) {
    // device order of the top-level objects: JSON order, except inside clustered runs (see above)
    std::vector<int> order(d.n_objects);
    for (int i = 0; i < d.n_objects; ++i) order[i] = i;
    struct ClusteredRun { int first, count, ctab; };
    std::vector<ClusteredRun> clustered;
    for (int i = 0; i < d.n_objects;) {
        const rt_prim& p = d.objects[i];
        const bool plain = p.flags == 0 && (p.kind == RT_PRIM_SPHERE || p.kind == RT_PRIM_POKEBALL);
        if (!plain) { ++i; continue; }
        int j = i; while (j < d.n_objects && d.objects[j].flags == 0 && d.objects[j].kind == p.kind) ++j;
        bool finite = true;
        for (int k = i; k < j; ++k) for (int a = 0; a < 4; ++a) finite = finite && std::isfinite(d.objects[k].geom[a]);
        if (j - i >= kClusterMinRun && finite) {
            std::vector<int> trial(order);
            split_run(trial, i, j, d.objects);
            std::vector<float4> bounds;
            for (int c = i; c < j; c += kClusterSize) bounds.push_back(bound_of(d.objects, trial.data() + c, std::min(kClusterSize, j - c)));
            // Expected tests per query if a line crosses a bound with probability ~ (R_cluster / R_run)^2.  Rays are
            // aimed at the scene, so the estimate is optimistic; cluster only when it promises < 0.35 n
            // (measured: 64 pokeballs on a grid and 1 000 scattered spheres gain, a dense 75-sphere lattice loses 4 %).
            const float4 whole = bound_of(d.objects, trial.data() + i, j - i);
            double est = (double)bounds.size();
            for (const float4& b : bounds) est += kClusterSize * std::min(1.0, (double)b.w * b.w / ((double)whole.w * whole.w));
            if (est < 0.35 * (j - i)) {
                order.swap(trial);
                clustered.push_back({i, j - i, (int)s.clusters.size()});
                s.clustered = true;
                s.clusters.insert(s.clusters.end(), bounds.begin(), bounds.end());
            }
        }
        i = j;
    }
    for (int i = 0; i < d.n_objects; ++i) {
        const rt_prim& p = d.objects[order[i]];
        s.obj_geom.push_back(make_float4(p.geom[0], p.geom[1], p.geom[2], p.geom[3]));
        s.obj_meta.push_back(meta_of(p));
        s.obj_orig.push_back(order[i]);
    }
    for (int i = 0; i < d.n_leaves; ++i) {
        const rt_prim& p = d.leaves[i];
        s.leaf_geom.push_back(make_float4(p.geom[0], p.geom[1], p.geom[2], p.geom[3]));
        s.leaf_meta.push_back(meta_of(p));
    }
    for (int i = 0; i < d.n_pokeballs; ++i) {
        const rt_pokeball& b = d.pokeballs[i];
        s.pokeballs.push_back(make_float4(b.btn_dir[0], b.btn_dir[1], b.btn_dir[2], b.cos_outer));
        s.pokeballs.push_back(make_float4(b.up[0], b.up[1], b.up[2], b.cos_inner));
        s.pokeballs.push_back(make_float4(b.belt_half, 0.f, 0.f, 0.f));
    }
    for (int i = 0; i < d.n_affines; ++i)
        for (int r = 0; r < 3; ++r) s.affines.push_back(make_float4(d.affines[i].m[r][0], d.affines[i].m[r][1], d.affines[i].m[r][2], d.affines[i].m[r][3]));
    for (int i = 0; i < d.n_ops; ++i) s.ops.push_back(d.ops[i]);
    for (int i = 0; i < d.n_materials; ++i) {
        const rt_material& m = d.materials[i];
        s.materials.push_back(make_float4(m.albedo[0], m.albedo[1], m.albedo[2], m.ks));
        s.materials.push_back(make_float4(m.ambient[0], m.ambient[1], m.ambient[2], m.kr));
        s.materials.push_back(make_float4(m.kt, m.shininess, m.ior, m.kd));
    }
    // directional lights first (shading.cpp:46 before :79), marked by w = 1 and carrying the unit vector TO the light
    for (int i = 0; i < d.n_dir_lights; ++i) {
        s.lights.push_back(make_float4(d.dir_lights[i].to_light[0], d.dir_lights[i].to_light[1], d.dir_lights[i].to_light[2], 1.f));
        s.lights.push_back(make_float4(d.dir_lights[i].radiance[0], d.dir_lights[i].radiance[1], d.dir_lights[i].radiance[2], 0.f));
    }
    for (int i = 0; i < d.n_lights; ++i) {
        s.lights.push_back(make_float4(d.lights[i].pos[0], d.lights[i].pos[1], d.lights[i].pos[2], 0.f));
        s.lights.push_back(make_float4(d.lights[i].intensity[0], d.lights[i].intensity[1], d.lights[i].intensity[2], 0.f));
    }
    for (int i = 0; i < d.n_fallbacks; ++i) s.fallbacks.push_back(make_float4(d.fallbacks[4 * i], d.fallbacks[4 * i + 1], d.fallbacks[4 * i + 2], 0.f));
    // runs of consecutive plain primitives of one kind (flags == 0), everything else is a 1-object generic run
    for (int i = 0; i < d.n_objects;) {
        const rt_prim& p = d.objects[i];
        const bool plain = p.flags == 0 && p.kind != RT_PRIM_PROGRAM;
        if (!plain) {                      // maximal run of generic objects (CSG programs, transform-wrapped / affine primitives)
            int j = i; while (j < d.n_objects && !(d.objects[j].flags == 0 && d.objects[j].kind != RT_PRIM_PROGRAM)) ++j;
            int ctab = 0;
            {
                // one conservative WORLD bounding sphere per object (R = inf: unknown): the scans cull the bounds against the
                // warp's ray bundle (or, for incoherent bundles, scan them with the candidate mask) before the object itself
                // is looked at
                ctab = (int)s.clusters.size() + 1;
                for (int k = i; k < j; ++k) {
                    const rt_prim& q = d.objects[k];
                    float4 b = make_float4(0.f, 0.f, 0.f, std::numeric_limits<float>::infinity());
                    const bool finite = std::isfinite(q.geom[0]) && std::isfinite(q.geom[1]) && std::isfinite(q.geom[2]) && std::isfinite(q.geom[3]);
                    if (q.kind == RT_PRIM_PROGRAM) { if (q.geom[3] >= 0.f && finite) b = make_float4(q.geom[0], q.geom[1], q.geom[2], q.geom[3]); }
                    else if (q.kind != RT_PRIM_HALFSPACE && !(q.flags & RT_FLAG_AFFINE) && finite) {
                        int one = k;                // baked similarity transforms: geom is the world-space sphere
                        b = bound_of(d.objects, &one, 1);
                    }
                    s.clusters.push_back(b);
                }
            }
            s.runs.push_back(make_int4(rtd::RUN_GENERIC, i, j - i, ctab));
            i = j; continue;
        }
        int j = i; while (j < d.n_objects && d.objects[j].flags == 0 && d.objects[j].kind == p.kind) ++j;
        const int kind = p.kind == RT_PRIM_SPHERE ? rtd::RUN_SPHERE : p.kind == RT_PRIM_HALFSPACE ? rtd::RUN_HALFSPACE : rtd::RUN_POKEBALL;
        int ctab = 0;
        for (const auto& c : clustered) if (c.first == i && c.count == j - i) ctab = c.ctab + 1;
        s.runs.push_back(make_int4(kind, i, j - i, ctab));
        i = j;
    }
    // Uniform grid (rt_grid.cuh) for scenes whose rays are mostly incoherent (the caller decides): every plain sphere /
    // pokeball — all of them sit in unclustered runs here — is binned by its device index.  (The kernels stay the ones
    // without bounds: the out-of-line scan of the K kernels costs this kind of scene 16 %.)
    constexpr int kGridMinSpheres = 24;
    if (want_grid && !s.clustered && d.n_objects <= 32 * rtd::kGridMaxWords) {
        std::vector<float> sph; std::vector<int> bit;
        for (size_t i = 0; i < s.obj_geom.size(); ++i) {
            const int kind = s.obj_meta[i].x & 0xff, flags = s.obj_meta[i].x >> 8;
            if (flags == 0 && (kind == RT_PRIM_SPHERE || kind == RT_PRIM_POKEBALL)) {
                sph.insert(sph.end(), {s.obj_geom[i].x, s.obj_geom[i].y, s.obj_geom[i].z, s.obj_geom[i].w});
                bit.push_back((int)i);
            }
        }
        if ((int)bit.size() >= kGridMinSpheres) rtd::grid_build(sph.data(), bit.data(), (int)bit.size(), d.n_objects, s.grid, s.grid_masks, grid_cells_per_sphere);
    }
    // Mask table of sphere_mask: (c, |c|^2 - r^2 - 2^-16 (|c|^2 + r^2)) for every entry of obj_geom | leaf_geom | clusters
    // (entries that are not sphere-like are never scanned through it): the primitive's own share of the scan's rounding
    // margin is folded into w, so the margin is per primitive.  `scale` = max(|c|^2 + r^2) is only reported.
    double scale = 0.0;
    auto add = [&](const float4& g, bool scanned) {
        const double cc = (double)g.x * g.x + (double)g.y * g.y + (double)g.z * g.z, rr = (double)g.w * g.w;
        s.masks.push_back(make_float4(g.x, g.y, g.z, (float)((cc - rr) - (double)rtd::kMaskMargin * (cc + rr))));
        if (scanned && std::isfinite(cc + rr)) scale = std::max(scale, cc + rr);
    };
    auto sphere_like = [](const int4& m) { return (m.x & 0xff) != RT_PRIM_HALFSPACE && !((m.x >> 8) & RT_FLAG_AFFINE); };
    for (size_t i = 0; i < s.obj_geom.size(); ++i) add(s.obj_geom[i], sphere_like(s.obj_meta[i]));
    for (size_t i = 0; i < s.leaf_geom.size(); ++i) add(s.leaf_geom[i], sphere_like(s.leaf_meta[i]));
    for (const float4& g : s.clusters) add(g, true);
    s.mask_scale = (float)std::min(scale * (1.0 + 1e-6), 3.0e38);
    // Round-level culling table (rt_cull.cuh, rt_device.cuh cand_build): one entry per top-level object (device order) and
    // per CSG leaf: (centre, radius) of a plain sphere / pokeball; the world bound of a generic object (CSG program, wrapped
    // primitive; the run's bound table); (n, n.p0) of a plain half-space, flagged in cull_plane_bits; radius = inf where
    // nothing can be said (always a candidate).  Both sections are padded to a multiple of 32 entries with w < 0 (never).
    const float inf = std::numeric_limits<float>::infinity();
    auto pad32 = [&]() { while (s.cull_geom.size() % 32) s.cull_geom.push_back(make_float4(0.f, 0.f, 0.f, -1.f)); };
    auto finite4 = [](const float4& g) { return std::isfinite(g.x) && std::isfinite(g.y) && std::isfinite(g.z) && std::isfinite(g.w); };
    std::vector<int> bound_of_obj(s.obj_geom.size(), -1);          // generic objects: index of their bound in s.clusters
    for (const int4& r : s.runs) if (r.x == rtd::RUN_GENERIC && r.w > 0) for (int k = 0; k < r.z; ++k) bound_of_obj[(size_t)(r.y + k)] = r.w - 1 + k;
    std::vector<size_t> planes;
    for (size_t i = 0; i < s.obj_geom.size(); ++i) {
        const float4 g = s.obj_geom[i]; const int4 m = s.obj_meta[i];
        const int kind = m.x & 0xff, flags = m.x >> 8;
        float4 e = make_float4(0.f, 0.f, 0.f, inf);
        if (flags == 0 && (kind == RT_PRIM_SPHERE || kind == RT_PRIM_POKEBALL)) { if (finite4(g)) e = make_float4(g.x, g.y, g.z, std::fabs(g.w)); }
        else if (flags == 0 && kind == RT_PRIM_HALFSPACE) {
            const double nn = (double)g.x * g.x + (double)g.y * g.y + (double)g.z * g.z;
            if (finite4(g) && std::fabs(nn - 1.0) < 1e-5) { e = g; planes.push_back(i); }      // the plane cull assumes a unit normal
        } else if (bound_of_obj[i] >= 0) {
            const float4 b = s.clusters[(size_t)bound_of_obj[i]];
            if (finite4(b) && b.w >= 0.f) e = b;
        }
        s.cull_geom.push_back(e);
    }
    pad32(); s.cull_leaf_word = (int)(s.cull_geom.size() / 32);
    for (size_t i = 0; i < s.leaf_geom.size(); ++i) {
        const float4 g = s.leaf_geom[i]; const int4 m = s.leaf_meta[i];
        const bool plain = (m.x & 0xff) == RT_PRIM_SPHERE && !((m.x >> 8) & (RT_FLAG_AFFINE | RT_FLAG_NEVER)) && finite4(g);
        s.cull_geom.push_back(plain ? make_float4(g.x, g.y, g.z, std::fabs(g.w)) : make_float4(0.f, 0.f, 0.f, inf));
    }
    pad32(); s.cull_words = (int)(s.cull_geom.size() / 32);
    // three bit planes of W = max(cull_words, 1) words each: [0, W) entry is culled as a plane; [W, 2W) object is a plain
    // half-space (kind, whatever its normal's length); [2W, 3W) object is generic (CSG program, wrapped / affine primitive):
    // the masked scans classify their candidates from these words instead of loading each candidate's meta record
    const size_t Wc = (size_t)std::max(s.cull_words, 1);
    s.cull_plane_bits.assign(3 * Wc, 0u);
    for (size_t i : planes) s.cull_plane_bits[i >> 5] |= 1u << (i & 31);
    for (size_t i = 0; i < s.obj_meta.size(); ++i) {
        const int kind = s.obj_meta[i].x & 0xff, flags = s.obj_meta[i].x >> 8;
        if (flags == 0 && kind == RT_PRIM_HALFSPACE) s.cull_plane_bits[Wc + (i >> 5)] |= 1u << (i & 31);
        if (flags != 0 || kind == RT_PRIM_PROGRAM) s.cull_plane_bits[2 * Wc + (i >> 5)] |= 1u << (i & 31);
    }
}


// ---- contexts ---------------------------------------------------------------------------------------------------
int init_context(Context& cx, int device) {
    CU(cudaSetDevice(device));
    cudaDeviceProp prop; CU(cudaGetDeviceProperties(&prop, device));
    cx.device = device; cx.sms = prop.multiProcessorCount;
    CU(cudaStreamCreateWithFlags(&cx.stream, cudaStreamNonBlocking));
    CU(cudaEventCreate(&cx.ev0)); CU(cudaEventCreate(&cx.ev1));
    CU(cudaEventCreateWithFlags(&cx.ev_last, cudaEventDisableTiming));
    CU(cudaMalloc((void**)&cx.counters, 4 * sizeof(unsigned long long)));
    CU(cudaMalloc((void**)&cx.tile_counter, sizeof(unsigned)));
    CU(cudaMalloc((void**)&cx.mt_resume, mt::kN * sizeof(uint32_t)));
    uint32_t seed[mt::kN]; mt::seed_state(12345u, seed);                  // tracer.cpp:284
    CU(cudaMemcpy(cx.mt_resume, seed, sizeof(seed), cudaMemcpyHostToDevice));
    const int dyn_smem = (int)kSceneSmemBudget + 4096;
    CU(cudaFuncSetAttribute(rtd::render_standard_kernel<true, false, false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, dyn_smem));
    CU(cudaFuncSetAttribute(rtd::render_standard_kernel<true, false, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, dyn_smem));
    CU(cudaFuncSetAttribute(rtd::render_standard_kernel<true, false, true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, dyn_smem));
    CU(cudaFuncSetAttribute(rtd::render_standard_kernel<true, false, true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, dyn_smem));
    CU(cudaFuncSetAttribute(rtd::render_standard_kernel<true, true, false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, dyn_smem));
    CU(cudaFuncSetAttribute(rtd::render_standard_kernel<true, true, true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, dyn_smem));
    CU(cudaFuncSetAttribute(rtd::render_wavefront_kernel<true, false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, dyn_smem));
    CU(cudaFuncSetAttribute(rtd::render_wavefront_kernel<true, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, dyn_smem));
    CU(cudaFuncSetAttribute(rtd::render_wavefront_kernel<true, true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, dyn_smem));
    CU(cudaFuncSetAttribute(rtd::paper_gbuffer_kernel<true, false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, dyn_smem));
    CU(cudaFuncSetAttribute(rtd::paper_gbuffer_kernel<true, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, dyn_smem));
    CU(cudaFuncSetAttribute(rtd::paper_gbuffer_kernel<true, true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, dyn_smem));
    CU(cudaFuncSetAttribute(rtd::trace_rays_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, dyn_smem));
    CU(cudaFuncSetAttribute(rtd::mt_jump_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, rtd::kJumpSmemBytes));
    return RT_OK;
}

void destroy_context(Context& cx) {
    std::lock_guard<std::mutex> lk(cx.mu);
    if (cx.device < 0) return;
    cudaSetDevice(cx.device);
    cudaDeviceSynchronize();
    for (rt_scene* s : cx.live) { cudaFree(s->arena.ptr); s->arena = Arena{nullptr, 0}; s->alive = false; }
    cx.live.clear();
    cudaFree(cx.mt_table); cudaFree(cx.mt_resume); cudaFree(cx.counters); cudaFree(cx.tile_counter); cudaFree(cx.d_jump_poly);
    cudaFree(cx.d_rgba8); cudaFree(cx.d_rgbf); cudaFree(cx.d_ids); cudaFree(cx.d_gbuf); cudaFree(cx.d_pack); cudaFree(cx.d_wave);
    for (const Arena& a : cx.arenas) cudaFree(a.ptr);
    if (cx.ev0) cudaEventDestroy(cx.ev0);
    if (cx.ev1) cudaEventDestroy(cx.ev1);
    if (cx.ev_last) cudaEventDestroy(cx.ev_last);
    if (cx.stream) cudaStreamDestroy(cx.stream);
    cx.arenas.clear();
    static_cast<ContextState&>(cx) = ContextState{};
}

// The context of `device`, created on first use (g_mu held by the caller).
int context_of(int device, Context** out) {
    if (device < 0 || device >= kMaxDevices) return fail(RT_ERR_ARG, "device ordinal out of range");
    if (!g_ctxs[device]) g_ctxs[device] = new Context();
    if (g_ctxs[device]->device < 0) {
        const int rc = init_context(*g_ctxs[device], device);
        if (rc) { g_ctxs[device]->device = device; destroy_context(*g_ctxs[device]); return rc; }
    }
    *out = g_ctxs[device];
    return RT_OK;
}

int device_count(int* count) {
    *count = 0;
    cudaError_t e = cudaGetDeviceCount(count);
    if (e != cudaSuccess || *count == 0)
        return fail(RT_ERR_CUDA, std::string("no CUDA device: ") + (e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0") +
                                     " (this library has no CPU fallback)");
    return RT_OK;
}

Context* current_context() { std::lock_guard<std::mutex> lk(g_mu); return g_cur; }

// ---- scenes -------------------------------------------------------------------------------------------------------
// Every device array of a scene lives in ONE arena (256-byte aligned sections), filled by one H2D copy; arenas of freed
// scenes are kept per context and reused, so a fresh scene per frame (the CLI, the end-to-end bench leg) costs no
// cudaMalloc / cudaFree.
struct Blob {
    std::vector<unsigned char> bytes;
    template <class T> size_t add(const std::vector<T>& v) {
        const size_t off = (bytes.size() + 255) & ~(size_t)255;
        bytes.resize(off + std::max<size_t>(v.size(), 1) * sizeof(T), 0);
        if (!v.empty()) std::memcpy(bytes.data() + off, v.data(), v.size() * sizeof(T));
        return off;
    }
};

int arena_take(Context& cx, size_t bytes, Arena* out) {
    size_t best = cx.arenas.size();
    for (size_t i = 0; i < cx.arenas.size(); ++i)
        if (cx.arenas[i].bytes >= bytes && (best == cx.arenas.size() || cx.arenas[i].bytes < cx.arenas[best].bytes)) best = i;
    if (best != cx.arenas.size()) { *out = cx.arenas[best]; cx.arenas.erase(cx.arenas.begin() + (long)best); return RT_OK; }
    size_t cap = 4096; while (cap < bytes) cap *= 2;
    out->ptr = nullptr; out->bytes = cap;
    CU(cudaMalloc((void**)&out->ptr, cap));
    return RT_OK;
}
void arena_give(Context& cx, Arena a) {
    if (!a.ptr) return;
    if (cx.arenas.size() >= 8) { cudaFree(a.ptr); return; }
    cx.arenas.push_back(a);
}

void free_scene_locked(rt_scene* s) {
    Context& cx = *s->cx;
    if (!s->alive) { delete s; return; }
    cx.live.erase(std::remove(cx.live.begin(), cx.live.end(), s), cx.live.end());
    if (cx.const_owner == s) cx.const_owner = nullptr;
    if (cx.in_flight) { cudaSetDevice(cx.device); cudaEventSynchronize(cx.ev_last); cx.in_flight = false; }   // its arena may be reused at once
    arena_give(cx, s->arena);
    delete s;
}

int upload_scene(Context& cx, const rt_scene_desc* d, rt_scene** out) {
    if (d->width <= 0 || d->height <= 0) return fail(RT_ERR_ARG, "rt_scene_upload: empty screen");
    if (d->recursion > RT_MAX_RECURSION) return fail(RT_ERR_LIMIT, "medium.recursion exceeds RT_MAX_RECURSION");
    std::lock_guard<std::mutex> lk(cx.mu);
    CU(cudaSetDevice(cx.device));
    auto* s = new rt_scene();
    s->cx = &cx;
    // bounce-heavy scenes (the wavefront kernel's): most rays are secondary rays that belong to no bundle -> uniform grid
    int n_bounce0 = 0;
    for (int i = 0; i < d->n_materials; ++i) n_bounce0 += (d->materials[i].kr > 0.f || d->materials[i].kt > 0.f) ? 1 : 0;
    bool want_grid = d->recursion >= 2 && d->n_materials > 0 && 2 * n_bounce0 >= d->n_materials;
    if (g_opts.wavefront.load() >= 0) want_grid = g_opts.wavefront.load() == 1;
    if (g_opts.grid.load() >= 0) want_grid = g_opts.grid.load() == 1;       // test hook
    build_host_arrays(*d, *s, want_grid, g_opts.grid_cells.load() > 0 ? (double)g_opts.grid_cells.load() : 2.0);     // measured on synthetic-B 1K: 1 / 2 / 4 / 8 / 16 / 32 cells per sphere = 29.7 / 28.5 / 29.5 / 29.8 / 34.5 / 41.4 ms (no grid: 31.7)
    Blob blob;
    const size_t o_obj_geom = blob.add(s->obj_geom), o_obj_meta = blob.add(s->obj_meta), o_leaf_geom = blob.add(s->leaf_geom),
                 o_leaf_meta = blob.add(s->leaf_meta), o_pokeballs = blob.add(s->pokeballs), o_affines = blob.add(s->affines),
                 o_ops = blob.add(s->ops), o_materials = blob.add(s->materials), o_lights = blob.add(s->lights),
                 o_fallbacks = blob.add(s->fallbacks), o_runs = blob.add(s->runs), o_clusters = blob.add(s->clusters),
                 o_obj_orig = blob.add(s->obj_orig), o_masks = blob.add(s->masks), o_cull_geom = blob.add(s->cull_geom),
                 o_cull_plane_bits = blob.add(s->cull_plane_bits), o_grid_masks = blob.add(s->grid_masks);
    int rc = arena_take(cx, blob.bytes.size(), &s->arena);
    if (rc) { delete s; return rc; }
    cx.live.push_back(s);
    {
        cudaError_t e = cudaMemcpy(s->arena.ptr, blob.bytes.data(), blob.bytes.size(), cudaMemcpyHostToDevice);
        if (e != cudaSuccess) { free_scene_locked(s); return fail(RT_ERR_CUDA, std::string("scene upload: ") + cudaGetErrorString(e)); }
    }
    unsigned char* base = s->arena.ptr;
    rtd::SceneHeader& h = s->hdr;
    for (int k = 0; k < 3; ++k) { h.eye[k] = d->eye[k]; h.P[k] = d->screen_p[k]; h.bg[k] = d->background[k]; }
    h.lx = d->lx; h.ly = d->ly; h.nx = d->width; h.ny = d->height;
    h.inv_nx = 1.0 / (double)d->width; h.inv_ny = 1.0 / (double)d->height;
    h.medium_index = d->medium_index; h.recursion = d->recursion;
    h.n_objects = d->n_objects; h.n_lights = d->n_dir_lights + d->n_lights; h.n_runs = (int)s->runs.size();
    h.obj_geom = (const float4*)(base + o_obj_geom); h.obj_meta = (const int4*)(base + o_obj_meta);
    h.leaf_geom = (const float4*)(base + o_leaf_geom); h.leaf_meta = (const int4*)(base + o_leaf_meta);
    h.pokeballs = (const float4*)(base + o_pokeballs); h.affines = (const float4*)(base + o_affines);
    h.ops = (const uint32_t*)(base + o_ops); h.materials = (const float4*)(base + o_materials);
    h.lights = (const float4*)(base + o_lights); h.fallbacks = (const float4*)(base + o_fallbacks);
    h.runs = (const int4*)(base + o_runs); h.clusters = (const float4*)(base + o_clusters); h.obj_orig = (const int*)(base + o_obj_orig);
    h.n_clusters = (int)s->clusters.size();
    h.masks = (const float4*)(base + o_masks); h.mask_scale = s->mask_scale;
    h.grid = s->grid; h.grid_masks = (const uint32_t*)(base + o_grid_masks);
    s->width = d->width; s->height = d->height;
    h.n_leaves = d->n_leaves; h.n_ops = d->n_ops;
    h.off_obj_meta = s->clustered ? 0 : d->n_objects;                // scenes with clustered runs read top-level geometry / meta from global memory
    h.off_leaf_geom = 2 * h.off_obj_meta;
    h.off_leaf_meta = h.off_leaf_geom + d->n_leaves; h.off_runs = h.off_leaf_meta + d->n_leaves;
    h.off_lights = h.off_runs + (int)s->runs.size();
    h.off_clusters = h.off_lights + 2 * h.n_lights;
    h.off_masks = h.off_clusters + h.n_clusters;
    h.off_ops_bytes = 16 * (h.off_masks + (int)s->masks.size());
    h.smem_bytes = h.off_ops_bytes + 4 * d->n_ops;
    h.n_grid_cells = (int)(s->grid_masks.size() / rtd::kGridMaxWords);
    h.off_grid_bytes = (h.smem_bytes + 15) & ~15;
    if (h.n_grid_cells) h.smem_bytes = h.off_grid_bytes + 16 * h.n_grid_cells;
    s->fits_const = (size_t)h.smem_bytes <= kSceneSmemBudget;
    // Wavefront kernel only where most surfaces spawn secondary rays: its per-sample state traffic costs ~10 % on scenes
    // whose ray trees are mostly depth 1 (snorlax with one reflective sphere), and gains ~6 % on the all-glass lattice.
    int n_bounce = 0;
    for (int i = 0; i < d->n_materials; ++i) n_bounce += (d->materials[i].kr > 0.f || d->materials[i].kt > 0.f) ? 1 : 0;
    struct { int flat, cull, wavefront, global_scene; } opts = {g_opts.flat.load(), g_opts.cull.load(), g_opts.wavefront.load(), g_opts.global_scene.load()};
    s->flat = d->recursion >= 1 && (n_bounce == 0 || d->recursion == 1);
    if (opts.flat == 0) s->flat = false;                                  // test hook: generic loop on a flat scene
    s->bounces = d->recursion >= 2 && d->n_materials > 0 && 2 * n_bounce >= d->n_materials;
    if (opts.wavefront >= 0) s->bounces = opts.wavefront == 1;            // test hook: either kernel on any scene
    if (opts.global_scene == 1) s->fits_const = false;                    // test hook: the global-memory path on any scene
    // Round-level culling: masks of (1 + kMaxListLights) bundles behind the staged scene (or alone, on the global path).
    // Round-level masks pay when a scan is more than a handful of tests and the table is short enough for a CTA to cull per
    // round: not for scenes of a few primitives (pokeballs: 4 objects), not for clustered scenes (their scans already walk a
    // spatial structure; culling 1 300 entries per bundle and round cost more than it saved: synthetic-A 2K 23.8 -> 41 ms),
    // not for the wavefront kernel (only its camera pass could use them: synthetic-B 1K 32.1 -> 37.3 ms).
    constexpr int kMaxCullWords = 8, kMinCullEntries = 10;
    bool cull = s->cull_words > 0 && s->cull_words <= kMaxCullWords && !s->clustered && !s->bounces && d->n_objects + d->n_leaves >= kMinCullEntries;
    if (opts.cull == 0) cull = false;
    if (opts.cull == 1) cull = s->cull_words > 0 && s->cull_words <= rtd::kCandWordsMax && !s->clustered;        // test hook: masks wherever they can be built
    h.cull_geom = (const float4*)(base + o_cull_geom); h.cull_words = cull ? s->cull_words : 0;
    h.cull_plane_bits = (const uint32_t*)(base + o_cull_plane_bits); h.cull_leaf_word = s->cull_leaf_word;
    h.off_cand_bytes = s->fits_const ? ((h.smem_bytes + 15) & ~15) : 0;
    s->smem_bytes = (size_t)h.off_cand_bytes + ((cull && !RT_CAND_STATIC) ? sizeof(rtd::CandShared) + (size_t)(1 + rtd::kMaxListLights) * s->cull_words * 4 : 0);
    *out = s;
    return RT_OK;
}

// Make `s` the scene held in the device's constant memory.  (The copy is staged from pageable memory before the call
// returns and is ordered before the kernels that follow on `st`.)
int bind_scene(Context& cx, rt_scene* s, cudaStream_t st) {
    if (cx.const_owner == s) return RT_OK;
    CU(cudaMemcpyToSymbolAsync(rtd::c_hdr, &s->hdr, sizeof(s->hdr), 0, cudaMemcpyHostToDevice, st));
    cx.const_owner = s;
    return RT_OK;
}

// One frame in flight per device (see Context): work on a stream other than the last frame's waits for that frame.
int frame_begin(Context& cx, cudaStream_t st) {
    if (cx.in_flight && cx.last_stream != st) CU(cudaStreamWaitEvent(st, cx.ev_last, 0));
    return RT_OK;
}
int frame_end(Context& cx, cudaStream_t st) {
    CU(cudaEventRecord(cx.ev_last, st));
    cx.last_stream = st; cx.in_flight = true;
    return RT_OK;
}

// The jitter checkpoint table of the tiles a rank owns among the first `tiles` tiles (scene / resolution independent):
// entry k = generator state at tile rank + k * world.  A rank keeps 1/world of the table (16384^2: 4.3 GB -> 0.54 GB on 8 GPUs).
int ensure_mt_table(Context& cx, long long tiles, int rank, int world) {
    if (rank != cx.mt_rank || world != cx.mt_world) {          // another partition: rebuilt from the seed (7.5 ms at 4K)
        if (cx.in_flight) { CU(cudaEventSynchronize(cx.ev_last)); cx.in_flight = false; }
        cx.mt_tiles = 0; cx.mt_rank = rank; cx.mt_world = world;
        if (cx.mt_resume_tiles) {
            uint32_t seed[mt::kN]; mt::seed_state(12345u, seed);
            CU(cudaMemcpy(cx.mt_resume, seed, sizeof(seed), cudaMemcpyHostToDevice));
            cx.mt_resume_tiles = 0;
        }
    }
    if (tiles <= cx.mt_tiles) return RT_OK;
    if (cx.in_flight) { CU(cudaEventSynchronize(cx.ev_last)); cx.in_flight = false; }      // the table may move
    const long long local = (tiles - rank + world - 1) / world;
    if (local > cx.mt_capacity) {
        uint32_t* nt = nullptr;
        CU(cudaMalloc((void**)&nt, (size_t)local * mt::kN * sizeof(uint32_t)));
        const long long have = cx.mt_tiles > rank ? (cx.mt_tiles - rank + world - 1) / world : 0;
        if (have) CU(cudaMemcpy(nt, cx.mt_table, (size_t)have * mt::kN * sizeof(uint32_t), cudaMemcpyDeviceToDevice));
        cudaFree(cx.mt_table);
        cx.mt_table = nt; cx.mt_capacity = local;
    }
    const int sequential = g_opts.mt_sequential.load();
    if (sequential == 1) {                                         // test hook: the twist-by-twist walk from the resume state
        if (cx.mt_resume_tiles != cx.mt_tiles) return fail(RT_ERR_STATE, "mt_sequential: the table was built by jumps; call rt_shutdown first");
        rtd::mt_checkpoint_kernel<<<1, 640, 0, cx.stream>>>(cx.mt_table, cx.mt_resume, cx.mt_tiles, tiles - cx.mt_tiles, tiles - cx.mt_tiles, rank, world);
        CU(cudaGetLastError());
        CU(cudaStreamSynchronize(cx.stream));
        cx.mt_resume_tiles = tiles;
    } else {
        // Segments of kSegTiles tiles (512 twists); their start states by jump-ahead, top-down: at level l the states at
        // multiples of 2^(l+1) segments are known and give those at odd multiples of 2^l (polynomial l + 9: 2^l * 512 twists).
        constexpr long long kSegTiles = 64;
        constexpr int kSegLog2Twists = 9;
        static_assert(kSegTiles * mt::kTwistsPerTile == (1ll << kSegLog2Twists), "segment length");
        const long long n_seg = (tiles + kSegTiles - 1) / kSegTiles;
        if (!cx.d_jump_poly) {
            CU(cudaMalloc((void**)&cx.d_jump_poly, sizeof(mt::kJumpPoly)));
            CU(cudaMemcpy(cx.d_jump_poly, mt::kJumpPoly, sizeof(mt::kJumpPoly), cudaMemcpyHostToDevice));
        }
        int top = 0; while ((1ll << top) < n_seg) ++top;
        if (top + kSegLog2Twists > mt::kJumpPolys) return fail(RT_ERR_LIMIT, "frame too large for the jump table");
        uint32_t* seg = nullptr;
        CU(cudaMalloc((void**)&seg, (size_t)n_seg * mt::kN * sizeof(uint32_t)));
        uint32_t seed[mt::kN]; mt::seed_state(12345u, seed);                  // tracer.cpp:284
        {
            cudaError_t e0 = cudaMemcpyAsync(seg, seed, sizeof(seed), cudaMemcpyHostToDevice, cx.stream);
            if (e0 != cudaSuccess) { cudaFree(seg); return fail(RT_ERR_CUDA, std::string("jitter checkpoint table: ") + cudaGetErrorString(e0)); }
        }
        for (int l = top - 1; l >= 0; --l) {
            const long long stride = 1ll << l, pairs = (n_seg + 2 * stride - 1) / (2 * stride);
            rtd::mt_jump_kernel<<<(unsigned)pairs, 640, rtd::kJumpSmemBytes, cx.stream>>>(cx.d_jump_poly + (size_t)(l + kSegLog2Twists) * mt::kN, seg, stride, n_seg);
        }
        rtd::mt_checkpoint_kernel<<<(unsigned)n_seg, 640, 0, cx.stream>>>(cx.mt_table, seg, 0, tiles, kSegTiles, rank, world);
        cudaError_t e = cudaGetLastError();
        if (e == cudaSuccess) e = cudaStreamSynchronize(cx.stream);
        cudaFree(seg);
        if (e != cudaSuccess) return fail(RT_ERR_CUDA, std::string("jitter checkpoint table: ") + cudaGetErrorString(e));
    }
    cx.mt_tiles = tiles;
    return RT_OK;
}

template <class T> int grow(T** p, size_t* cap, size_t bytes) {
    if (*cap >= bytes) return RT_OK;
    cudaFree(*p); *p = nullptr; *cap = 0;
    CU(cudaMalloc((void**)p, bytes));
    *cap = bytes;
    return RT_OK;
}

// paper_gbuffer_kernel variant: <scene staged in shared memory, clustered runs, round-level masks>
using PaperKernel = void (*)(rtd::GBuf, int*, unsigned long long*, int, int);
// the two planes of a frame's G-buffer inside one allocation (rt_kernels.cuh)
rtd::GBuf gbuf_planes(unsigned char* base, long long npx) { rtd::GBuf g; g.nt = (float4*)base; g.ml = (int*)(base + (size_t)npx * sizeof(float4)); return g; }
PaperKernel paper_kernel_of(const rt_scene* s) {
    const bool mm = s->hdr.cull_words != 0;
    return s->clustered ? (s->fits_const ? rtd::paper_gbuffer_kernel<true, true, false> : rtd::paper_gbuffer_kernel<false, true, false>)
         : s->fits_const ? (mm ? rtd::paper_gbuffer_kernel<true, false, true> : rtd::paper_gbuffer_kernel<true, false, false>)
                         : (mm ? rtd::paper_gbuffer_kernel<false, false, true> : rtd::paper_gbuffer_kernel<false, false, false>);
}

// Launch the frame on `st` (asynchronous).  rgba8 / rgbf / ids / rgb8 are device-accessible and optional; rgb8 is the packed
// host-direct output (mapped host memory).  Counters are reset here and read back by the caller.
int launch_frame(Context& cx, rt_scene* s, int mode, int rank, int world, uchar4* rgba8, float4* rgbf, int* ids, unsigned char* rgb8,
                 cudaStream_t st, int* n_tiles_out) {
    const long long npx = (long long)s->width * s->height;
    int rc;
    if (mode == RT_MODE_STANDARD) {
        const long long tiles = (npx + mt::kPixelsPerTile - 1) / mt::kPixelsPerTile;
        rc = ensure_mt_table(cx, tiles, rank, world); if (rc) return rc;
    } else {
        rc = grow(&cx.d_gbuf, &cx.gbuf_cap, (size_t)npx * rtd::kGBufBytesPerPixel); if (rc) return rc;
    }
    rc = frame_begin(cx, st); if (rc) return rc;
    rc = bind_scene(cx, s, st); if (rc) return rc;
    CU(cudaMemsetAsync(cx.counters, 0, 4 * sizeof(unsigned long long), st));
    if (mode == RT_MODE_PAPER) {
        const unsigned blocks = (unsigned)((npx + 255) / 256);
        const rtd::GBuf gb = gbuf_planes(cx.d_gbuf, npx);
        paper_kernel_of(s)<<<blocks, 256, s->smem_bytes, st>>>(gb, ids, cx.counters, rank, world);
        rtd::paper_post_kernel<<<blocks, 256, 0, st>>>(gb, rgba8, rgbf, rgb8, s->width, s->height, rank, world);
        CU(cudaGetLastError());
        if (n_tiles_out) *n_tiles_out = (s->height + rtd::kPaperRows - 1) / rtd::kPaperRows;
        return frame_end(cx, st);
    }
    const long long tiles = (npx + mt::kPixelsPerTile - 1) / mt::kPixelsPerTile;
    const long long local = tiles > rank ? (tiles - rank + world - 1) / world : 0;
    rtd::RenderParams prm{};
    prm.mt_table = cx.mt_table; prm.rgba8 = rgba8; prm.rgbf = rgbf; prm.ids = ids; prm.rgb8 = rgb8;
    prm.counters = cx.counters; prm.tile_counter = cx.tile_counter;
    prm.n_local_tiles = (int)local; prm.rank = rank; prm.world = world; prm.n_pixels = npx;
    CU(cudaMemsetAsync(cx.tile_counter, 0, sizeof(unsigned), st));
    // kernel variant: <scene staged in shared memory, clustered runs, [flat], round-level masks>; deep ray trees -> wavefront kernel
    const bool kk = s->clustered, cc = s->fits_const, mm = s->hdr.cull_words != 0;      // (clustered scenes never have masks)
    void (*kernel)(rtd::RenderParams) = nullptr;
    if (s->bounces) {
        kernel = kk ? (cc ? rtd::render_wavefront_kernel<true, true, false> : rtd::render_wavefront_kernel<false, true, false>)
               : cc ? (mm ? rtd::render_wavefront_kernel<true, false, true> : rtd::render_wavefront_kernel<true, false, false>)
                    : (mm ? rtd::render_wavefront_kernel<false, false, true> : rtd::render_wavefront_kernel<false, false, false>);
    } else if (s->flat) {
        kernel = kk ? (cc ? rtd::render_standard_kernel<true, true, true, false> : rtd::render_standard_kernel<false, true, true, false>)
               : cc ? (mm ? rtd::render_standard_kernel<true, false, true, true> : rtd::render_standard_kernel<true, false, true, false>)
                    : (mm ? rtd::render_standard_kernel<false, false, true, true> : rtd::render_standard_kernel<false, false, true, false>);
    } else {
        kernel = kk ? (cc ? rtd::render_standard_kernel<true, true, false, false> : rtd::render_standard_kernel<false, true, false, false>)
               : cc ? (mm ? rtd::render_standard_kernel<true, false, false, true> : rtd::render_standard_kernel<true, false, false, false>)
                    : (mm ? rtd::render_standard_kernel<false, false, false, true> : rtd::render_standard_kernel<false, false, false, false>);
    }
    const size_t smem = s->smem_bytes;
    int per_sm = 0;
    const int threads = rtd::kRenderThreads;
    CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, threads, smem));
    const long long units = s->bounces ? (local + rtd::kWaveTiles - 1) / rtd::kWaveTiles : local;      // the wavefront kernel takes tiles in batches
    const long long grid = std::max<long long>(1, std::min<long long>(units, (long long)cx.sms * std::max(per_sm, 1)));
    if (s->bounces) {
        const size_t stride = (size_t)rtd::wave_stride(s->hdr.recursion);
        int rcw = grow(&cx.d_wave, &cx.wave_cap, (size_t)grid * rtd::kWaveSamples * stride * sizeof(float4)); if (rcw) return rcw;
        prm.wave_state = cx.d_wave;
    }
    if (local > 0) {
        kernel<<<(unsigned)grid, threads, smem, st>>>(prm);
        CU(cudaGetLastError());
    }
    if (n_tiles_out) *n_tiles_out = (int)local;
    return frame_end(cx, st);
}

int fill_stats(Context& cx, rt_scene* s, int mode, int n_tiles, float kernel_ms, double total_ms, rt_stats* stats) {
    if (!stats) return RT_OK;
    unsigned long long c[4];
    CU(cudaMemcpy(c, cx.counters, sizeof(c), cudaMemcpyDeviceToHost));
    stats->kernel_ms = kernel_ms; stats->total_ms = total_ms;
    stats->primary_rays = c[0]; stats->shadow_rays = c[1]; stats->secondary_rays = c[2]; stats->prim_tests = c[3];
    stats->width = s->width; stats->height = s->height; stats->samples_per_pixel = mode == RT_MODE_PAPER ? 1 : 8;
    stats->n_tiles = n_tiles;
    return RT_OK;
}

int check_args(rt_scene* s, int mode, int rank, int world) {
    if (!s) return fail(RT_ERR_ARG, "render: null scene");
    if (!s->cx || !s->alive) return fail(RT_ERR_STATE, "render: rt_shutdown was called after this scene was uploaded");
    if (mode != RT_MODE_STANDARD && mode != RT_MODE_PAPER) return fail(RT_ERR_ARG, "render: unknown mode");
    if (world < 1 || rank < 0 || rank >= world) return fail(RT_ERR_ARG, "render: need 0 <= rank < world");
    return RT_OK;
}

// The device address of a host buffer the kernels can store into directly: pinned memory mapped into the device's address
// space (rt_host_frame_alloc / _open, cudaHostAlloc, cudaHostRegister with the Mapped flag), 4-byte aligned.  nullptr for
// ordinary (pageable) memory: the frame is then staged on the device and copied.  (Current device = the context's.)
unsigned char* mapped_host_pointer(const void* p) {
    if (!p) return nullptr;
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { (void)cudaGetLastError(); return nullptr; }
    if (a.type != cudaMemoryTypeHost || !a.devicePointer || ((uintptr_t)a.devicePointer & 3u)) return nullptr;
    return (unsigned char*)a.devicePointer;
}

// ---- pinned, device-mapped host frames (rt_host_frame_*) ---------------------------------------------------------
struct HostFrame { size_t bytes; std::string shm; bool owner; bool registered; };
std::map<void*, HostFrame> g_host_frames;      // (g_mu)

}  // namespace

extern "C" {

const char* rt_version(void) { return "rt_b200 0.2 (sm_100a)"; }

int rt_init(int device) {
    std::lock_guard<std::mutex> lk(g_mu);
    int count = 0, rc = device_count(&count); if (rc) return rc;
    if (device < 0) { CU(cudaGetDevice(&device)); }
    if (g_cur && g_cur->device == device) return RT_OK;
    if (g_cur) return fail(RT_ERR_STATE, "rt_init: already initialised on another device; call rt_shutdown first");
    Context* cx = nullptr;
    rc = context_of(device, &cx); if (rc) return rc;
    g_cur = cx;
    return RT_OK;
}

int rt_init_devices(const int* devices, int n_devices) {
    if (n_devices < 1 || n_devices > kMaxDevices) return fail(RT_ERR_ARG, "rt_init_devices: need 1 <= n_devices <= 64");
    std::lock_guard<std::mutex> lk(g_mu);
    int count = 0, rc = device_count(&count); if (rc) return rc;
    for (int i = 0; i < n_devices; ++i) {
        const int dev = devices ? devices[i] : i;
        if (dev < 0 || dev >= count) return fail(RT_ERR_ARG, "rt_init_devices: device " + std::to_string(dev) + " of " + std::to_string(count) + " does not exist");
        for (int k = 0; k < i; ++k) if ((devices ? devices[k] : k) == dev) return fail(RT_ERR_ARG, "rt_init_devices: device listed twice");
        Context* cx = nullptr;
        rc = context_of(dev, &cx); if (rc) return rc;
        if (i == 0 && !g_cur) g_cur = cx;
    }
    CU(cudaSetDevice(g_cur->device));
    return RT_OK;
}

void rt_shutdown(void) {
    std::lock_guard<std::mutex> lk(g_mu);
    for (int d = 0; d < kMaxDevices; ++d) if (g_ctxs[d]) destroy_context(*g_ctxs[d]);
    g_cur = nullptr;
}

int rt_scene_upload(const rt_scene_desc* d, rt_scene** out) {
    if (!d || !out) return fail(RT_ERR_ARG, "rt_scene_upload: null argument");
    if (!current_context()) { int rc = rt_init(-1); if (rc) return rc; }
    return upload_scene(*current_context(), d, out);
}

void rt_scene_free(rt_scene* s) {
    if (!s) return;
    std::lock_guard<std::mutex> lk(s->cx->mu);
    free_scene_locked(s);
}

int rt_debug_set_option(const char* name, int value) {
    if (!name) return fail(RT_ERR_ARG, "rt_debug_set_option: null name");
    std::lock_guard<std::mutex> lk(g_mu);
    const std::string n(name);
    std::atomic<int>* slot = n == "flat" ? &g_opts.flat : n == "cull" ? &g_opts.cull : n == "wavefront" ? &g_opts.wavefront
              : n == "global_scene" ? &g_opts.global_scene : n == "mt_sequential" ? &g_opts.mt_sequential : n == "grid" ? &g_opts.grid : n == "grid_cells" ? &g_opts.grid_cells : nullptr;
    if (!slot) return fail(RT_ERR_ARG, "rt_debug_set_option: unknown option " + n);
    slot->store(value < 0 ? -1 : (slot == &g_opts.grid_cells ? value : (value ? 1 : 0)));      // grid_cells: cells per sphere
    return RT_OK;
}

// Host-only view of the device layout rt_scene_upload would build (no GPU needed): the permutation of the top-level
// objects, the runs (kind, first, count, 1 + first cluster | 0) and the cluster bounds.  Any output may be NULL.
int rt_debug_device_layout(const rt_scene_desc* d, int* order /*n_objects*/, int* n_runs, int* runs4 /*<= n_objects x 4*/,
                           int* n_clusters, float* clusters4 /*<= n_objects + n_objects/16 + 1, x 4*/) {
    if (!d) return fail(RT_ERR_ARG, "rt_debug_device_layout: null scene");
    rt_scene s;
    build_host_arrays(*d, s);
    if (order) for (int i = 0; i < d->n_objects; ++i) order[i] = s.obj_orig[i];
    if (n_runs) *n_runs = (int)s.runs.size();
    if (runs4) for (size_t i = 0; i < s.runs.size(); ++i) { runs4[4 * i] = s.runs[i].x; runs4[4 * i + 1] = s.runs[i].y; runs4[4 * i + 2] = s.runs[i].z; runs4[4 * i + 3] = s.runs[i].w; }
    if (n_clusters) *n_clusters = (int)s.clusters.size();
    if (clusters4) for (size_t i = 0; i < s.clusters.size(); ++i) { clusters4[4 * i] = s.clusters[i].x; clusters4[4 * i + 1] = s.clusters[i].y; clusters4[4 * i + 2] = s.clusters[i].z; clusters4[4 * i + 3] = s.clusters[i].w; }
    return RT_OK;
}

int rt_debug_mask_table(const rt_scene_desc* d, int* n, float* masks4, float* scale) {
    if (!d) return fail(RT_ERR_ARG, "rt_debug_mask_table: null scene");
    rt_scene s;
    build_host_arrays(*d, s);
    if (n) *n = (int)s.masks.size();
    if (scale) *scale = s.mask_scale;
    if (masks4) for (size_t i = 0; i < s.masks.size(); ++i) { masks4[4 * i] = s.masks[i].x; masks4[4 * i + 1] = s.masks[i].y; masks4[4 * i + 2] = s.masks[i].z; masks4[4 * i + 3] = s.masks[i].w; }
    return RT_OK;
}

int rt_scene_size(const rt_scene* s, int* w, int* h) {
    if (!s) return fail(RT_ERR_ARG, "rt_scene_size: null scene");
    if (w) *w = s->width;
    if (h) *h = s->height;
    return RT_OK;
}

int rt_render_device(rt_scene* s, int mode, int rank, int world, void* rgba8_device, void* rgbf_device, void* ids_device,
                     void* stream, rt_stats* stats) {
    int rc = check_args(s, mode, rank, world); if (rc) return rc;
    if (!rgba8_device) return fail(RT_ERR_ARG, "rt_render_device: rgba8_device is required");
    Context& cx = *s->cx;
    std::lock_guard<std::mutex> lk(cx.mu);
    CU(cudaSetDevice(cx.device));
    cudaStream_t st = stream ? (cudaStream_t)stream : cx.stream;
    const auto t0 = std::chrono::steady_clock::now();
    int n_tiles = 0;
    if (mode == RT_MODE_STANDARD) {         // table growth (one-off) must not sit between the events
        const long long npx = (long long)s->width * s->height;
        rc = ensure_mt_table(cx, (npx + mt::kPixelsPerTile - 1) / mt::kPixelsPerTile, rank, world); if (rc) return rc;
    }
    rc = frame_begin(cx, st); if (rc) return rc;
    rc = bind_scene(cx, s, st); if (rc) return rc;
    CU(cudaEventRecord(cx.ev0, st));
    rc = launch_frame(cx, s, mode, rank, world, (uchar4*)rgba8_device, (float4*)rgbf_device, (int*)ids_device, nullptr, st, &n_tiles);
    if (rc) return rc;
    CU(cudaEventRecord(cx.ev1, st));
    if (stats) {
        CU(cudaEventSynchronize(cx.ev1));
        float ms = 0.f; CU(cudaEventElapsedTime(&ms, cx.ev0, cx.ev1));
        const double total = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
        return fill_stats(cx, s, mode, n_tiles, ms, total, stats);
    }
    return RT_OK;
}

}  // extern "C"

// ---- Paper frame without any allocation (rt_render_paper_device_noalloc) ----------------------------------------
// Scratch of the entry point below: 256 bytes of device counters (cleared by every call), then the G-buffer of
// paper_gbuffer_kernel / paper_post_kernel for at most RT_NOALLOC_MAX_PIXELS pixels.  The array lives in the module's
// global segment of every device that loads the library (25 MiB, until the process exits).  In host code the name below
// is only a stand-in for the symbol: the device address comes from cudaGetSymbolAddress, once per device.
static constexpr size_t kNoAllocHead = 256;
static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoAllocHead + (size_t)RT_NOALLOC_MAX_PIXELS * rtd::kGBufBytesPerPixel];
static unsigned char* g_noalloc_ptr[kMaxDevices] = {};      // device address of g_noalloc_scratch, per device ordinal

namespace {
// Set-up, once per device, with that device current (the caller holds the context's mutex and has called cudaSetDevice).
int noalloc_scratch(int device, unsigned char** out) {
    if (device < 0 || device >= kMaxDevices) return fail(RT_ERR_LIMIT, "rt_render_paper_device_noalloc: device ordinal beyond 63");
    if (!g_noalloc_ptr[device]) {
        void* p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if (e != cudaSuccess || !p) return fail(RT_ERR_CUDA, std::string("cudaGetSymbolAddress(g_noalloc_scratch): ") + cudaGetErrorString(e));
        g_noalloc_ptr[device] = (unsigned char*)p;
    }
    *out = g_noalloc_ptr[device];
    return RT_OK;
}
}  // namespace

extern "C" {

int rt_render_paper_device_noalloc(rt_scene* s, int rank, int world, void* rgba8_device, void* stream) {
    int rc = check_args(s, RT_MODE_PAPER, rank, world); if (rc) return rc;
    if (!rgba8_device) return fail(RT_ERR_ARG, "rt_render_paper_device_noalloc: rgba8_device is required");
    const long long npx = (long long)s->width * s->height;
    if (npx > (long long)RT_NOALLOC_MAX_PIXELS)
        return fail(RT_ERR_LIMIT, "rt_render_paper_device_noalloc: " + std::to_string(s->width) + "x" + std::to_string(s->height) + " = " + std::to_string(npx) +
                                  " pixels exceed the static scratch of RT_NOALLOC_MAX_PIXELS = " + std::to_string((long long)RT_NOALLOC_MAX_PIXELS) +
                                  " pixels; use rt_render_device");
    Context& cx = *s->cx;
    std::lock_guard<std::mutex> lk(cx.mu);
    CU(cudaSetDevice(cx.device));
    unsigned char* scratch = nullptr;
    rc = noalloc_scratch(cx.device, &scratch); if (rc) return rc;
    auto* counters = (unsigned long long*)scratch;
    const rtd::GBuf gbuf = gbuf_planes(scratch + kNoAllocHead, npx);
    cudaStream_t st = stream ? (cudaStream_t)stream : cx.stream;
    rc = frame_begin(cx, st); if (rc) return rc;
    rc = bind_scene(cx, s, st); if (rc) return rc;
    // the counters must start at zero; the G-buffer need not: pass 1 writes every entry pass 2 reads (own rows + halo)
    CU(cudaMemsetAsync(counters, 0, kNoAllocHead, st));
    const unsigned blocks = (unsigned)((npx + 255) / 256);
    paper_kernel_of(s)<<<blocks, 256, s->smem_bytes, st>>>(gbuf, nullptr, counters, rank, world);
    rtd::paper_post_kernel<<<blocks, 256, 0, st>>>(gbuf, (uchar4*)rgba8_device, nullptr, nullptr, s->width, s->height, rank, world);
    CU(cudaGetLastError());
    return frame_end(cx, st);
}

int rt_peer_frame_create(size_t bytes, void** device_ptr, unsigned char handle[RT_IPC_HANDLE_BYTES]) {
    static_assert(sizeof(cudaIpcMemHandle_t) == RT_IPC_HANDLE_BYTES, "IPC handle size");
    if (!device_ptr || !handle || bytes == 0) return fail(RT_ERR_ARG, "rt_peer_frame_create: null argument or zero size");
    Context* cx = current_context();
    if (!cx) return fail(RT_ERR_STATE, "rt_peer_frame_create: rt_init has not been called");
    std::lock_guard<std::mutex> lk(cx->mu);
    CU(cudaSetDevice(cx->device));
    void* p = nullptr;
    CU(cudaMalloc(&p, bytes));
    cudaError_t e = cudaMemset(p, 0, bytes);
    cudaIpcMemHandle_t h;
    if (e == cudaSuccess) e = cudaIpcGetMemHandle(&h, p);
    if (e != cudaSuccess) { cudaFree(p); return fail(RT_ERR_CUDA, std::string("rt_peer_frame_create: ") + cudaGetErrorString(e)); }
    std::memcpy(handle, &h, sizeof h);
    *device_ptr = p;
    return RT_OK;
}

int rt_peer_frame_open(const unsigned char handle[RT_IPC_HANDLE_BYTES], void** device_ptr) {
    if (!device_ptr || !handle) return fail(RT_ERR_ARG, "rt_peer_frame_open: null argument");
    Context* cx = current_context();
    if (!cx) return fail(RT_ERR_STATE, "rt_peer_frame_open: rt_init has not been called");
    std::lock_guard<std::mutex> lk(cx->mu);
    CU(cudaSetDevice(cx->device));
    cudaIpcMemHandle_t h;
    std::memcpy(&h, handle, sizeof h);
    void* p = nullptr;
    cudaError_t e = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) { (void)cudaGetLastError(); return fail(RT_ERR_CUDA, std::string("cudaIpcOpenMemHandle: ") + cudaGetErrorString(e)); }
    *device_ptr = p;
    return RT_OK;
}

int rt_peer_frame_close(void* device_ptr, int is_owner) {
    if (!device_ptr) return RT_OK;
    Context* cx = current_context();
    if (!cx) return fail(RT_ERR_STATE, "rt_peer_frame_close: rt_init has not been called");
    std::lock_guard<std::mutex> lk(cx->mu);
    CU(cudaSetDevice(cx->device));
    CU(cudaDeviceSynchronize());
    if (is_owner) CU(cudaFree(device_ptr)); else CU(cudaIpcCloseMemHandle(device_ptr));
    return RT_OK;
}

// ---- host frames ---------------------------------------------------------------------------------------------------
int rt_host_frame_alloc(size_t bytes, const char* shm_name, void** host_ptr) {
    if (!host_ptr || bytes == 0) return fail(RT_ERR_ARG, "rt_host_frame_alloc: null argument or zero size");
    *host_ptr = nullptr;
    if (!current_context()) { int rc = rt_init(-1); if (rc) return rc; }
    std::lock_guard<std::mutex> lk(g_mu);
    CU(cudaSetDevice(g_cur->device));
    void* p = nullptr;
    HostFrame f{bytes, shm_name ? shm_name : "", true, false};
    if (!shm_name) {
        CU(cudaHostAlloc(&p, bytes, cudaHostAllocPortable | cudaHostAllocMapped));
    } else {
        const int fd = shm_open(shm_name, O_CREAT | O_EXCL | O_RDWR, 0600);
        if (fd < 0) return fail(RT_ERR_IO, std::string("rt_host_frame_alloc: shm_open(") + shm_name + "): " + std::strerror(errno));
        if (ftruncate(fd, (off_t)bytes) != 0) { const int e = errno; close(fd); shm_unlink(shm_name); return fail(RT_ERR_IO, std::string("rt_host_frame_alloc: ftruncate: ") + std::strerror(e)); }
        p = mmap(nullptr, bytes, PROT_READ | PROT_WRITE, MAP_SHARED | MAP_POPULATE, fd, 0);
        close(fd);
        if (p == MAP_FAILED) { const int e = errno; shm_unlink(shm_name); return fail(RT_ERR_IO, std::string("rt_host_frame_alloc: mmap: ") + std::strerror(e)); }
        cudaError_t e = cudaHostRegister(p, bytes, cudaHostRegisterPortable | cudaHostRegisterMapped);
        if (e != cudaSuccess) { (void)cudaGetLastError(); munmap(p, bytes); shm_unlink(shm_name); return fail(RT_ERR_CUDA, std::string("rt_host_frame_alloc: cudaHostRegister: ") + cudaGetErrorString(e)); }
        f.registered = true;
    }
    g_host_frames[p] = f;
    *host_ptr = p;
    return RT_OK;
}

int rt_host_frame_open(const char* shm_name, size_t bytes, void** host_ptr) {
    if (!host_ptr || !shm_name || bytes == 0) return fail(RT_ERR_ARG, "rt_host_frame_open: null argument or zero size");
    *host_ptr = nullptr;
    if (!current_context()) { int rc = rt_init(-1); if (rc) return rc; }
    std::lock_guard<std::mutex> lk(g_mu);
    CU(cudaSetDevice(g_cur->device));
    const int fd = shm_open(shm_name, O_RDWR, 0600);
    if (fd < 0) return fail(RT_ERR_IO, std::string("rt_host_frame_open: shm_open(") + shm_name + "): " + std::strerror(errno));
    void* p = mmap(nullptr, bytes, PROT_READ | PROT_WRITE, MAP_SHARED | MAP_POPULATE, fd, 0);
    close(fd);
    if (p == MAP_FAILED) return fail(RT_ERR_IO, std::string("rt_host_frame_open: mmap: ") + std::strerror(errno));
    cudaError_t e = cudaHostRegister(p, bytes, cudaHostRegisterPortable | cudaHostRegisterMapped);
    if (e != cudaSuccess) { (void)cudaGetLastError(); munmap(p, bytes); return fail(RT_ERR_CUDA, std::string("rt_host_frame_open: cudaHostRegister: ") + cudaGetErrorString(e)); }
    g_host_frames[p] = HostFrame{bytes, shm_name, false, true};
    *host_ptr = p;
    return RT_OK;
}

int rt_host_frame_free(void* host_ptr) {
    if (!host_ptr) return RT_OK;
    std::lock_guard<std::mutex> lk(g_mu);
    auto it = g_host_frames.find(host_ptr);
    if (it == g_host_frames.end()) return fail(RT_ERR_ARG, "rt_host_frame_free: not a host frame of this library");
    const HostFrame f = it->second;
    g_host_frames.erase(it);
    for (int d = 0; d < kMaxDevices; ++d) if (g_ctxs[d] && g_ctxs[d]->device >= 0) { cudaSetDevice(d); cudaDeviceSynchronize(); }     // no kernel may still store into it
    if (g_cur) cudaSetDevice(g_cur->device);
    if (!f.registered) { CU(cudaFreeHost(host_ptr)); return RT_OK; }
    cudaError_t e = cudaHostUnregister(host_ptr);
    munmap(host_ptr, f.bytes);
    if (f.owner) shm_unlink(f.shm.c_str());
    if (e != cudaSuccess) return fail(RT_ERR_CUDA, std::string("rt_host_frame_free: cudaHostUnregister: ") + cudaGetErrorString(e));
    return RT_OK;
}

int rt_render(rt_scene* s, int mode, int rank, int world, uint8_t* rgb8, float* rgbf, int32_t* ids, rt_stats* stats) {
    int rc = check_args(s, mode, rank, world); if (rc) return rc;
    Context& cx = *s->cx;
    std::lock_guard<std::mutex> lk(cx.mu);
    CU(cudaSetDevice(cx.device));
    const auto t0 = std::chrono::steady_clock::now();
    cudaStream_t st = cx.stream;
    const long long npx = (long long)s->width * s->height;
    const size_t id_bytes = (size_t)npx * (mode == RT_MODE_PAPER ? 1 : 8) * sizeof(int);
    // Host-direct output: when rgb8 is pinned, device-mapped host memory (rt_host_frame_*, cudaHostAlloc, ...) the kernel
    // stores the packed pixels straight into it — no device frame, no pack pass, no D2H copy afterwards; with world > 1
    // the pixels of other ranks are left untouched (several devices / processes share one frame).
    unsigned char* direct = mapped_host_pointer(rgb8);
    const bool staged = rgb8 && !direct;
    if (staged) { rc = grow(&cx.d_rgba8, &cx.rgba8_cap, (size_t)npx * sizeof(uchar4)); if (rc) return rc; }
    if (rgbf) { rc = grow(&cx.d_rgbf, &cx.rgbf_cap, (size_t)npx * sizeof(float4)); if (rc) return rc; }
    if (ids) { rc = grow(&cx.d_ids, &cx.ids_cap, id_bytes); if (rc) return rc; }
    if (staged || rgbf) { rc = grow(&cx.d_pack, &cx.pack_cap, (size_t)npx * (rgbf ? 12 : 3)); if (rc) return rc; }
    if (mode == RT_MODE_STANDARD) { rc = ensure_mt_table(cx, (npx + mt::kPixelsPerTile - 1) / mt::kPixelsPerTile, rank, world); if (rc) return rc; }
    rc = frame_begin(cx, st); if (rc) return rc;
    rc = bind_scene(cx, s, st); if (rc) return rc;
    if (world > 1) {   // staged outputs: pixels of other ranks are defined (zero / -1) so that slices can be summed or compared
        if (staged) CU(cudaMemsetAsync(cx.d_rgba8, 0, (size_t)npx * sizeof(uchar4), st));
        if (rgbf) CU(cudaMemsetAsync(cx.d_rgbf, 0, (size_t)npx * sizeof(float4), st));
        if (ids) CU(cudaMemsetAsync(cx.d_ids, 0xff, id_bytes, st));
    }
    int n_tiles = 0;
    CU(cudaEventRecord(cx.ev0, st));
    rc = launch_frame(cx, s, mode, rank, world, staged ? cx.d_rgba8 : nullptr, rgbf ? cx.d_rgbf : nullptr, ids ? cx.d_ids : nullptr, direct, st, &n_tiles);
    if (rc) return rc;
    CU(cudaEventRecord(cx.ev1, st));
    const unsigned blocks = (unsigned)((npx + 255) / 256);
    if (staged) {
        rtd::pack_rgb_kernel<<<blocks, 256, 0, st>>>(cx.d_rgba8, cx.d_pack, npx);
        CU(cudaMemcpyAsync(rgb8, cx.d_pack, (size_t)npx * 3, cudaMemcpyDeviceToHost, st));
        if (rgbf) CU(cudaStreamSynchronize(st));                 // d_pack is reused below
    }
    if (rgbf) {
        rtd::unpack_rgbf_kernel<<<blocks, 256, 0, st>>>(cx.d_rgbf, (float*)cx.d_pack, npx);
        CU(cudaMemcpyAsync(rgbf, cx.d_pack, (size_t)npx * 12, cudaMemcpyDeviceToHost, st));
    }
    if (ids) CU(cudaMemcpyAsync(ids, cx.d_ids, id_bytes, cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    CU(cudaGetLastError());
    cx.in_flight = false;
    float ms = 0.f; CU(cudaEventElapsedTime(&ms, cx.ev0, cx.ev1));
    const double total = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    return fill_stats(cx, s, mode, n_tiles, ms, total, stats);
}

// Tracer::render on several GPUs of one box from ONE process (tracer.cpp:247-305; `RAY_GPUS=8 ray scene.json out.png`).
// Device i renders the tiles i, i + n, ... of the frame (a replica of the scene each: KBs) and stores its pixels as packed
// RGB8 straight into the caller's frame when that is a host frame (rt_host_frame_alloc), else into a pinned frame of the
// library that is copied out at the end.  No peer access, no collective, no device-side exchange.
int rt_render_multi(const rt_scene_desc* desc, int mode, const int* devices, int n_devices, uint8_t* rgb8, rt_stats* stats) {
    if (!desc || !rgb8) return fail(RT_ERR_ARG, "rt_render_multi: null argument");
    if (mode != RT_MODE_STANDARD && mode != RT_MODE_PAPER) return fail(RT_ERR_ARG, "rt_render_multi: unknown mode");
    const auto t0 = std::chrono::steady_clock::now();
    int rc = rt_init_devices(devices, n_devices); if (rc) return rc;
    std::vector<Context*> cxs((size_t)n_devices);
    { std::lock_guard<std::mutex> lk(g_mu); for (int i = 0; i < n_devices; ++i) cxs[(size_t)i] = g_ctxs[devices ? devices[i] : i]; }
    const size_t bytes = (size_t)desc->width * desc->height * 3;
    std::vector<rt_scene*> scenes((size_t)n_devices, nullptr);
    auto cleanup = [&]() { for (rt_scene* s : scenes) if (s) rt_scene_free(s); };
    for (int i = 0; i < n_devices; ++i)
        if ((rc = upload_scene(*cxs[(size_t)i], desc, &scenes[(size_t)i]))) { cleanup(); return rc; }
    // the frame every device stores into
    void* own_frame = nullptr;
    unsigned char* frame = rgb8;
    { CU(cudaSetDevice(cxs[0]->device)); }
    if (!mapped_host_pointer(rgb8)) {
        rc = rt_host_frame_alloc(bytes, nullptr, &own_frame);
        if (rc) { cleanup(); return rc; }
        frame = (unsigned char*)own_frame;
    }
    std::vector<int> n_tiles((size_t)n_devices, 0);
    for (int i = 0; i < n_devices && !rc; ++i) {
        Context& cx = *cxs[(size_t)i];
        std::lock_guard<std::mutex> lk(cx.mu);
        cudaError_t e = cudaSetDevice(cx.device);
        unsigned char* direct = e == cudaSuccess ? mapped_host_pointer(frame) : nullptr;
        if (!direct) { rc = fail(RT_ERR_CUDA, "rt_render_multi: the host frame is not mapped on device " + std::to_string(cx.device)); break; }
        if (mode == RT_MODE_STANDARD) {
            const long long npx = (long long)desc->width * desc->height;
            rc = ensure_mt_table(cx, (npx + mt::kPixelsPerTile - 1) / mt::kPixelsPerTile, i, n_devices); if (rc) break;
        }
        if ((rc = frame_begin(cx, cx.stream)) || (rc = bind_scene(cx, scenes[(size_t)i], cx.stream))) break;
        if (cudaEventRecord(cx.ev0, cx.stream) != cudaSuccess) { rc = fail(RT_ERR_CUDA, "rt_render_multi: cudaEventRecord"); break; }
        rc = launch_frame(cx, scenes[(size_t)i], mode, i, n_devices, nullptr, nullptr, nullptr, direct, cx.stream, &n_tiles[(size_t)i]);
        if (!rc && cudaEventRecord(cx.ev1, cx.stream) != cudaSuccess) rc = fail(RT_ERR_CUDA, "rt_render_multi: cudaEventRecord");
    }
    // wait for every device (also after a failure: nothing may still be storing into the frame)
    rt_stats total{}; float kernel_ms = 0.f;
    for (int i = 0; i < n_devices; ++i) {
        Context& cx = *cxs[(size_t)i];
        std::lock_guard<std::mutex> lk(cx.mu);
        cudaSetDevice(cx.device);
        cudaError_t e = cudaStreamSynchronize(cx.stream);
        cx.in_flight = false;
        if (e != cudaSuccess && !rc) rc = fail(RT_ERR_CUDA, std::string("rt_render_multi: device ") + std::to_string(cx.device) + ": " + cudaGetErrorString(e));
        if (rc || !stats) continue;
        float ms = 0.f; rt_stats one{};
        if (cudaEventElapsedTime(&ms, cx.ev0, cx.ev1) == cudaSuccess) kernel_ms = std::max(kernel_ms, ms);
        if (fill_stats(cx, scenes[(size_t)i], mode, n_tiles[(size_t)i], ms, 0.0, &one) == RT_OK) {
            total.primary_rays += one.primary_rays; total.shadow_rays += one.shadow_rays; total.secondary_rays += one.secondary_rays;
            total.prim_tests += one.prim_tests; total.n_tiles += one.n_tiles;
            total.width = one.width; total.height = one.height; total.samples_per_pixel = one.samples_per_pixel;
        }
    }
    cudaSetDevice(cxs[0]->device);
    if (!rc && own_frame) std::memcpy(rgb8, own_frame, bytes);
    if (own_frame) rt_host_frame_free(own_frame);
    cleanup();
    if (rc) return rc;
    if (stats) {
        total.kernel_ms = kernel_ms;
        total.total_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
        *stats = total;
    }
    return RT_OK;
}

int rt_trace_rays(rt_scene* s, int n, const double* rays, double* out) {
    if (!s || !rays || !out || n < 0) return fail(RT_ERR_ARG, "rt_trace_rays: bad argument");
    if (n == 0) return RT_OK;
    Context& cx = *s->cx;
    std::lock_guard<std::mutex> lk(cx.mu);
    CU(cudaSetDevice(cx.device));
    cudaStream_t st = cx.stream;
    int rc = frame_begin(cx, st); if (rc) return rc;
    rc = bind_scene(cx, s, st); if (rc) return rc;
    double *d_in = nullptr, *d_out = nullptr;
    CU(cudaMalloc((void**)&d_in, (size_t)n * 22 * sizeof(double)));          // one allocation: rays in, records out
    d_out = d_in + (size_t)n * 6;
    {
        cudaError_t e0 = cudaMemcpyAsync(d_in, rays, (size_t)n * 6 * sizeof(double), cudaMemcpyHostToDevice, st);
        if (e0 != cudaSuccess) { cudaFree(d_in); return fail(RT_ERR_CUDA, std::string("rt_trace_rays: ") + cudaGetErrorString(e0)); }
    }
    if (s->fits_const) rtd::trace_rays_kernel<true><<<(n + 127) / 128, 128, s->smem_bytes, st>>>(d_in, d_out, n);
    else rtd::trace_rays_kernel<false><<<(n + 127) / 128, 128, s->smem_bytes, st>>>(d_in, d_out, n);
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaMemcpyAsync(out, d_out, (size_t)n * 16 * sizeof(double), cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    cudaFree(d_in);
    if (e != cudaSuccess) return fail(RT_ERR_CUDA, std::string("rt_trace_rays: ") + cudaGetErrorString(e));
    return RT_OK;
}

int rt_debug_jitter(uint64_t first_sample, int n, double* out) {
    if (!out || n < 0) return fail(RT_ERR_ARG, "rt_debug_jitter: bad argument");
    if (!current_context()) { int rc = rt_init(-1); if (rc) return rc; }
    if (n == 0) return RT_OK;
    Context& cx = *current_context();
    std::lock_guard<std::mutex> lk(cx.mu);
    CU(cudaSetDevice(cx.device));
    const long long tiles = (long long)((first_sample + (uint64_t)n + mt::kSamplesPerTile - 1) / mt::kSamplesPerTile);
    int rc = ensure_mt_table(cx, tiles, 0, 1); if (rc) return rc;
    double* d_out = nullptr;
    CU(cudaMalloc((void**)&d_out, (size_t)n * 2 * sizeof(double)));
    rtd::jitter_debug_kernel<<<1, 256, 0, cx.stream>>>(cx.mt_table, first_sample, n, d_out);
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaMemcpyAsync(out, d_out, (size_t)n * 2 * sizeof(double), cudaMemcpyDeviceToHost, cx.stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(cx.stream);
    cudaFree(d_out);
    if (e != cudaSuccess) return fail(RT_ERR_CUDA, std::string("rt_debug_jitter: ") + cudaGetErrorString(e));
    return RT_OK;
}

int rt_measure_fp32_peak(double* tflops) {
    if (!tflops) return fail(RT_ERR_ARG, "rt_measure_fp32_peak: null argument");
    if (!current_context()) { int rc = rt_init(-1); if (rc) return rc; }
    Context& cx = *current_context();
    std::lock_guard<std::mutex> lk(cx.mu);
    CU(cudaSetDevice(cx.device));
    float* sink = nullptr; CU(cudaMalloc((void**)&sink, sizeof(float)));
    const int iters = 4096, blocks = cx.sms * 8, threads = 256;
    double best = 0.0;
    for (int rep = 0; rep < 5; ++rep) {
        CU(cudaEventRecord(cx.ev0, cx.stream));
        rtd::ffma_peak_kernel<<<blocks, threads, 0, cx.stream>>>(sink, iters, 1.0000001f, 1e-9f);
        CU(cudaEventRecord(cx.ev1, cx.stream));
        CU(cudaEventSynchronize(cx.ev1));
        float ms = 0.f; CU(cudaEventElapsedTime(&ms, cx.ev0, cx.ev1));
        const double flops = 2.0 * 8 * 16 * (double)iters * blocks * threads;
        if (rep > 0) best = std::max(best, flops / (ms * 1e-3) * 1e-12);
    }
    cudaFree(sink);
    *tflops = best;
    return RT_OK;
}

}  // extern "C"
