// rt_api.cu — C ABI (include/rt_b200.h) over the sm_100a kernels.  One process renders on one
// GPU; multi-GPU jobs run one process per GPU and gather the slices themselves (NCCL).
// There is no CPU fallback: every compute entry point fails with RT_ERR_CUDA without a device.
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <mutex>
#include <string>
#include <vector>

#include "rt_kernels.cuh"
#include "rt_internal.hpp"

namespace {

int fail(int code, const std::string& msg) { rt_set_error(msg); return code; }
#define CU(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) \
    return fail(RT_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(e_)); } while (0)

struct Context {
    bool ready{false};
    int device{-1}, sms{0};
    cudaStream_t stream{nullptr};
    cudaEvent_t ev0{nullptr}, ev1{nullptr};
    // jitter-stream checkpoints (scene independent)
    uint32_t* mt_table{nullptr}; long long mt_tiles{0}, mt_capacity{0};
    uint32_t* mt_resume{nullptr}; long long mt_resume_tiles{0};    // sequential walk (RT_MT_SEQUENTIAL): state after mt_resume_tiles tiles
    uint32_t* d_jump_poly{nullptr};                                // mt::kJumpPoly on the device
    unsigned long long* counters{nullptr}; unsigned* tile_counter{nullptr};
    const void* const_owner{nullptr};     // which scene currently occupies constant memory
    // frame buffers of the host-output path (rt_render): owned by the context and grown on demand, so that a
    // fresh scene (the CLI, the bench's end-to-end leg) does not pay cudaMalloc/cudaFree of ~100 MB per frame
    uchar4* d_rgba8{nullptr}; float4* d_rgbf{nullptr}; int* d_ids{nullptr}; rtd::GBuf* d_gbuf{nullptr}; unsigned char* d_pack{nullptr};
    size_t rgba8_cap{0}, rgbf_cap{0}, ids_cap{0}, gbuf_cap{0}, pack_cap{0};     // bytes
    float4* d_wave{nullptr}; size_t wave_cap{0};                                  // wavefront kernel: per-CTA sample state
    std::mutex mu;
};
Context g_ctx;
// Test hooks (rt_debug_set_option): -1 = the library's own choice, 0 / 1 = forced.  Never read from the environment.
struct DebugOptions { int flat{-1}, cull{-1}, wavefront{-1}, global_scene{-1}, mt_sequential{-1}; };
DebugOptions g_opts;
constexpr size_t kSceneSmemBudget = 80 * 1024;   // + 25 KB of MT19937 buffers: two CTAs per SM still fit in 227 KB

template <class T> int upload(const std::vector<T>& h, T** d) {
    *d = nullptr;
    size_t bytes = std::max<size_t>(h.size(), 1) * sizeof(T);
    CU(cudaMalloc((void**)d, bytes));
    if (!h.empty()) CU(cudaMemcpy(*d, h.data(), h.size() * sizeof(T), cudaMemcpyHostToDevice));
    return RT_OK;
}

}  // namespace

struct rt_scene {
    rtd::SceneHeader hdr{};
    bool fits_const{false};   // scanned arrays fit the shared-memory budget (staged per CTA)
    bool flat{false};         // no ray can spawn a child (no reflective / refractive material, or recursion == 1)
    bool bounces{false};      // most materials reflect or refract and medium.recursion >= 2: deep, uneven ray trees
    bool clustered{false};    // a long run of plain spheres / pokeballs was reordered into spatial clusters (kernel variant K)
    size_t smem_bytes{0};
    std::vector<float4> obj_geom, leaf_geom, pokeballs, affines, materials, lights, fallbacks, clusters, masks, cull_geom;
    std::vector<uint32_t> cull_plane_bits;
    int cull_words{0}, cull_leaf_word{0};
    uint32_t* d_cull_plane_bits{nullptr};
    float mask_scale{0.f};
    std::vector<int4> obj_meta, leaf_meta, runs;
    std::vector<int> obj_orig;
    std::vector<uint32_t> ops;
    float4 *d_obj_geom{nullptr}, *d_leaf_geom{nullptr}, *d_pokeballs{nullptr}, *d_affines{nullptr}, *d_materials{nullptr},
           *d_lights{nullptr}, *d_fallbacks{nullptr}, *d_clusters{nullptr}, *d_masks{nullptr}, *d_cull_geom{nullptr};
    int* d_obj_orig{nullptr};
    int4 *d_obj_meta{nullptr}, *d_leaf_meta{nullptr}, *d_runs{nullptr};
    uint32_t* d_ops{nullptr};
    int width{0}, height{0};
};

namespace {

void free_scene_locked(rt_scene* s);

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

void build_host_arrays(const rt_scene_desc& d, rt_scene& s) {
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
    s.cull_plane_bits.assign((size_t)std::max(s.cull_words, 1), 0u);
    for (size_t i : planes) s.cull_plane_bits[i >> 5] |= 1u << (i & 31);
}

}  // namespace

extern "C" {

const char* rt_version(void) { return "rt_b200 0.1 (sm_100a)"; }

int rt_init(int device) {
    std::lock_guard<std::mutex> lk(g_ctx.mu);
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0)
        return fail(RT_ERR_CUDA, std::string("no CUDA device: ") + (e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0") +
                                     " (this library has no CPU fallback)");
    if (device < 0) { CU(cudaGetDevice(&device)); }
    if (g_ctx.ready && g_ctx.device == device) return RT_OK;
    if (g_ctx.ready) return fail(RT_ERR_STATE, "rt_init: already initialised on another device; call rt_shutdown first");
    CU(cudaSetDevice(device));
    cudaDeviceProp prop; CU(cudaGetDeviceProperties(&prop, device));
    g_ctx.device = device; g_ctx.sms = prop.multiProcessorCount;
    CU(cudaStreamCreateWithFlags(&g_ctx.stream, cudaStreamNonBlocking));
    CU(cudaEventCreate(&g_ctx.ev0)); CU(cudaEventCreate(&g_ctx.ev1));
    CU(cudaMalloc((void**)&g_ctx.counters, 4 * sizeof(unsigned long long)));
    CU(cudaMalloc((void**)&g_ctx.tile_counter, sizeof(unsigned)));
    CU(cudaMalloc((void**)&g_ctx.mt_resume, mt::kN * sizeof(uint32_t)));
    uint32_t seed[mt::kN]; mt::seed_state(12345u, seed);                  // tracer.cpp:284
    CU(cudaMemcpy(g_ctx.mt_resume, seed, sizeof(seed), cudaMemcpyHostToDevice));
    g_ctx.mt_tiles = 0; g_ctx.mt_capacity = 0; g_ctx.mt_table = nullptr; g_ctx.const_owner = nullptr;
    g_ctx.mt_resume_tiles = 0; g_ctx.d_jump_poly = nullptr;
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
    g_ctx.ready = true;
    return RT_OK;
}

void rt_shutdown(void) {
    std::lock_guard<std::mutex> lk(g_ctx.mu);
    if (!g_ctx.ready) return;
    cudaSetDevice(g_ctx.device);
    cudaFree(g_ctx.mt_table); cudaFree(g_ctx.mt_resume); cudaFree(g_ctx.counters); cudaFree(g_ctx.tile_counter);
    cudaFree(g_ctx.d_jump_poly); g_ctx.d_jump_poly = nullptr; g_ctx.mt_resume_tiles = 0;
    cudaFree(g_ctx.d_rgba8); cudaFree(g_ctx.d_rgbf); cudaFree(g_ctx.d_ids); cudaFree(g_ctx.d_gbuf); cudaFree(g_ctx.d_pack);
    g_ctx.d_rgba8 = nullptr; g_ctx.d_rgbf = nullptr; g_ctx.d_ids = nullptr; g_ctx.d_gbuf = nullptr; g_ctx.d_pack = nullptr;
    g_ctx.rgba8_cap = g_ctx.rgbf_cap = g_ctx.ids_cap = g_ctx.gbuf_cap = g_ctx.pack_cap = 0;
    cudaFree(g_ctx.d_wave); g_ctx.d_wave = nullptr; g_ctx.wave_cap = 0;
    cudaEventDestroy(g_ctx.ev0); cudaEventDestroy(g_ctx.ev1); cudaStreamDestroy(g_ctx.stream);
    g_ctx.ready = false; g_ctx.device = -1; g_ctx.mt_table = nullptr; g_ctx.mt_tiles = 0; g_ctx.mt_capacity = 0;
    g_ctx.mt_resume = nullptr; g_ctx.counters = nullptr; g_ctx.tile_counter = nullptr; g_ctx.const_owner = nullptr;
    g_ctx.stream = nullptr; g_ctx.ev0 = g_ctx.ev1 = nullptr;
}

int rt_scene_upload(const rt_scene_desc* d, rt_scene** out) {
    if (!d || !out) return fail(RT_ERR_ARG, "rt_scene_upload: null argument");
    if (!g_ctx.ready) { int rc = rt_init(-1); if (rc) return rc; }
    if (d->width <= 0 || d->height <= 0) return fail(RT_ERR_ARG, "rt_scene_upload: empty screen");
    if (d->recursion > RT_MAX_RECURSION) return fail(RT_ERR_LIMIT, "medium.recursion exceeds RT_MAX_RECURSION");
    std::lock_guard<std::mutex> lk(g_ctx.mu);
    CU(cudaSetDevice(g_ctx.device));
    auto* s = new rt_scene();
    build_host_arrays(*d, *s);
    int rc = RT_OK;
    if ((rc = upload(s->obj_geom, &s->d_obj_geom)) || (rc = upload(s->obj_meta, &s->d_obj_meta)) ||
        (rc = upload(s->leaf_geom, &s->d_leaf_geom)) || (rc = upload(s->leaf_meta, &s->d_leaf_meta)) ||
        (rc = upload(s->pokeballs, &s->d_pokeballs)) || (rc = upload(s->affines, &s->d_affines)) ||
        (rc = upload(s->ops, &s->d_ops)) || (rc = upload(s->materials, &s->d_materials)) ||
        (rc = upload(s->lights, &s->d_lights)) || (rc = upload(s->fallbacks, &s->d_fallbacks)) ||
        (rc = upload(s->runs, &s->d_runs)) || (rc = upload(s->clusters, &s->d_clusters)) ||
        (rc = upload(s->obj_orig, &s->d_obj_orig)) || (rc = upload(s->masks, &s->d_masks)) ||
        (rc = upload(s->cull_geom, &s->d_cull_geom)) || (rc = upload(s->cull_plane_bits, &s->d_cull_plane_bits))) { free_scene_locked(s); return rc; }
    rtd::SceneHeader& h = s->hdr;
    for (int k = 0; k < 3; ++k) { h.eye[k] = d->eye[k]; h.P[k] = d->screen_p[k]; h.bg[k] = d->background[k]; }
    h.lx = d->lx; h.ly = d->ly; h.nx = d->width; h.ny = d->height;
    h.medium_index = d->medium_index; h.recursion = d->recursion;
    h.n_objects = d->n_objects; h.n_lights = d->n_dir_lights + d->n_lights; h.n_runs = (int)s->runs.size();
    h.obj_geom = s->d_obj_geom; h.obj_meta = s->d_obj_meta; h.leaf_geom = s->d_leaf_geom; h.leaf_meta = s->d_leaf_meta;
    h.pokeballs = s->d_pokeballs; h.affines = s->d_affines; h.ops = s->d_ops; h.materials = s->d_materials;
    h.lights = s->d_lights; h.fallbacks = s->d_fallbacks; h.runs = s->d_runs; h.clusters = s->d_clusters; h.obj_orig = s->d_obj_orig;
    h.n_clusters = (int)s->clusters.size();
    h.masks = s->d_masks; h.mask_scale = s->mask_scale;
    s->width = d->width; s->height = d->height;
    h.n_leaves = d->n_leaves; h.n_ops = d->n_ops;
    h.off_leaf_geom = s->clustered ? 0 : d->n_objects;               // scenes with clustered runs read top-level geometry from global memory
    h.off_leaf_meta = h.off_leaf_geom + d->n_leaves; h.off_runs = h.off_leaf_meta + d->n_leaves;
    h.off_lights = h.off_runs + (int)s->runs.size();
    h.off_clusters = h.off_lights + 2 * h.n_lights;
    h.off_masks = h.off_clusters + h.n_clusters;
    h.off_ops_bytes = 16 * (h.off_masks + (int)s->masks.size());
    h.smem_bytes = h.off_ops_bytes + 4 * d->n_ops;
    s->fits_const = (size_t)h.smem_bytes <= kSceneSmemBudget;
    // Wavefront kernel only where most surfaces spawn secondary rays: its per-sample state traffic costs ~10 % on scenes
    // whose ray trees are mostly depth 1 (snorlax with one reflective sphere), and gains ~6 % on the all-glass lattice.
    int n_bounce = 0;
    for (int i = 0; i < d->n_materials; ++i) n_bounce += (d->materials[i].kr > 0.f || d->materials[i].kt > 0.f) ? 1 : 0;
    s->flat = d->recursion >= 1 && (n_bounce == 0 || d->recursion == 1);
    if (g_opts.flat == 0) s->flat = false;                                  // test hook: generic loop on a flat scene
    s->bounces = d->recursion >= 2 && d->n_materials > 0 && 2 * n_bounce >= d->n_materials;
    if (g_opts.wavefront >= 0) s->bounces = g_opts.wavefront == 1;          // test hook: either kernel on any scene
    if (g_opts.global_scene == 1) s->fits_const = false;                    // test hook: the global-memory path on any scene
    // Round-level culling: masks of (1 + kMaxListLights) bundles behind the staged scene (or alone, on the global path).
    // Scenes whose table is too long for a CTA to cull per round (> kMaxCullWords x 32 entries) scan without masks.
    // Round-level masks pay when a scan is more than a handful of tests and the table is short enough for a CTA to cull per
    // round: not for scenes of a few primitives (pokeballs: 4 objects), not for clustered scenes (their scans already walk a
    // spatial structure; culling 1 300 entries per bundle and round cost more than it saved: synthetic-A 2K 23.8 -> 41 ms),
    // not for the wavefront kernel (only its camera pass could use them: synthetic-B 1K 32.1 -> 37.3 ms).
    constexpr int kMaxCullWords = 8, kMinCullEntries = 10;
    bool cull = s->cull_words > 0 && s->cull_words <= kMaxCullWords && !s->clustered && !s->bounces && d->n_objects + d->n_leaves >= kMinCullEntries;
    if (g_opts.cull == 0) cull = false;
    if (g_opts.cull == 1) cull = s->cull_words > 0 && s->cull_words <= 96 && !s->clustered;        // test hook: masks wherever they can be built
    h.cull_geom = s->d_cull_geom; h.cull_words = cull ? s->cull_words : 0; h.cull_plane_bits = s->d_cull_plane_bits; h.cull_leaf_word = s->cull_leaf_word;
    h.off_cand_bytes = s->fits_const ? ((h.smem_bytes + 15) & ~15) : 0;
    s->smem_bytes = (size_t)h.off_cand_bytes + (cull ? sizeof(rtd::CandShared) + (size_t)(1 + rtd::kMaxListLights) * s->cull_words * 4 : 0);    *out = s;
    return RT_OK;
}

void rt_scene_free(rt_scene* s) {
    if (!s) return;
    std::lock_guard<std::mutex> lk(g_ctx.mu);
    free_scene_locked(s);
}

int rt_debug_set_option(const char* name, int value) {
    if (!name) return fail(RT_ERR_ARG, "rt_debug_set_option: null name");
    std::lock_guard<std::mutex> lk(g_ctx.mu);
    const std::string n(name);
    int* slot = n == "flat" ? &g_opts.flat : n == "cull" ? &g_opts.cull : n == "wavefront" ? &g_opts.wavefront
              : n == "global_scene" ? &g_opts.global_scene : n == "mt_sequential" ? &g_opts.mt_sequential : nullptr;
    if (!slot) return fail(RT_ERR_ARG, "rt_debug_set_option: unknown option " + n);
    *slot = value < 0 ? -1 : (value ? 1 : 0);
    return RT_OK;
}

}  // extern "C"

namespace {
void free_scene_locked(rt_scene* s) {
    if (g_ctx.const_owner == s) g_ctx.const_owner = nullptr;
    cudaFree(s->d_obj_geom); cudaFree(s->d_obj_meta); cudaFree(s->d_leaf_geom); cudaFree(s->d_leaf_meta);
    cudaFree(s->d_pokeballs); cudaFree(s->d_affines); cudaFree(s->d_ops); cudaFree(s->d_materials);
    cudaFree(s->d_lights); cudaFree(s->d_fallbacks); cudaFree(s->d_runs); cudaFree(s->d_clusters); cudaFree(s->d_obj_orig); cudaFree(s->d_masks); cudaFree(s->d_cull_geom); cudaFree(s->d_cull_plane_bits);
    delete s;
}
}  // namespace

extern "C" {

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

}  // extern "C"

namespace {

// Make `s` the scene held in constant memory (header always; arrays when they fit).
int bind_scene(rt_scene* s, cudaStream_t st) {
    if (g_ctx.const_owner == s) return RT_OK;
    CU(cudaMemcpyToSymbolAsync(rtd::c_hdr, &s->hdr, sizeof(s->hdr), 0, cudaMemcpyHostToDevice, st));
    CU(cudaStreamSynchronize(st));      // host vectors are pageable: be done before returning
    g_ctx.const_owner = s;
    return RT_OK;
}

// Extend the jitter checkpoint table to cover `tiles` tiles (scene/resolution independent).
int ensure_mt_table(long long tiles) {
    if (tiles <= g_ctx.mt_tiles) return RT_OK;
    if (tiles > g_ctx.mt_capacity) {
        long long cap = std::max(tiles, g_ctx.mt_capacity * 2);
        uint32_t* nt = nullptr;
        CU(cudaMalloc((void**)&nt, (size_t)cap * mt::kN * sizeof(uint32_t)));
        if (g_ctx.mt_tiles) CU(cudaMemcpy(nt, g_ctx.mt_table, (size_t)g_ctx.mt_tiles * mt::kN * sizeof(uint32_t), cudaMemcpyDeviceToDevice));
        cudaFree(g_ctx.mt_table);
        g_ctx.mt_table = nt; g_ctx.mt_capacity = cap;
    }
    if (g_opts.mt_sequential == 1) {                               // test hook: the twist-by-twist walk from the resume state
        if (g_ctx.mt_resume_tiles != g_ctx.mt_tiles) return fail(RT_ERR_STATE, "mt_sequential: the table was built by jumps; call rt_shutdown first");
        rtd::mt_checkpoint_kernel<<<1, 640, 0, g_ctx.stream>>>(g_ctx.mt_table, g_ctx.mt_resume, g_ctx.mt_tiles, tiles - g_ctx.mt_tiles, tiles - g_ctx.mt_tiles);
        CU(cudaGetLastError());
        CU(cudaStreamSynchronize(g_ctx.stream));
        g_ctx.mt_resume_tiles = tiles;
    } else {
        // Segments of kSegTiles tiles (512 twists); their start states by jump-ahead, top-down: at level l the states at
        // multiples of 2^(l+1) segments are known and give those at odd multiples of 2^l (polynomial l + 9: 2^l * 512 twists).
        constexpr long long kSegTiles = 64;
        constexpr int kSegLog2Twists = 9;
        static_assert(kSegTiles * mt::kTwistsPerTile == (1ll << kSegLog2Twists), "segment length");
        const long long n_seg = (tiles + kSegTiles - 1) / kSegTiles;
        if (!g_ctx.d_jump_poly) {
            CU(cudaMalloc((void**)&g_ctx.d_jump_poly, sizeof(mt::kJumpPoly)));
            CU(cudaMemcpy(g_ctx.d_jump_poly, mt::kJumpPoly, sizeof(mt::kJumpPoly), cudaMemcpyHostToDevice));
            CU(cudaFuncSetAttribute(rtd::mt_jump_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, rtd::kJumpSmemBytes));
        }
        int top = 0; while ((1ll << top) < n_seg) ++top;
        if (top + kSegLog2Twists > mt::kJumpPolys) return fail(RT_ERR_LIMIT, "frame too large for the jump table");
        uint32_t* seg = nullptr;
        CU(cudaMalloc((void**)&seg, (size_t)n_seg * mt::kN * sizeof(uint32_t)));
        uint32_t seed[mt::kN]; mt::seed_state(12345u, seed);                  // tracer.cpp:284
        {
            cudaError_t e0 = cudaMemcpyAsync(seg, seed, sizeof(seed), cudaMemcpyHostToDevice, g_ctx.stream);
            if (e0 != cudaSuccess) { cudaFree(seg); return fail(RT_ERR_CUDA, std::string("jitter checkpoint table: ") + cudaGetErrorString(e0)); }
        }
        for (int l = top - 1; l >= 0; --l) {
            const long long stride = 1ll << l, pairs = (n_seg + 2 * stride - 1) / (2 * stride);
            rtd::mt_jump_kernel<<<(unsigned)pairs, 640, rtd::kJumpSmemBytes, g_ctx.stream>>>(g_ctx.d_jump_poly + (size_t)(l + kSegLog2Twists) * mt::kN, seg, stride, n_seg);
        }
        rtd::mt_checkpoint_kernel<<<(unsigned)n_seg, 640, 0, g_ctx.stream>>>(g_ctx.mt_table, seg, 0, tiles, kSegTiles);
        cudaError_t e = cudaGetLastError();
        if (e == cudaSuccess) e = cudaStreamSynchronize(g_ctx.stream);
        cudaFree(seg);
        if (e != cudaSuccess) return fail(RT_ERR_CUDA, std::string("jitter checkpoint table: ") + cudaGetErrorString(e));
    }
    g_ctx.mt_tiles = tiles;
    return RT_OK;
}

}  // namespace

namespace {

template <class T> int grow(T** p, size_t* cap, size_t bytes) {
    if (*cap >= bytes) return RT_OK;
    cudaFree(*p); *p = nullptr; *cap = 0;
    CU(cudaMalloc((void**)p, bytes));
    *cap = bytes;
    return RT_OK;
}

int ensure_frame_buffers(rt_scene* s, bool want_f, bool want_ids, int mode) {
    const size_t npx = (size_t)s->width * s->height;
    int rc = grow(&g_ctx.d_rgba8, &g_ctx.rgba8_cap, npx * sizeof(uchar4)); if (rc) return rc;
    if (want_f && (rc = grow(&g_ctx.d_rgbf, &g_ctx.rgbf_cap, npx * sizeof(float4)))) return rc;
    if (want_ids && (rc = grow(&g_ctx.d_ids, &g_ctx.ids_cap, npx * (mode == RT_MODE_PAPER ? 1 : 8) * sizeof(int)))) return rc;
    if (mode == RT_MODE_PAPER && (rc = grow(&g_ctx.d_gbuf, &g_ctx.gbuf_cap, npx * sizeof(rtd::GBuf)))) return rc;
    return RT_OK;
}

// Launch the frame on `st`.  Counters are reset and read back by the caller.
int launch_frame(rt_scene* s, int mode, int rank, int world, uchar4* rgba8, float4* rgbf, int* ids, cudaStream_t st, int* n_tiles_out) {
    const long long npx = (long long)s->width * s->height;
    int rc = bind_scene(s, st); if (rc) return rc;
    CU(cudaMemsetAsync(g_ctx.counters, 0, 4 * sizeof(unsigned long long), st));
    if (mode == RT_MODE_PAPER) {
        { int rc2 = grow(&g_ctx.d_gbuf, &g_ctx.gbuf_cap, (size_t)npx * sizeof(rtd::GBuf)); if (rc2) return rc2; }
        const unsigned blocks = (unsigned)((npx + 255) / 256);
        // kernel variant: <scene staged in shared memory, scene has clustered runs>
        // kernel variant: <scene staged in shared memory, clustered runs, round-level masks>
        const bool mm = s->hdr.cull_words != 0;
        auto paper = s->clustered ? (s->fits_const ? rtd::paper_gbuffer_kernel<true, true, false> : rtd::paper_gbuffer_kernel<false, true, false>)
                   : s->fits_const ? (mm ? rtd::paper_gbuffer_kernel<true, false, true> : rtd::paper_gbuffer_kernel<true, false, false>)
                                   : (mm ? rtd::paper_gbuffer_kernel<false, false, true> : rtd::paper_gbuffer_kernel<false, false, false>);
        paper<<<blocks, 256, s->smem_bytes, st>>>(g_ctx.d_gbuf, ids, g_ctx.counters, rank, world);
        rtd::paper_post_kernel<<<blocks, 256, 0, st>>>(g_ctx.d_gbuf, rgba8, rgbf, s->width, s->height, rank, world);
        CU(cudaGetLastError());
        if (n_tiles_out) *n_tiles_out = (s->height + rtd::kPaperRows - 1) / rtd::kPaperRows;
        return RT_OK;
    }
    const long long tiles = (npx + mt::kPixelsPerTile - 1) / mt::kPixelsPerTile;
    rc = ensure_mt_table(tiles); if (rc) return rc;
    const long long local = tiles > rank ? (tiles - rank + world - 1) / world : 0;
    rtd::RenderParams prm{};
    prm.mt_table = g_ctx.mt_table; prm.rgba8 = rgba8; prm.rgbf = rgbf; prm.ids = ids;
    prm.counters = g_ctx.counters; prm.tile_counter = g_ctx.tile_counter;
    prm.n_local_tiles = (int)local; prm.rank = rank; prm.world = world; prm.n_pixels = npx;
    CU(cudaMemsetAsync(g_ctx.tile_counter, 0, sizeof(unsigned), st));
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
    const long long grid = std::max<long long>(1, std::min<long long>(units, (long long)g_ctx.sms * std::max(per_sm, 1)));
    if (s->bounces) {
        const size_t stride = (size_t)rtd::wave_stride(s->hdr.recursion);
        int rcw = grow(&g_ctx.d_wave, &g_ctx.wave_cap, (size_t)grid * rtd::kWaveSamples * stride * sizeof(float4)); if (rcw) return rcw;
        prm.wave_state = g_ctx.d_wave;
    }
    if (local > 0) {
        kernel<<<(unsigned)grid, threads, smem, st>>>(prm);
        CU(cudaGetLastError());
    }
    if (n_tiles_out) *n_tiles_out = (int)local;
    return RT_OK;
}

int fill_stats(rt_scene* s, int mode, int n_tiles, float kernel_ms, double total_ms, rt_stats* stats) {
    if (!stats) return RT_OK;
    unsigned long long c[4];
    CU(cudaMemcpy(c, g_ctx.counters, sizeof(c), cudaMemcpyDeviceToHost));
    stats->kernel_ms = kernel_ms; stats->total_ms = total_ms;
    stats->primary_rays = c[0]; stats->shadow_rays = c[1]; stats->secondary_rays = c[2]; stats->prim_tests = c[3];
    stats->width = s->width; stats->height = s->height; stats->samples_per_pixel = mode == RT_MODE_PAPER ? 1 : 8;
    stats->n_tiles = n_tiles;
    return RT_OK;
}

int check_args(rt_scene* s, int mode, int rank, int world) {
    if (!s) return fail(RT_ERR_ARG, "render: null scene");
    if (!g_ctx.ready) return fail(RT_ERR_STATE, "render: rt_init has not been called");
    if (mode != RT_MODE_STANDARD && mode != RT_MODE_PAPER) return fail(RT_ERR_ARG, "render: unknown mode");
    if (world < 1 || rank < 0 || rank >= world) return fail(RT_ERR_ARG, "render: need 0 <= rank < world");
    return RT_OK;
}

}  // namespace

extern "C" {

int rt_render_device(rt_scene* s, int mode, int rank, int world, void* rgba8_device, void* rgbf_device, void* ids_device,
                     void* stream, rt_stats* stats) {
    int rc = check_args(s, mode, rank, world); if (rc) return rc;
    if (!rgba8_device) return fail(RT_ERR_ARG, "rt_render_device: rgba8_device is required");
    std::lock_guard<std::mutex> lk(g_ctx.mu);
    CU(cudaSetDevice(g_ctx.device));
    cudaStream_t st = stream ? (cudaStream_t)stream : g_ctx.stream;
    const auto t0 = std::chrono::steady_clock::now();
    int n_tiles = 0;
    if (mode == RT_MODE_STANDARD) {         // table growth (one-off) must not sit between the events
        const long long npx = (long long)s->width * s->height;
        rc = ensure_mt_table((npx + mt::kPixelsPerTile - 1) / mt::kPixelsPerTile); if (rc) return rc;
    }
    rc = bind_scene(s, st); if (rc) return rc;
    CU(cudaEventRecord(g_ctx.ev0, st));
    rc = launch_frame(s, mode, rank, world, (uchar4*)rgba8_device, (float4*)rgbf_device, (int*)ids_device, st, &n_tiles);
    if (rc) return rc;
    CU(cudaEventRecord(g_ctx.ev1, st));
    if (stats) {
        CU(cudaEventSynchronize(g_ctx.ev1));
        float ms = 0.f; CU(cudaEventElapsedTime(&ms, g_ctx.ev0, g_ctx.ev1));
        const double total = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
        return fill_stats(s, mode, n_tiles, ms, total, stats);
    }
    return RT_OK;
}

}  // extern "C"

// ---- Paper frame without any allocation (rt_render_paper_device_noalloc) ----------------------------------------
// Scratch of the entry point below: 256 bytes of device counters (cleared by every call), then the G-buffer of
// paper_gbuffer_kernel / paper_post_kernel for at most RT_NOALLOC_MAX_PIXELS pixels.  The array lives in the module's
// global segment of every device that loads the library (40 MiB, until the process exits).  In host code the name below
// is only a stand-in for the symbol: the device address comes from cudaGetSymbolAddress, once per device.
static constexpr size_t kNoAllocHead = 256;
static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoAllocHead + (size_t)RT_NOALLOC_MAX_PIXELS * sizeof(rtd::GBuf)];
static constexpr int kNoAllocMaxDevices = 64;
static unsigned char* g_noalloc_ptr[kNoAllocMaxDevices] = {};      // device address of g_noalloc_scratch, per device ordinal

namespace {
// Set-up, once per device, with that device current (the caller holds g_ctx.mu and has called cudaSetDevice).
int noalloc_scratch(int device, unsigned char** out) {
    if (device < 0 || device >= kNoAllocMaxDevices) return fail(RT_ERR_LIMIT, "rt_render_paper_device_noalloc: device ordinal beyond 63");
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
    std::lock_guard<std::mutex> lk(g_ctx.mu);
    CU(cudaSetDevice(g_ctx.device));
    unsigned char* scratch = nullptr;
    rc = noalloc_scratch(g_ctx.device, &scratch); if (rc) return rc;
    auto* counters = (unsigned long long*)scratch;
    auto* gbuf = (rtd::GBuf*)(scratch + kNoAllocHead);
    cudaStream_t st = stream ? (cudaStream_t)stream : g_ctx.stream;
    rc = bind_scene(s, st); if (rc) return rc;
    // the counters must start at zero; the G-buffer need not: pass 1 writes every entry pass 2 reads (own rows + halo)
    CU(cudaMemsetAsync(counters, 0, kNoAllocHead, st));
    const unsigned blocks = (unsigned)((npx + 255) / 256);
    const bool mm = s->hdr.cull_words != 0;
    auto paper = s->clustered ? (s->fits_const ? rtd::paper_gbuffer_kernel<true, true, false> : rtd::paper_gbuffer_kernel<false, true, false>)
               : s->fits_const ? (mm ? rtd::paper_gbuffer_kernel<true, false, true> : rtd::paper_gbuffer_kernel<true, false, false>)
                               : (mm ? rtd::paper_gbuffer_kernel<false, false, true> : rtd::paper_gbuffer_kernel<false, false, false>);
    paper<<<blocks, 256, s->smem_bytes, st>>>(gbuf, nullptr, counters, rank, world);
    rtd::paper_post_kernel<<<blocks, 256, 0, st>>>(gbuf, (uchar4*)rgba8_device, nullptr, s->width, s->height, rank, world);
    CU(cudaGetLastError());
    return RT_OK;
}

int rt_peer_frame_create(size_t bytes, void** device_ptr, unsigned char handle[RT_IPC_HANDLE_BYTES]) {
    static_assert(sizeof(cudaIpcMemHandle_t) == RT_IPC_HANDLE_BYTES, "IPC handle size");
    if (!device_ptr || !handle || bytes == 0) return fail(RT_ERR_ARG, "rt_peer_frame_create: null argument or zero size");
    if (!g_ctx.ready) return fail(RT_ERR_STATE, "rt_peer_frame_create: rt_init has not been called");
    std::lock_guard<std::mutex> lk(g_ctx.mu);
    CU(cudaSetDevice(g_ctx.device));
    void* p = nullptr;
    CU(cudaMalloc(&p, bytes));
    CU(cudaMemset(p, 0, bytes));
    cudaIpcMemHandle_t h;
    cudaError_t e = cudaIpcGetMemHandle(&h, p);
    if (e != cudaSuccess) { cudaFree(p); return fail(RT_ERR_CUDA, std::string("cudaIpcGetMemHandle: ") + cudaGetErrorString(e)); }
    std::memcpy(handle, &h, sizeof h);
    *device_ptr = p;
    return RT_OK;
}

int rt_peer_frame_open(const unsigned char handle[RT_IPC_HANDLE_BYTES], void** device_ptr) {
    if (!device_ptr || !handle) return fail(RT_ERR_ARG, "rt_peer_frame_open: null argument");
    if (!g_ctx.ready) return fail(RT_ERR_STATE, "rt_peer_frame_open: rt_init has not been called");
    std::lock_guard<std::mutex> lk(g_ctx.mu);
    CU(cudaSetDevice(g_ctx.device));
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
    if (!g_ctx.ready) return fail(RT_ERR_STATE, "rt_peer_frame_close: rt_init has not been called");
    std::lock_guard<std::mutex> lk(g_ctx.mu);
    CU(cudaSetDevice(g_ctx.device));
    CU(cudaDeviceSynchronize());
    if (is_owner) CU(cudaFree(device_ptr)); else CU(cudaIpcCloseMemHandle(device_ptr));
    return RT_OK;
}

int rt_render(rt_scene* s, int mode, int rank, int world, uint8_t* rgb8, float* rgbf, int32_t* ids, rt_stats* stats) {
    int rc = check_args(s, mode, rank, world); if (rc) return rc;
    std::lock_guard<std::mutex> lk(g_ctx.mu);
    CU(cudaSetDevice(g_ctx.device));
    const auto t0 = std::chrono::steady_clock::now();
    cudaStream_t st = g_ctx.stream;
    const long long npx = (long long)s->width * s->height;
    rc = ensure_frame_buffers(s, rgbf != nullptr, ids != nullptr, mode); if (rc) return rc;
    if (mode == RT_MODE_STANDARD) { rc = ensure_mt_table((npx + mt::kPixelsPerTile - 1) / mt::kPixelsPerTile); if (rc) return rc; }
    rc = bind_scene(s, st); if (rc) return rc;
    if (world > 1) {   // pixels of other ranks: defined (zero / -1) so that slices can be summed or compared
        CU(cudaMemsetAsync(g_ctx.d_rgba8, 0, (size_t)npx * sizeof(uchar4), st));
        if (rgbf) CU(cudaMemsetAsync(g_ctx.d_rgbf, 0, (size_t)npx * sizeof(float4), st));
        if (ids) CU(cudaMemsetAsync(g_ctx.d_ids, 0xff, (size_t)npx * (mode == RT_MODE_PAPER ? 1 : 8) * sizeof(int), st));
    }
    int n_tiles = 0;
    CU(cudaEventRecord(g_ctx.ev0, st));
    rc = launch_frame(s, mode, rank, world, g_ctx.d_rgba8, rgbf ? g_ctx.d_rgbf : nullptr, ids ? g_ctx.d_ids : nullptr, st, &n_tiles);
    if (rc) return rc;
    CU(cudaEventRecord(g_ctx.ev1, st));
    const unsigned blocks = (unsigned)((npx + 255) / 256);
    if (rgb8) {
        const size_t need = (size_t)npx * 3;
        rc = grow(&g_ctx.d_pack, &g_ctx.pack_cap, std::max(need, (size_t)npx * 12)); if (rc) return rc;
        rtd::pack_rgb_kernel<<<blocks, 256, 0, st>>>(g_ctx.d_rgba8, g_ctx.d_pack, npx);
        CU(cudaMemcpyAsync(rgb8, g_ctx.d_pack, need, cudaMemcpyDeviceToHost, st));
        CU(cudaStreamSynchronize(st));
    }
    if (rgbf) {
        rc = grow(&g_ctx.d_pack, &g_ctx.pack_cap, (size_t)npx * 12); if (rc) return rc;
        rtd::unpack_rgbf_kernel<<<blocks, 256, 0, st>>>(g_ctx.d_rgbf, (float*)g_ctx.d_pack, npx);
        CU(cudaMemcpyAsync(rgbf, g_ctx.d_pack, (size_t)npx * 12, cudaMemcpyDeviceToHost, st));
        CU(cudaStreamSynchronize(st));
    }
    if (ids) CU(cudaMemcpyAsync(ids, g_ctx.d_ids, (size_t)npx * (mode == RT_MODE_PAPER ? 1 : 8) * sizeof(int), cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    CU(cudaGetLastError());
    float ms = 0.f; CU(cudaEventElapsedTime(&ms, g_ctx.ev0, g_ctx.ev1));
    const double total = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    return fill_stats(s, mode, n_tiles, ms, total, stats);
}

int rt_trace_rays(rt_scene* s, int n, const double* rays, double* out) {
    if (!s || !rays || !out || n < 0) return fail(RT_ERR_ARG, "rt_trace_rays: bad argument");
    if (!g_ctx.ready) return fail(RT_ERR_STATE, "rt_trace_rays: rt_init has not been called");
    if (n == 0) return RT_OK;
    std::lock_guard<std::mutex> lk(g_ctx.mu);
    CU(cudaSetDevice(g_ctx.device));
    cudaStream_t st = g_ctx.stream;
    int rc = bind_scene(s, st); if (rc) return rc;
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
    if (!g_ctx.ready) { int rc = rt_init(-1); if (rc) return rc; }
    if (n == 0) return RT_OK;
    std::lock_guard<std::mutex> lk(g_ctx.mu);
    CU(cudaSetDevice(g_ctx.device));
    const long long tiles = (long long)((first_sample + (uint64_t)n + mt::kSamplesPerTile - 1) / mt::kSamplesPerTile);
    int rc = ensure_mt_table(tiles); if (rc) return rc;
    double* d_out = nullptr;
    CU(cudaMalloc((void**)&d_out, (size_t)n * 2 * sizeof(double)));
    rtd::jitter_debug_kernel<<<1, 256, 0, g_ctx.stream>>>(g_ctx.mt_table, first_sample, n, d_out);
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaMemcpyAsync(out, d_out, (size_t)n * 2 * sizeof(double), cudaMemcpyDeviceToHost, g_ctx.stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(g_ctx.stream);
    cudaFree(d_out);
    if (e != cudaSuccess) return fail(RT_ERR_CUDA, std::string("rt_debug_jitter: ") + cudaGetErrorString(e));
    return RT_OK;
}

int rt_measure_fp32_peak(double* tflops) {
    if (!tflops) return fail(RT_ERR_ARG, "rt_measure_fp32_peak: null argument");
    if (!g_ctx.ready) { int rc = rt_init(-1); if (rc) return rc; }
    std::lock_guard<std::mutex> lk(g_ctx.mu);
    CU(cudaSetDevice(g_ctx.device));
    float* sink = nullptr; CU(cudaMalloc((void**)&sink, sizeof(float)));
    const int iters = 4096, blocks = g_ctx.sms * 8, threads = 256;
    double best = 0.0;
    for (int rep = 0; rep < 5; ++rep) {
        CU(cudaEventRecord(g_ctx.ev0, g_ctx.stream));
        rtd::ffma_peak_kernel<<<blocks, threads, 0, g_ctx.stream>>>(sink, iters, 1.0000001f, 1e-9f);
        CU(cudaEventRecord(g_ctx.ev1, g_ctx.stream));
        CU(cudaEventSynchronize(g_ctx.ev1));
        float ms = 0.f; CU(cudaEventElapsedTime(&ms, g_ctx.ev0, g_ctx.ev1));
        const double flops = 2.0 * 8 * 16 * (double)iters * blocks * threads;
        if (rep > 0) best = std::max(best, flops / (ms * 1e-3) * 1e-12);
    }
    cudaFree(sink);
    *tflops = best;
    return RT_OK;
}

}  // extern "C"
