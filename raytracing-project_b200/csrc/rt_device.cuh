// rt_device.cuh — device-side scene layout and ray queries (sm_100a, FP32).
//
// Layout in HBM / constant memory:
//   geom[]  float4 per object/leaf  (sphere: c.xyz, r | half-space: n.xyz, n.p0)
//   meta[]  int4   per object/leaf  (kind | flags<<8, material, aux, xf)
// split so that the closest-hit / any-hit scans read 16 B per primitive and touch meta only for
// the winner.  Each CTA stages the scanned arrays into shared memory once (broadcast LDS.128 per
// primitive per warp); scenes beyond the shared budget use the same arrays through the global
// read-only path.  (A first version kept them in __constant__ memory: ncu showed every load going
// through the address-divergence unit, 31 % ADU utilisation, because the scans sit inside divergent
// per-lane loops where the index is not provably uniform.)  `runs` groups consecutive plain primitives of one kind (JSON order
// is preserved, scene.cpp:15-21) so the scan is a few tight loops instead of a per-object switch.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/rt_b200.h"
#include "rt_cull.cuh"
#include "rt_grid.cuh"

namespace rtd {

struct SceneHeader {
    double eye[3], P[3], lx, ly;
    double inv_nx, inv_ny;      // RN(1 / nx), RN(1 / ny): div_by() below
    int nx, ny;
    float bg[3]; float medium_index;
    int recursion, n_objects, n_lights, n_runs, n_leaves, n_ops, n_clusters;
    // shared-memory image (units of 16 B from the start of dynamic shared memory; ops in bytes)
    int off_obj_meta, off_leaf_geom, off_leaf_meta, off_runs, off_lights, off_clusters, off_masks, off_ops_bytes, smem_bytes;
    float mask_scale;           // unused by the device since the margin moved into the table (kept for rt_debug_mask_table)
    const float4* obj_geom; const int4* obj_meta;
    const float4* leaf_geom; const int4* leaf_meta;
    const float4* pokeballs;    // 3 float4 each
    const float4* affines;      // 3 float4 each (rows)
    const uint32_t* ops;
    const float4* materials;    // 3 float4 each: albedo.rgb ks | ambient.rgb kr | kt shininess ior kd
    const float4* lights;       // 2 float4 each: pos | intensity
    const float4* fallbacks;
    const int4* runs;           // kind, first, count, 1 + first cluster (0: run is not clustered)
    const float4* clusters;     // bounding sphere of every 16 consecutive primitives of a clustered run
    const int* obj_orig;        // JSON index of device object i (clustered runs are reordered spatially)
    const float4* masks;        // (c.xyz, |c|^2 - r^2 - 2^-16 (|c|^2 + r^2)) of every entry of obj_geom, leaf_geom, clusters (in that order)
    // round-level culling (rt_cull.cuh): (centre, radius | inf) of objects | leaves — a generic object is represented by its
    // world bound, a plain half-space by (n, n.p0) with its bit set in cull_plane_bits — both sections padded to a multiple of
    // 32 with w < 0; candidate bit of object i: i, of leaf i: cull_leaf_word * 32 + i
    const float4* cull_geom;
    const uint32_t* cull_plane_bits;                  // one bit per entry: the entry is a half-space surface
    int cull_words, cull_leaf_word;                   // words per bundle (0: scene not culled), first word of the leaf section
    int off_cand_bytes;                               // CandShared + masks in dynamic shared memory
    // uniform grid over the plain spheres / pokeballs (rt_grid.cuh): scenes scanned by incoherent rays; grid.words == 0: none
    GridHeader grid;
    const uint32_t* grid_masks;                       // 4 candidate words per cell (+ the ALWAYS cell), bit i = device object i
    int off_grid_bytes, n_grid_cells;                 // staged copy: byte offset in the shared-memory image; cells incl. the ALWAYS cell
};

enum { RUN_SPHERE = 0, RUN_HALFSPACE = 1, RUN_POKEBALL = 2, RUN_GENERIC = 3 };

// A/B build switch: 1 = the candidate scan evaluates the exact discriminant on the geometry arrays (13 FMA-pipe
// instructions per primitive) instead of the conservative expanded form on the mask table (7).
#ifndef RT_MASK_EXACT
#define RT_MASK_EXACT 0
#endif

__constant__ SceneHeader c_hdr;
extern __shared__ __align__(16) unsigned char s_scene[];

// Scene view.  C = true: the arrays every ray scans (object/leaf geometry, leaf meta, CSG ops, runs,
// lights) were staged into shared memory by stage_scene() — a warp reading one primitive is a single
// broadcast LDS.128, independent of control-flow divergence.  C = false (scene larger than the shared
// budget): the same arrays through the global read-only path.  Per-hit data (object meta, pokeball
// tables, affine maps, materials) is indexed per lane and always comes from global memory.
template <bool C> struct SV {
    static __device__ __forceinline__ float4 obj_geom(int i)  { return C ? reinterpret_cast<const float4*>(s_scene)[i] : __ldg(c_hdr.obj_geom + i); }
    // Scenes with bounds (K) do not stage the top-level geometry: the scan reads the mask table, geometry is only needed to
    // finish the few candidates, and without it the 1 100-primitive scene keeps 4 CTAs per SM (43 -> 26 KB per CTA).
    template <bool K> static __device__ __forceinline__ float4 obj_geom_k(int i) { return (C && !K) ? reinterpret_cast<const float4*>(s_scene)[i] : __ldg(c_hdr.obj_geom + i); }
    // object meta: staged next to the geometry (scenes without bounds: the masked scans read it per candidate), else global
#ifndef RT_META_SMEM
#define RT_META_SMEM 1
#endif
    template <bool K> static __device__ __forceinline__ int4 obj_meta_k(int i) {
        return (RT_META_SMEM && C && !K) ? reinterpret_cast<const int4*>(s_scene)[c_hdr.off_obj_meta + i] : __ldg(c_hdr.obj_meta + i);
    }
    static __device__ __forceinline__ float4 leaf_geom(int i) { return C ? reinterpret_cast<const float4*>(s_scene)[c_hdr.off_leaf_geom + i] : __ldg(c_hdr.leaf_geom + i); }
    static __device__ __forceinline__ int4   leaf_meta(int i) { return C ? reinterpret_cast<const int4*>(s_scene)[c_hdr.off_leaf_meta + i] : __ldg(c_hdr.leaf_meta + i); }
    static __device__ __forceinline__ float4 pokeball(int i, int k) { return __ldg(c_hdr.pokeballs + i * 3 + k); }
    static __device__ __forceinline__ float4 affine(int i, int k)   { return __ldg(c_hdr.affines + i * 3 + k); }
    static __device__ __forceinline__ uint32_t op(int i)      { return C ? reinterpret_cast<const uint32_t*>(s_scene + c_hdr.off_ops_bytes)[i] : __ldg(c_hdr.ops + i); }
    static __device__ __forceinline__ float4 light(int i, int k)    { return C ? reinterpret_cast<const float4*>(s_scene)[c_hdr.off_lights + i * 2 + k] : __ldg(c_hdr.lights + i * 2 + k); }
    static __device__ __forceinline__ int4   run(int i)       { return C ? reinterpret_cast<const int4*>(s_scene)[c_hdr.off_runs + i] : __ldg(c_hdr.runs + i); }
    static __device__ __forceinline__ float4 fallback(int i)  { return __ldg(c_hdr.fallbacks + i); }
    static __device__ __forceinline__ float4 cluster(int i)   { return C ? reinterpret_cast<const float4*>(s_scene)[c_hdr.off_clusters + i] : __ldg(c_hdr.clusters + i); }
    static __device__ __forceinline__ int    obj_orig(int i)  { return __ldg(c_hdr.obj_orig + i); }
    // the four candidate words of grid cell i (rt_grid.cuh): staged behind the CSG ops when the scene is staged
    static __device__ __forceinline__ GridWords grid_cell(int i) {
        const uint4 v = C ? reinterpret_cast<const uint4*>(s_scene + c_hdr.off_grid_bytes)[i] : __ldg(reinterpret_cast<const uint4*>(c_hdr.grid_masks) + i);
        GridWords g; g.x = v.x; g.y = v.y; g.z = v.z; g.w = v.w; return g;
    }
    // base pointers for sphere_mask (the mask table mirrors obj_geom | leaf_geom | clusters)
    static __device__ __forceinline__ const float4* mask_ptr(int i)    { return C ? reinterpret_cast<const float4*>(s_scene) + c_hdr.off_masks + i : c_hdr.masks + i; }
#if RT_MASK_EXACT
    static __device__ __forceinline__ const float4* obj_ptr(int i)     { return C ? reinterpret_cast<const float4*>(s_scene) + i : c_hdr.obj_geom + i; }
    static __device__ __forceinline__ const float4* leaf_ptr(int i)    { return C ? reinterpret_cast<const float4*>(s_scene) + c_hdr.off_leaf_geom + i : c_hdr.leaf_geom + i; }
    static __device__ __forceinline__ const float4* cluster_ptr(int i) { return C ? reinterpret_cast<const float4*>(s_scene) + c_hdr.off_clusters + i : c_hdr.clusters + i; }
#else
    static __device__ __forceinline__ const float4* obj_ptr(int i)     { return mask_ptr(i); }
    static __device__ __forceinline__ const float4* leaf_ptr(int i)    { return mask_ptr(c_hdr.n_objects + i); }
    static __device__ __forceinline__ const float4* cluster_ptr(int i) { return mask_ptr(c_hdr.n_objects + c_hdr.n_leaves + i); }
#endif
};
// Copy the scanned arrays into shared memory (all threads of the CTA; ends with a barrier).
template <bool C> __device__ __forceinline__ void stage_scene() {
    if (!C) return;
    float4* dst = reinterpret_cast<float4*>(s_scene);
    const int t = threadIdx.x, n = blockDim.x;
    for (int i = t; i < c_hdr.off_obj_meta; i += n) {                                         // 0 entries when the scene has bounds
        dst[i] = __ldg(c_hdr.obj_geom + i);
        reinterpret_cast<int4*>(dst)[c_hdr.off_obj_meta + i] = __ldg(c_hdr.obj_meta + i);
    }
    for (int i = t; i < c_hdr.n_leaves; i += n) {
        dst[c_hdr.off_leaf_geom + i] = __ldg(c_hdr.leaf_geom + i);
        reinterpret_cast<int4*>(dst)[c_hdr.off_leaf_meta + i] = __ldg(c_hdr.leaf_meta + i);
    }
    for (int i = t; i < c_hdr.n_runs; i += n) reinterpret_cast<int4*>(dst)[c_hdr.off_runs + i] = __ldg(c_hdr.runs + i);
    for (int i = t; i < 2 * c_hdr.n_lights; i += n) dst[c_hdr.off_lights + i] = __ldg(c_hdr.lights + i);
    for (int i = t; i < c_hdr.n_clusters; i += n) dst[c_hdr.off_clusters + i] = __ldg(c_hdr.clusters + i);
    const int n_masks = c_hdr.n_objects + c_hdr.n_leaves + c_hdr.n_clusters;
    for (int i = t; i < n_masks; i += n) dst[c_hdr.off_masks + i] = __ldg(c_hdr.masks + i);
    uint32_t* ops = reinterpret_cast<uint32_t*>(s_scene + c_hdr.off_ops_bytes);
    for (int i = t; i < c_hdr.n_ops; i += n) ops[i] = __ldg(c_hdr.ops + i);
    uint4* gc = reinterpret_cast<uint4*>(s_scene + c_hdr.off_grid_bytes);
    for (int i = t; i < c_hdr.n_grid_cells; i += n) gc[i] = __ldg(reinterpret_cast<const uint4*>(c_hdr.grid_masks) + i);
    __syncthreads();
}
// materials are indexed per lane (divergent) -> always the global read-only path
__device__ __forceinline__ float4 material(int m, int k) { return __ldg(c_hdr.materials + m * 3 + k); }

constexpr float kInf = __builtin_huge_valf();
// ---- warp-level helpers --------------------------------------------------------------------------
constexpr unsigned kFullWarp = 0xffffffffu;
// min / max over the warp into a uniform register (CREDUX); a NaN in any lane gives NaN
__device__ __forceinline__ float warp_max_nan(float v) {
    float r; asm volatile("redux.sync.max.NaN.f32 %0, %1, 0xffffffff;" : "=f"(r) : "f"(v)); return r;
}
__device__ __forceinline__ float warp_min_nan(float v) {
    float r; asm volatile("redux.sync.min.NaN.f32 %0, %1, 0xffffffff;" : "=f"(r) : "f"(v)); return r;
}

// ---- small vector helpers ---------------------------------------------------------------------
struct V3 { float x, y, z; };
__device__ __forceinline__ V3 mk(float x, float y, float z) { V3 v; v.x = x; v.y = y; v.z = z; return v; }
__device__ __forceinline__ V3 operator+(V3 a, V3 b) { return mk(a.x + b.x, a.y + b.y, a.z + b.z); }
__device__ __forceinline__ V3 operator-(V3 a, V3 b) { return mk(a.x - b.x, a.y - b.y, a.z - b.z); }
__device__ __forceinline__ V3 operator*(V3 a, float s) { return mk(a.x * s, a.y * s, a.z * s); }
__device__ __forceinline__ V3 neg(V3 a) { return mk(-a.x, -a.y, -a.z); }
__device__ __forceinline__ float dot(V3 a, V3 b) { return fmaf(a.x, b.x, fmaf(a.y, b.y, a.z * b.z)); }
__device__ __forceinline__ V3 madd(V3 a, V3 d, float t) { return mk(fmaf(d.x, t, a.x), fmaf(d.y, t, a.y), fmaf(d.z, t, a.z)); }
// Vec4::normalized (core.h:58-64): length <= 1e-6 falls back to (0,1,0)
#ifndef RT_RSQRT
#define RT_RSQRT 1          // length and reciprocal length from ONE MUFU.RSQ (sqrt is x * rsqrt(x) under -prec-sqrt=false anyway)
#endif
__device__ __forceinline__ V3 normalized(V3 a) {
    float l2 = dot(a, a);
#if RT_RSQRT
    const float inv = rsqrtf(l2);
    if (l2 * inv > 1e-6f) return mk(a.x * inv, a.y * inv, a.z * inv);        // (l2 == 0: 0 * inf = NaN fails the test like L = 0)
#else
    float L = sqrtf(l2);
    if (L > 1e-6f) { float inv = 1.0f / L; return mk(a.x * inv, a.y * inv, a.z * inv); }
#endif
    return mk(0.f, 1.f, 0.f);
}

}  // namespace rtd
#include "rt_csg.cuh"
namespace rtd {

struct Hit {
    float t; V3 n; int mat; int code; bool front;
};

// per-thread ray statistics (rt_stats)
struct Counters { unsigned primary, shadow, secondary, tests; };

// ---- candidate masks of one round (rt_cull.cuh) ---------------------------------------------------------
// A CTA traces 256 samples = 32 neighbouring pixels per round.  Before the camera rays are traced, and again before their
// shadow rays are, the CTA culls EVERY cullable entry of the scene — top-level objects (CSG programs and wrapped primitives
// by their world bound, half-spaces as planes) and CSG leaves: the table c_hdr.cull_geom, (centre, radius) or radius = inf
// for entries that cannot be culled — against the
// round's bundle, one entry per thread, and leaves one candidate bit per entry in shared memory: bundle 0 = the camera
// rays, bundle 1 + l = the shadow rays towards light l (l < kMaxListLights).  The scans then take their candidates from
// these words instead of testing every primitive per ray.  The boxes the bundles are built from are reduced with CREDUX
// per warp and one shared-memory atomic per warp and component (floats mapped to order-preserving ints).
constexpr int kMaxListLights = 8;
struct CandShared {
    int dir_lo[3], dir_hi[3];              // camera rays: box of the directions
    int org_lo[3], org_hi[3];              // shadow rays: box of the origins
    int eps_max, nan_flags;                // largest shadow epsilon; bit 0 / 1: a NaN in the first / second box
    int pad[2];
    // followed by (1 + kMaxListLights) * c_hdr.cull_words mask words
};
// The boxes and the mask words live in STATIC shared memory: their addresses are compile-time constants (behind the staged
// scene they cost a constant-bank load and an add at every use: 2.6 % of the snorlax kernel's instructions).
constexpr int kCandWordsMax = 16;          // words per bundle: scenes of up to 512 cullable entries (the library picks <= 8)
#ifndef RT_CAND_STATIC
#define RT_CAND_STATIC 1
#endif
#if RT_CAND_STATIC
__device__ __forceinline__ CandShared* cand_shared() { __shared__ CandShared S; return &S; }
__device__ __forceinline__ uint32_t* cand_words(int bundle) {
    __shared__ uint32_t words[(1 + kMaxListLights) * kCandWordsMax];
    return words + bundle * kCandWordsMax;
}
#else
__device__ __forceinline__ CandShared* cand_shared() { return reinterpret_cast<CandShared*>(s_scene + c_hdr.off_cand_bytes); }
__device__ __forceinline__ uint32_t* cand_words(int bundle) {
    return reinterpret_cast<uint32_t*>(s_scene + c_hdr.off_cand_bytes + sizeof(CandShared)) + bundle * c_hdr.cull_words;
}
#endif
__device__ __forceinline__ int f2ord(float f) { const int i = __float_as_int(f); return i ^ ((i >> 31) & 0x7fffffff); }
__device__ __forceinline__ float ord2f(int i) { return __int_as_float(i ^ ((i >> 31) & 0x7fffffff)); }
constexpr int kOrdPosInf = 0x7f800000, kOrdNegInf = (int)0x807fffff;      // f2ord(+inf), f2ord(-inf)

// All threads of the CTA: empty boxes (before the first round of a kernel; ends with a barrier).
__device__ __forceinline__ void cand_init() {
    if (!c_hdr.cull_words) return;
    CandShared* S = cand_shared();
    if (threadIdx.x < 3) { S->dir_lo[threadIdx.x] = kOrdPosInf; S->dir_hi[threadIdx.x] = kOrdNegInf; S->org_lo[threadIdx.x] = kOrdPosInf; S->org_hi[threadIdx.x] = kOrdNegInf; }
    if (threadIdx.x == 3) { S->eps_max = 0; S->nan_flags = 0; }
    __syncthreads();
}
// One warp: add the vectors v of its lanes with `use` to a box (lo, hi in shared memory).
__device__ __forceinline__ void cand_box_add(bool use, V3 v, int* lo, int* hi, int* nan_flags, int nan_bit) {
    const float ax = warp_min_nan(use ? v.x : kInf), ay = warp_min_nan(use ? v.y : kInf), az = warp_min_nan(use ? v.z : kInf);
    const float bx = warp_max_nan(use ? v.x : -kInf), by = warp_max_nan(use ? v.y : -kInf), bz = warp_max_nan(use ? v.z : -kInf);
    if ((threadIdx.x & 31) == 0 && ax <= bx) {                   // (false for an empty warp and for NaN)
        atomicMin(lo + 0, f2ord(ax)); atomicMin(lo + 1, f2ord(ay)); atomicMin(lo + 2, f2ord(az));
        atomicMax(hi + 0, f2ord(bx)); atomicMax(hi + 1, f2ord(by)); atomicMax(hi + 2, f2ord(bz));
    }
    if ((threadIdx.x & 31) == 0 && !(ax == ax && ay == ay && az == az)) atomicOr(nan_flags, nan_bit);
}
// All threads of the CTA, after a barrier that follows the box reductions: bundles [b0, b0 + nb) x all entry words are
// dealt to the warps; a warp builds the bundle it is working on (the same arithmetic in every lane), culls 32 entries and
// lane 0 stores the ballot.  b = 0: camera rays; b >= 1: shadow rays towards light b - 1.  Ends with a barrier.
#ifndef RT_CAND_NOINLINE
#define RT_CAND_NOINLINE 1        // one out-of-line copy of cand_build: camera and shadow masks share it (snorlax 4K 11.87 -> 11.35 ms)
#endif
#if RT_CAND_NOINLINE
__device__ __noinline__ unsigned cand_build(int b0, int nb) {
#else
__device__ __forceinline__ unsigned cand_build(int b0, int nb) {
#endif
    // (the number of bundle tests is RETURNED: a reference parameter of an out-of-line function pins the caller's whole
    // counter struct to local memory, and every ++cnt.x in the kernel becomes a load and a store)
    unsigned tests = 0;
    CandShared* S = cand_shared();
    const int W = c_hdr.cull_words, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
    const bool nan = b0 == 0 ? (S->nan_flags & 1) : (S->nan_flags & 2);
    float lo[3], hi[3];
    const int* box_lo = b0 == 0 ? S->dir_lo : S->org_lo;            // (lo[3] and hi[3] are adjacent in both boxes)
#pragma unroll
    for (int k = 0; k < 3; ++k) { lo[k] = ord2f(box_lo[k]); hi[k] = ord2f(box_lo[3 + k]); }
    const bool empty = !(lo[0] <= hi[0]);                         // no ray contributed: nothing will be asked
    const float eps_max = __int_as_float(S->eps_max);
    Bundle B; B.wide = 1; B.t_hi = 0.f; B.h = 0.f; B.kd = 0.f; B.inv1md = 1.f; B.mx = B.my = B.mz = 0.f; B.cx = B.cy = B.cz = 0.f;
    if (!empty && !nan) {
        // A task = one bundle x `chunk` consecutive words: the bundle is set up once per task, and the tasks are sized so that
        // they spread evenly over the warps (snorlax: 5 lights x 2 words -> 5 tasks of 2 words instead of 10 tasks on 8 warps).
#ifndef RT_CAND_CHUNK
#define RT_CAND_CHUNK 0          // measured: snorlax 4K 10.50 (off) vs 10.72 ms (on); with the leaf-word skip: see DESIGN section 7
#endif
        const int chunk = RT_CAND_CHUNK ? max(1, min(W, (nb * W + nwarps - 1) / nwarps)) : 1;
        const int per_bundle = (W + chunk - 1) / chunk;
        int bb = 0, part = threadIdx.x >> 5;                       // task k = bb * per_bundle + part
        while (part >= per_bundle) { part -= per_bundle; ++bb; }
        while (bb < nb) {                                          // (empty and no NaN: no ray will ask)
            const int b = b0 + bb;
            if (b == 0) { const float eye[3] = {(float)c_hdr.eye[0], (float)c_hdr.eye[1], (float)c_hdr.eye[2]}; bundle_from_dirbox(B, eye, lo, hi); }
            else {
                const float4 lp = __ldg(c_hdr.lights + 2 * (b - 1));
                const float L[3] = {lp.x, lp.y, lp.z};
                bundle_from_box_light(B, lo, hi, eps_max, L, lp.w != 0.f);
            }
            const int w_end = min(W, (part + 1) * chunk);
            // A task that covers the whole bundle culls the CSG leaves only if one of the bundle's generic objects (the
            // programs that own them) stayed a candidate: nobody reads the leaf words otherwise.
            bool leaves_needed = !(RT_CAND_CHUNK && chunk == W);
            for (int w = part * chunk; w < w_end; ++w) {
                if (w >= c_hdr.cull_leaf_word && !leaves_needed) break;
                const float4 g = __ldg(c_hdr.cull_geom + 32 * w + lane);
                const bool plane = (__ldg(c_hdr.cull_plane_bits + w) >> lane) & 1u;
                Sphere4 q; q.x = g.x; q.y = g.y; q.z = g.z; q.w = g.w;
                // padding entries have w < 0; CSG leaves (words from cull_leaf_word on) are culled on the whole line
                // (a bundle too wide to cull — a round across a silhouette, a light inside the box — keeps every entry)
                // (a plane's w is its offset n.p0 and may well be negative: only non-plane entries use w < 0 as "padding")
                const bool cand = plane ? (B.wide || !bundle_rejects_plane(B, q)) : (g.w >= 0.f && (B.wide || !bundle_rejects(B, q, B.t_hi, w < c_hdr.cull_leaf_word)));
                ++tests;
                const uint32_t bits = __ballot_sync(kFullWarp, cand);
                if (lane == 0) cand_words(b)[w] = bits;
                if (RT_CAND_CHUNK && w < c_hdr.cull_leaf_word && (bits & __ldg(c_hdr.cull_plane_bits + 2 * W + w))) leaves_needed = true;
            }
            part += nwarps;
            while (part >= per_bundle) { part -= per_bundle; ++bb; }
        }
    } else if (nan) {                                             // a NaN among the rays: keep every entry
        for (int k = threadIdx.x >> 5; k < nb * W; k += nwarps) {
            const int bb = k / W, w = k - bb * W;
            const bool plane = (__ldg(c_hdr.cull_plane_bits + w) >> lane) & 1u;
            const uint32_t bits = __ballot_sync(kFullWarp, plane || !(__ldg(c_hdr.cull_geom + 32 * w + lane).w < 0.f));
            if (lane == 0) cand_words(b0 + bb)[w] = bits;
        }
    }
    __syncthreads();
    return tests;
}
// Bundle index b as the scans take it (-1: light beyond kMaxListLights, scanned without a mask).  The scans are handed the
// INDEX, not a pointer: a generic 64-bit pointer into shared memory costs two registers across every scan, address-window
// arithmetic at every out-of-line call and generic loads inside the callee; cand_words(b) is an immediate LDS offset.
__device__ __forceinline__ int cand_id(int b) { return b <= kMaxListLights ? b : -1; }
// Candidate bits of entries [first, first + n), n <= 32, of a word array (nullptr: all of them).
__device__ __forceinline__ uint32_t cand_bits(const uint32_t* words, int first, int n) {
    const uint32_t all = n >= 32 ? 0xffffffffu : ((1u << n) - 1u);
    if (!words) return all;
    const int w = first >> 5, sh = first & 31;
    const uint32_t lo = words[w], hi = (sh + n > 32) ? words[w + 1] : 0u;
    return __funnelshift_r(lo, hi, sh) & all;
}

// ---- primitive tests -----------------------------------------------------------------------------
// Quadratic of Sphere::intersect (geometry.cpp:15-22) for a unit direction.  The reference evaluates
// disc = (oc.d)^2 - (oc.oc - r^2) in double; in FP32 that cancels catastrophically once |oc| >> r
// (an error of 1e-4 in t at |oc| = 17, more than the tracer's tmin).  The same quantity is therefore
// computed as r^2 - |oc - (oc.d) d|^2, whose rounding error scales with r instead of |oc|.
__device__ __forceinline__ float sphere_disc(float4 g, V3 o, V3 d, float& hb) {
    V3 oc = mk(o.x - g.x, o.y - g.y, o.z - g.z);
    hb = dot(oc, d);
    V3 l = mk(fmaf(-hb, d.x, oc.x), fmaf(-hb, d.y, oc.y), fmaf(-hb, d.z, oc.z));
    return fmaf(-l.x, l.x, fmaf(-l.y, l.y, fmaf(-l.z, l.z, g.w * g.w)));
}
// Sphere::intersect (geometry.cpp:12-37): near root, else far root, inclusive [tmin,tmax].
__device__ __forceinline__ bool sphere_t(float4 g, V3 o, V3 d, float tmin, float tmax, float& t) {
    float hb;
    float disc = sphere_disc(g, o, d, hb);
    if (disc < 0.f) return false;
    float s = sqrtf(disc);
    float tt = -hb - s;
    if (tt < tmin || tt > tmax) {
        tt = -hb + s;
        if (tt < tmin || tt > tmax) return false;
    }
    t = tt;
    return true;
}
// HalfSpace::intersect (geometry.cpp:90-106): g = (n, n.p0)
__device__ __forceinline__ bool plane_t(float4 g, V3 o, V3 d, float tmin, float tmax, float& t) {
    float nd = fmaf(g.x, d.x, fmaf(g.y, d.y, g.z * d.z));
    if (fabsf(nd) < 1e-12f) return false;
    float no = fmaf(g.x, o.x, fmaf(g.y, o.y, g.z * o.z));
    float tt = (g.w - no) / nd;
    if (tt < tmin || tt > tmax) return false;
    t = tt;
    return true;
}

// Candidate mask of a run of sphere-like primitives: bit k set if the LINE of the ray may touch primitive first+k.
// The mask is CONSERVATIVE (a superset of the primitives whose exact discriminant, sphere_disc, is >= 0; NaN counts as a
// candidate, as in `if (disc < 0) return false`); every caller finishes the candidates in order with exactly the
// arithmetic of sphere_t / leaf_interval and re-checks the discriminant there, so the result is bit-identical to
// testing the primitives one after the other.  That freedom buys the cheap expanded form of the reference's own
// quadratic (geometry.cpp:15-22): with the per-primitive constants m = (c, |c|^2 - r^2) of the mask table
//     hb = o.d - c.d,   s = (|c|^2 - r^2) - 2 o.c,   candidate <=> hb^2 - s >= |o|^2 - margin
// = 7 FMA-pipe instructions + 1 compare per primitive after one broadcast LDS.128, no data-dependent branch, four
// independent chains in flight.  The form cancels (rounding error ~ 2^-24 (|o|^2 + |c|^2), where sphere_disc's is ~ r^2),
// hence the margin 2^-16 (|o|^2 + max(|c|^2 + r^2)): it covers the rounding of both forms (< 70 of the 256 ulp of that scale) and
// | |d|^2 - 1 | <= 2^-19 of directions normalised with the 2-ulp MUFU reciprocal; tests/test_mask_margin.py measures
// the slack.  The primitive's share of the margin, 2^-16 (|c|^2 + r^2), is folded into the table's w on upload, so one huge
// sphere (the r = 1000 "floor") does not loosen the mask of the small ones (scenes/big_floor.json).  cnt <= 32.
// Out of line on purpose: the scans call it from up to ten places (object runs, cluster bounds, cluster members,
// generic-run bounds, CSG sphere runs; closest-hit and any-hit each); one copy keeps the kernel's hot code inside the
// instruction cache (synthetic-A 2K 42.1 -> 30.4 ms, snorlax 4K 17.5 -> 16.8 ms).
constexpr float kMaskMargin = 1.0f / 65536.0f;
__device__ __forceinline__ bool mask_miss(float4 m, V3 m2o, V3 d, float od, float bound) {
    const float hb = fmaf(-m.x, d.x, fmaf(-m.y, d.y, fmaf(-m.z, d.z, od)));
    const float s = fmaf(m.x, m2o.x, fmaf(m.y, m2o.y, fmaf(m.z, m2o.z, m.w)));
    return fmaf(hb, hb, -s) < bound;
}
#if RT_MASK_EXACT
template <bool C> __device__ __noinline__ uint32_t sphere_mask(const float4* base, int cnt, V3 o, V3 d) {
    uint32_t mask = 0;
    int k = 0;
    for (; k + 4 <= cnt; k += 4) {
        uint32_t nib = 0;
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const float4 g = C ? base[k + u] : __ldg(base + k + u);
            float hb;
            nib |= (sphere_disc(g, o, d, hb) < 0.f) ? 0u : (1u << u);
        }
        mask |= nib << k;
    }
    for (; k < cnt; ++k) {
        const float4 g = C ? base[k] : __ldg(base + k);
        float hb;
        mask |= (sphere_disc(g, o, d, hb) < 0.f) ? 0u : (1u << k);
    }
    return mask;
}
#else
template <bool C> __device__ __noinline__ uint32_t sphere_mask(const float4* base, int cnt, V3 o, V3 d) {
    const float oo = dot(o, o), od = dot(o, d);
    const float bound = fmaf(-kMaskMargin, oo, oo) - 1e-30f;              // |o|^2's share; the primitive's is in the table
    const V3 m2o = mk(-2.f * o.x, -2.f * o.y, -2.f * o.z);
    uint32_t mask = 0;
    int k = 0;
    for (; k + 4 <= cnt; k += 4) {
        uint32_t nib = 0;
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const float4 m = C ? base[k + u] : __ldg(base + k + u);
            nib |= mask_miss(m, m2o, d, od, bound) ? 0u : (1u << u);
        }
        mask |= nib << k;
    }
#pragma unroll 1
    for (; k < cnt; ++k) {
        const float4 m = C ? base[k] : __ldg(base + k);
        mask |= mask_miss(m, m2o, d, od, bound) ? 0u : (1u << k);
    }
    return mask;
}
#endif

// pick_region_material (geometry.cpp:163-180) on the unit offset u; returns 0..4 = top,bottom,belt,ring,button
template <bool C> __device__ __forceinline__ int pokeball_region(int pb, V3 u) {
    float4 a = SV<C>::pokeball(pb, 0), b = SV<C>::pokeball(pb, 1), c = SV<C>::pokeball(pb, 2);
    float cb = fminf(1.f, fmaxf(-1.f, fmaf(u.x, a.x, fmaf(u.y, a.y, u.z * a.z))));
    if (cb >= a.w) return (cb <= b.w) ? 3 : 4;
    float uy = fmaf(u.x, b.x, fmaf(u.y, b.y, u.z * b.z));
    if (fabsf(uy) <= c.x) return 2;
    return uy >= 0.f ? 0 : 1;
}

// world -> leaf-local ray for RT_FLAG_AFFINE leaves (Scaling::intersect, transform.cpp:98-108, composed
// with the rigid transforms around it).  Returns the stretch L = |A d| (world t = local t / L).
template <bool C> __device__ __forceinline__ float to_local(int xf, V3 o, V3 d, V3& ol, V3& dl) {
    float4 r0 = SV<C>::affine(xf, 0), r1 = SV<C>::affine(xf, 1), r2 = SV<C>::affine(xf, 2);
    ol = mk(fmaf(r0.x, o.x, fmaf(r0.y, o.y, fmaf(r0.z, o.z, r0.w))),
            fmaf(r1.x, o.x, fmaf(r1.y, o.y, fmaf(r1.z, o.z, r1.w))),
            fmaf(r2.x, o.x, fmaf(r2.y, o.y, fmaf(r2.z, o.z, r2.w))));
    V3 dd = mk(fmaf(r0.x, d.x, fmaf(r0.y, d.y, r0.z * d.z)),
               fmaf(r1.x, d.x, fmaf(r1.y, d.y, r1.z * d.z)),
               fmaf(r2.x, d.x, fmaf(r2.y, d.y, r2.z * d.z)));
    float L = sqrtf(dot(dd, dd));
    if (L > 1e-6f) { float inv = 1.f / L; dl = mk(dd.x * inv, dd.y * inv, dd.z * inv); return L; }
    dl = mk(0.f, 1.f, 0.f);                  // Ray ctor fallback (core.h:278, 95-101)
    return 1.f;
}
// local normal -> world: n_w = normalize(A^T n_l)  (transform.cpp:117-118 / 201-203 composed)
template <bool C> __device__ __forceinline__ V3 normal_to_world(int xf, V3 n) {
    float4 r0 = SV<C>::affine(xf, 0), r1 = SV<C>::affine(xf, 1), r2 = SV<C>::affine(xf, 2);
    return normalized(mk(fmaf(r0.x, n.x, fmaf(r1.x, n.y, r2.x * n.z)),
                         fmaf(r0.y, n.x, fmaf(r1.y, n.y, r2.y * n.z)),
                         fmaf(r0.z, n.x, fmaf(r1.z, n.y, r2.z * n.z))));
}

// Outward unit normal of a sphere-like leaf at world parameter t, plus the pokeball region.
template <bool C> __device__ __forceinline__ V3 sphere_outward(float4 g, int4 m, V3 o, V3 d, float t, int& region) {
    const int flags = (m.x >> 8);
    V3 out;
    if (flags & RT_FLAG_AFFINE) {
        V3 ol, dl; float L = to_local<C>(m.w, o, d, ol, dl);
        V3 pl = madd(ol, dl, t * L);
        float ir = 1.f / g.w;
        V3 ul = mk((pl.x - g.x) * ir, (pl.y - g.y) * ir, (pl.z - g.z) * ir);
        region = ((m.x & 0xff) == RT_PRIM_POKEBALL) ? pokeball_region<C>(m.z, ul) : 0;
        // orient in the local frame first (geometry.h:42-45), then map (the orientation commutes)
        V3 nl = dot(dl, ul) < 0.f ? ul : neg(ul);
        out = normal_to_world<C>(m.w, nl);
        return out;
    }
    V3 p = madd(o, d, t);
    float ir = 1.f / g.w;
    out = mk((p.x - g.x) * ir, (p.y - g.y) * ir, (p.z - g.z) * ir);
    region = ((m.x & 0xff) == RT_PRIM_POKEBALL) ? pokeball_region<C>(m.z, out) : 0;
    return out;
}

template <bool C> __device__ Ivl leaf_interval(int li, V3 o, V3 d, unsigned& tests) {
    Ivl r; r.ok = false; r.t0 = 0.f; r.t1 = 0.f; r.c0 = r.c1 = (uint32_t)li;
    float4 g = SV<C>::leaf_geom(li);
    int4 m = SV<C>::leaf_meta(li);
    const int kind = m.x & 0xff, flags = m.x >> 8;
    ++tests;
    if (flags & RT_FLAG_NEVER) return r;
    const bool xformed = flags & RT_FLAG_XFORMED;
    if (kind == RT_PRIM_HALFSPACE) {                       // HalfSpace::interval (geometry.cpp:117-147)
        float nd = fmaf(g.x, d.x, fmaf(g.y, d.y, g.z * d.z));
        float f0 = fmaf(g.x, o.x, fmaf(g.y, o.y, g.z * o.z)) - g.w;
        if (fabsf(nd) < 1e-12f) {
            if (!(f0 >= 0.f)) return r;
            r.t0 = -kInf; r.t1 = kInf;
        } else {
            float tp = -f0 / nd;
            if (nd > 0.f) { r.t0 = tp; r.t1 = kInf; } else { r.t0 = -kInf; r.t1 = tp; }
        }
        if (xformed) {                                     // transform.cpp:79-80: ends re-projected from p = ray origin
            if (!isfinite(r.t0)) r.t0 = 0.f;
            if (!isfinite(r.t1)) r.t1 = 0.f;
        }
        bool front = xformed ? (nd != 0.f) : (nd < 0.f);
        if (front) { r.c0 |= kFrontBit; r.c1 |= kFrontBit; }
        r.ok = true;
        return r;
    }
    // Sphere::interval / Pokeball::interval (geometry.cpp:48-78, 207-217)
    V3 ol = o, dl = d; float L = 1.f;
    if (flags & RT_FLAG_AFFINE) L = to_local<C>(m.w, o, d, ol, dl);
    float hb;
    float disc = sphere_disc(g, ol, dl, hb);
    if (disc < 0.f) return r;
    float s = sqrtf(disc);
    float ta = -hb - s, tb = -hb + s;
    float ir = 1.f / g.w;
    if (kind == RT_PRIM_POKEBALL) {
        V3 ua = mk((fmaf(dl.x, ta, ol.x) - g.x) * ir, (fmaf(dl.y, ta, ol.y) - g.y) * ir, (fmaf(dl.z, ta, ol.z) - g.z) * ir);
        V3 ub = mk((fmaf(dl.x, tb, ol.x) - g.x) * ir, (fmaf(dl.y, tb, ol.y) - g.y) * ir, (fmaf(dl.z, tb, ol.z) - g.z) * ir);
        r.c0 |= (uint32_t)pokeball_region<C>(m.z, ua) << 20;
        r.c1 |= (uint32_t)pokeball_region<C>(m.z, ub) << 20;
    }
    // front_face = d.outward < 0 with outward = (p - c)/r: at ta it is -s/r, at tb +s/r
    bool fa = xformed ? (s != 0.f) : (-s * ir < 0.f);
    bool fb = xformed ? (s != 0.f) : (s * ir < 0.f);
    if (fa) r.c0 |= kFrontBit;
    if (fb) r.c1 |= kFrontBit;
    float iL = 1.f / L;
    r.t0 = ta * iL; r.t1 = tb * iL;
    r.ok = true;
    return r;
}

// Out of line on purpose: inlined into both scans it cost 800 B of spills at 128 registers; as a call the kernel fits 64
// registers (4 CTAs = 32 warps per SM) with 1/4 of the spills (snorlax 4K 20.5 -> 17.1 ms, synthetic-B 547 -> 486 ms).
struct ProgResult { Ivl v; unsigned tests; };       // tests: primitives examined (returned by value: a reference would pin the caller's counter to memory)

// Candidates among the n <= 32 consecutive sphere-like entries [first, first + n) of a mask section for THIS ray: the
// round's candidate bits (the leaf section of bundle lb's words, or all of them for lb = -1), narrowed by the per-lane conservative mask over the mask
// table entries `masks` when many remain.  A superset of the entries whose exact discriminant is >= 0.
template <bool C> __device__ __forceinline__ uint32_t candidates(int lb, int first, const float4* masks, int n, V3 o, V3 d, unsigned& tests) {
    uint32_t cand = cand_bits(lb >= 0 ? cand_words(lb) + c_hdr.cull_leaf_word : nullptr, first, n);
    if (3 * __popc(cand) > n + 2) { tests += n; cand &= sphere_mask<C>(masks, n, o, d); }
    return cand;
}

// RT_OP_COMBINE_SPHERES: acc = (...(acc op s[first]) op s[first + 1]) ...) over n consecutive plain-sphere leaves.  OP >= 0:
// the operation is known at compile time (csg is ignored).
#ifndef RT_CSG_UNION_FOLD
#define RT_CSG_UNION_FOLD 1
#endif
#ifndef RT_FOLD_COLD
#define RT_FOLD_COLD 0           // A/B: 1 = the generic fold (intersection / difference) out of line
#endif
template <bool C, int OP>
__device__ __forceinline__ void fold_spheres(Ivl& acc, int csg_rt, int first, int n, V3 o, V3 d, int lb, unsigned& tests) {
    const int csg = OP >= 0 ? OP : csg_rt;
    for (int c0 = 0; c0 < n; c0 += 32) {
        const int cnt = min(32, n - c0), base = first + c0;
        // (leaf intervals live on the whole line, csg.cpp:61-163: the leaf bits were culled without a window)
        uint32_t mask = candidates<C>(lb, base, SV<C>::leaf_ptr(base), cnt, o, d, tests);
        int pos = 0;
        for (;;) {
            const int k = mask ? (__ffs(mask) - 1) : cnt;          // next candidate, or the end of the chunk
            float4 g = make_float4(0.f, 0.f, 0.f, 0.f); float hb = 0.f, disc = 0.f;
            if (k < cnt) {
                mask &= mask - 1; ++tests;
                g = SV<C>::leaf_geom(base + k);
                disc = sphere_disc(g, o, d, hb);
                if (disc < 0.f) continue;                          // the mask is conservative: k is a miss like pos..k-1
            }
            // ---- spheres pos..k-1 are missed: the one-child case of csg_combine, applied in order ----
            if (k > pos && acc.ok) {
                if (csg == RT_CSG_INTERSECTION) acc.ok = false;
                else if (isfinite(acc.t0) && isfinite(acc.t1) && acc.t0 <= acc.t1) origin_rewrite(acc);   // idempotent
                else {
                    Ivl none; none.ok = false; none.t0 = none.t1 = 0.f; none.c0 = none.c1 = 0;
                    for (int m = pos; m < k && acc.ok; ++m) {
                        if (isfinite(acc.t0) && isfinite(acc.t1) && acc.t0 <= acc.t1) origin_rewrite(acc);
                        else acc = csg_combine(csg, acc, none);
                    }
                }
            }
            if (k >= cnt) break;
            pos = k + 1;
            const float sq = sqrtf(disc), ir = 1.f / g.w;
            const bool xformed = (SV<C>::leaf_meta(base + k).x >> 8) & RT_FLAG_XFORMED;
            Ivl b; b.ok = true; b.t0 = -hb - sq; b.t1 = -hb + sq;
            b.c0 = b.c1 = (uint32_t)(base + k);
            if (xformed ? (sq != 0.f) : (-sq * ir < 0.f)) b.c0 |= kFrontBit;
            if (xformed ? (sq != 0.f) : (sq * ir < 0.f)) b.c1 |= kFrontBit;
            acc = csg_combine(csg, acc, b);
        }
    }
}

template <bool C>
__device__ __noinline__ void fold_spheres_cold(Ivl& acc, int csg, int first, int n, V3 o, V3 d, int lb, unsigned& tests) {
    fold_spheres<C, -1>(acc, csg, first, n, o, d, lb, tests);
}

// `lb`: the bundle whose mask holds the round's candidate bits of the CSG leaves, or -1.
template <bool C> __device__ __noinline__ ProgResult run_program(int first, int count, V3 o, V3 d, int lb) {
    unsigned tests = 0;
    Ivl acc; acc.ok = false; acc.t0 = acc.t1 = 0.f; acc.c0 = acc.c1 = 0;
    Ivl stack[RT_MAX_CSG_STACK]; int sp = 0;
    for (int i = 0; i < count; ++i) {
        const uint32_t w = SV<C>::op(first + i);
        const int opcode = w & 15, csg = (w >> 4) & 3; const uint32_t arg = w >> 8;
        switch (opcode) {
            case RT_OP_LEAF: acc = leaf_interval<C>((int)arg, o, d, tests); break;
            case RT_OP_COMBINE_LEAF: { Ivl b = leaf_interval<C>((int)arg, o, d, tests); acc = csg_combine(csg, acc, b); } break;
            case RT_OP_PUSH: stack[sp++] = acc; break;
            case RT_OP_COMBINE_POP: { --sp; acc = csg_combine(csg, stack[sp], acc); } break;
            case RT_OP_XFIX: xfix(acc, arg); break;
            case RT_OP_COMBINE_SPHERES: {            // acc = (...(acc op s[first]) op s[first+1]) ... ) over plain spheres
                const int first = (int)(arg & 0x3FFFu), n = (int)(arg >> 14);
                // (the union fold — a model built from spheres — gets its own copy with the operation known at compile time:
                // csg_combine's case analysis collapses to a third)
                if (RT_CSG_UNION_FOLD && csg == RT_CSG_UNION) fold_spheres<C, RT_CSG_UNION>(acc, csg, first, n, o, d, lb, tests);
                else if (RT_FOLD_COLD) fold_spheres_cold<C>(acc, csg, first, n, o, d, lb, tests);
                else fold_spheres<C, -1>(acc, csg, first, n, o, d, lb, tests);
            } break;
        }
    }
    ProgResult r; r.v = acc; r.tests = tests;
    return r;
}

}  // namespace rtd
