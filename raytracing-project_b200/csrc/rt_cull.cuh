// rt_cull.cuh — conservative culling of sphere-like primitives against a BUNDLE of rays.
//
// The 32 rays a warp traces together (4 neighbouring pixels x 8 samples, or their shadow rays towards one light) are
// nearly parallel and start close together.  Instead of every lane testing every primitive of the scene for its own ray
// (Scene::intersect / Scene::occluded are linear scans, scene.cpp:10-42), the warp first tests the primitives against the
// bundle as a whole, ONE PRIMITIVE PER LANE, and every lane then runs the reference's exact arithmetic only on the few
// primitives the bundle can touch.  Results are identical to the linear scan as long as the bundle test never rejects a
// primitive that the exact FP32 test of some lane would accept; this header holds that test as pure arithmetic (no scene
// access, no CUDA intrinsics) so that tests/host/cull_conservative.cpp can check exactly this property on the host.
//
// Bundle = central ray (m, c) with |c| = 1 up to rounding, origin spread h >= max_i |o_i - m|, direction spread
// delta >= max_i |d_i - c|.  A point of ray i at parameter t lies within h + |t| delta of the central point m + t c.
// If ray i touches the sphere (q, r) at parameter t, then with v = q - m, s = v.c, perp = |v - s c|:
//     |t - s| <= r + h + |t| delta   =>   |t| <= T := (|s| + r + h) / (1 - delta)
//     perp    <= r + h + T delta =: R,    t in [s - R, s + R]
// so the primitive is rejected iff perp > R (the LINE of no ray meets it: what CSG leaves need, their intervals live on the
// whole line, csg.cpp:61-163), or, for top-level objects whose hit must lie in (0, t_hi], iff s + R < 0 or s - R > t_hi.
// R carries a relative margin of 2e-3 and absolute margins of 1e-5 |s| and ~1e-6 of the scene coordinates: the FP32
// rounding of this test and of the exact test (sphere_disc: ~ 2^-22 |o - q|) are two orders of magnitude below that.
// Every comparison is written so that NaN and infinite radii (entries that must not be culled) come out as candidates.
#pragma once
#include <math.h>
#include <stdint.h>

#ifdef __CUDACC__
#define RT_CULL_FN __host__ __device__ __forceinline__
#else
#define RT_CULL_FN inline
#endif

namespace rtd {

struct Bundle {
    float mx, my, mz, h;          // central origin, origin spread (inflated)
    float cx, cy, cz, kd;         // central direction, 1.002 * delta + 1e-5
    float inv1md, t_hi;           // 1 / (1 - delta); far end of the widest query window in the bundle
    int wide, any;                // wide: delta >= 0.5, nothing is culled; any: at least one ray
};

struct Sphere4 { float x, y, z, w; };     // (centre, radius) — float4 on the device

constexpr float kCullRel = 1.002f;
constexpr float kCullWide = 0.5f;

// h2 = max_i |o_i - m|^2 and d2 = max_i |d_i - c|^2 as computed in FP32 by the lanes (m, c = one lane's own ray).
RT_CULL_FN void bundle_finish(Bundle& b, float h2, float d2, float t_hi) {
    // o_i - m is rounded like the exact test's o - q: an absolute error ~ 2^-24 of the coordinates themselves
    b.h = sqrtf(h2) * 1.0001f + 1e-6f * (fabsf(b.mx) + fabsf(b.my) + fabsf(b.mz)) + 1e-7f;
    const float delta = sqrtf(d2) * 1.0001f + 3e-7f;
    b.wide = !(delta < kCullWide);                       // NaN -> wide
    b.kd = kCullRel * delta + 1e-5f;
    b.inv1md = 1.0f / (1.0f - fminf(delta, kCullWide));
    b.t_hi = t_hi;
}

// true: no ray of the bundle can touch sphere g (segment: ... inside (0, t_hi]).  false: candidate.
RT_CULL_FN bool bundle_rejects(const Bundle& b, Sphere4 g, float t_hi, bool segment) {
    const float vx = g.x - b.mx, vy = g.y - b.my, vz = g.z - b.mz;
    const float s = fmaf(vx, b.cx, fmaf(vy, b.cy, vz * b.cz));
    const float lx = fmaf(-s, b.cx, vx), ly = fmaf(-s, b.cy, vy), lz = fmaf(-s, b.cz, vz);
    const float p2 = fmaf(lx, lx, fmaf(ly, ly, lz * lz));
    const float a = fabsf(g.w) + b.h;
    const float T = (fabsf(s) + a) * b.inv1md;
    const float R = fmaf(b.kd, T, kCullRel * a) + 1e-7f;
    bool rej = p2 > R * R;
    if (segment) rej = rej || (s + R < 0.f) || (s - R > t_hi);
    return rej;
}

}  // namespace rtd
