// rt_cull.cuh — conservative culling of sphere-like primitives against a BUNDLE of rays.
//
// The 256 camera rays a CTA traces in one round (32 neighbouring pixels x 8 samples), and their shadow rays towards one
// light, are nearly parallel and start close together.  Instead of every thread testing every primitive of the scene for
// its own ray (Scene::intersect / Scene::occluded are linear scans, scene.cpp:10-42), the CTA first tests the primitives
// against the bundle as a whole, ONE PRIMITIVE PER THREAD, into a candidate bitmask in shared memory, and every thread
// then runs the reference's exact arithmetic only on the few primitives the bundle can touch.  Results are identical to
// the linear scan as long as the bundle test never rejects a primitive that the exact FP32 test of some ray would accept;
// this header holds that test as pure arithmetic (no scene access, no CUDA intrinsics) so that
// tests/host/cull_conservative.cpp can check exactly this property on the host.
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

// Camera rays of one round: one origin (the eye), directions inside the box [lo, hi] (component-wise min / max over the
// rays).  The central direction is the normalised box centre; a ray direction differs from it by at most the half
// diagonal plus the centre's distance from the unit sphere.
RT_CULL_FN void bundle_from_dirbox(Bundle& b, const float eye[3], const float lo[3], const float hi[3]) {
    const float qx = 0.5f * (lo[0] + hi[0]), qy = 0.5f * (lo[1] + hi[1]), qz = 0.5f * (lo[2] + hi[2]);
    const float ex = 0.5f * (hi[0] - lo[0]), ey = 0.5f * (hi[1] - lo[1]), ez = 0.5f * (hi[2] - lo[2]);
    const float n = sqrtf(fmaf(qx, qx, fmaf(qy, qy, qz * qz)));
    const float inv = 1.0f / n;
    b.mx = eye[0]; b.my = eye[1]; b.mz = eye[2];
    b.cx = qx * inv; b.cy = qy * inv; b.cz = qz * inv;
    const float spread = sqrtf(fmaf(ex, ex, fmaf(ey, ey, ez * ez))) + fabsf(1.0f - n);
    bundle_finish(b, 0.f, spread * spread, INFINITY);      // (NaN / a zero centre give delta = NaN: wide)
}

// Shadow rays of one round towards one light: origins inside the box [lo, hi], every ray aimed from a point within
// eps_max of its origin (the hit point; the origin is offset by eps along the normal, shading.cpp:43,53) at the light.
// Point light at L: with g = h + eps_max every aim vector is within g of L - m, so the unit directions are within
// 2 g / |L - m| of the central one, and no ray is longer than |L - m| + g.  Directional light: one direction, no end.
RT_CULL_FN void bundle_from_box_light(Bundle& b, const float lo[3], const float hi[3], float eps_max, const float light[3], bool directional) {
    b.mx = 0.5f * (lo[0] + hi[0]); b.my = 0.5f * (lo[1] + hi[1]); b.mz = 0.5f * (lo[2] + hi[2]);
    const float ex = 0.5f * (hi[0] - lo[0]), ey = 0.5f * (hi[1] - lo[1]), ez = 0.5f * (hi[2] - lo[2]);
    // (the half extents are rounded: one ulp of the coordinates on top of the exact half diagonal)
    const float h2 = fmaf(ex, ex, fmaf(ey, ey, ez * ez));
    if (directional) {
        b.cx = light[0]; b.cy = light[1]; b.cz = light[2];
        bundle_finish(b, h2, 0.f, INFINITY);
        return;
    }
    const float ux = light[0] - b.mx, uy = light[1] - b.my, uz = light[2] - b.mz;
    const float D = sqrtf(fmaf(ux, ux, fmaf(uy, uy, uz * uz)));
    const float inv = 1.0f / D;
    b.cx = ux * inv; b.cy = uy * inv; b.cz = uz * inv;
    const float g = sqrtf(h2) * 1.0001f + eps_max + 1e-6f * (fabsf(b.mx) + fabsf(b.my) + fabsf(b.mz) + fabsf(light[0]) + fabsf(light[1]) + fabsf(light[2])) + 1e-7f;
    const float delta = 2.0f * g * inv * 1.001f + 1e-6f;
    bundle_finish(b, h2, delta * delta, fmaxf(D + g, 0.1f));      // (shading.cpp:86-92: the distance is clamped to 0.1 from below)
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

// Half-space surface (unit normal n, offset w = n.p0; HalfSpace::intersect, geometry.cpp:90-106): a ray hits it at
// t = (w - n.o) / (n.d) > 0 only if it starts below the plane and runs upwards, or starts above and runs downwards.  Rejected
// when every origin of the bundle is on one side and every direction leads away from the plane (or along it: |n.d| < 1e-12
// is a miss as well).  The far end of the window is not used.
RT_CULL_FN bool bundle_rejects_plane(const Bundle& b, Sphere4 g) {
    const float no = fmaf(g.x, b.mx, fmaf(g.y, b.my, g.z * b.mz)) - g.w;          // signed distance of the central origin
    const float nd = fmaf(g.x, b.cx, fmaf(g.y, b.cy, g.z * b.cz));
    // b.kd = 1.002 delta + 1e-5 >= |n.(d_i - c)| + rounding;  b.h >= |o_i - m| incl. its own rounding margin
    const float slack = b.h + 1e-5f * (fabsf(no) + fabsf(g.w)) + 1e-6f;
    const bool above = no - slack > 0.f, below = no + slack < 0.f;
    const bool up = nd - b.kd > 0.f, down = nd + b.kd < 0.f;
    return (above && up) || (below && down);
}

}  // namespace rtd
