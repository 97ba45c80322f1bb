// Host-side check that the bundle cull (csrc/rt_cull.cuh) is CONSERVATIVE: it never rejects a sphere that the device's
// exact FP32 test (sphere_disc / sphere_t of csrc/rt_device.cuh, restated below with the same fmaf sequence) accepts for
// some ray of the bundle.  Random bundles at several scene scales and spreads, built the three ways the device builds them
// (central ray + spreads; camera rays = eye + box of directions; shadow rays = box of origins + light); spheres are placed
// near-tangent to a ray of the bundle (both sides of the silhouette, relative gaps 1e-7 ... 1e-2), near the ends of the
// query window, and at random.
// Prints "<cases> <rejected> <violations>"; test infrastructure only (tests/test_cull.py builds and runs it).
#include <cmath>
#include <cstdio>
#include <random>
#include <vector>

#include "../../raytracing-project_b200/csrc/rt_cull.cuh"

struct V { float x, y, z; };
static float dotf(V a, V b) { return fmaf(a.x, b.x, fmaf(a.y, b.y, a.z * b.z)); }
// sphere_disc (rt_device.cuh): r^2 - |oc - (oc.d) d|^2
static float sphere_disc(rtd::Sphere4 g, V o, V d, float& hb) {
    V oc{o.x - g.x, o.y - g.y, o.z - g.z};
    hb = dotf(oc, d);
    V l{fmaf(-hb, d.x, oc.x), fmaf(-hb, d.y, oc.y), fmaf(-hb, d.z, oc.z)};
    return fmaf(-l.x, l.x, fmaf(-l.y, l.y, fmaf(-l.z, l.z, g.w * g.w)));
}

int main(int argc, char** argv) {
    const long long trials = argc > 1 ? atoll(argv[1]) : 180000;
    std::mt19937_64 rng(20261105);
    std::uniform_real_distribution<double> U(0.0, 1.0);
    std::normal_distribution<double> N(0.0, 1.0);
    auto unit = [&]() { double x = N(rng), y = N(rng), z = N(rng), L = std::sqrt(x * x + y * y + z * z); return std::vector<double>{x / L, y / L, z / L}; };
    const double scales[] = {1.0, 20.0, 300.0, 5000.0};
    const double hs[] = {0.0, 1e-4, 1e-2, 0.3, 3.0};
    const double deltas[] = {0.0, 1e-5, 1e-3, 0.03, 0.2, 0.45};
    const double gaps[] = {-1e-2, -1e-4, -1e-6, -1e-7, 0.0, 1e-7, 1e-6, 1e-4, 1e-2, 0.3, 3.0};
    long long cases = 0, rejected = 0, bad = 0;
    for (long long tr = 0; tr < trials; ++tr) {
        const double S = scales[tr % 4], hh = hs[(tr / 4) % 5] * (S > 100 ? 10 : 1), dl = deltas[(tr / 20) % 6];
        V o[32], d[32]; float tmax[32]; bool act[32];
        auto c0 = unit();
        const double m0[3] = {(2 * U(rng) - 1) * S, (2 * U(rng) - 1) * S, (2 * U(rng) - 1) * S};
        const bool finite_window = U(rng) < 0.5;
        for (int i = 0; i < 32; ++i) {
            auto u = unit(); const double ro = (i == 16 ? 0.0 : hh * std::cbrt(U(rng)));
            o[i] = {(float)(m0[0] + ro * u[0]), (float)(m0[1] + ro * u[1]), (float)(m0[2] + ro * u[2])};
            auto w = unit(); const double rd = (i == 16 ? 0.0 : dl * U(rng));
            double x = c0[0] + rd * w[0], y = c0[1] + rd * w[1], z = c0[2] + rd * w[2], L = std::sqrt(x * x + y * y + z * z);
            const double sc = (1.0 + (2 * U(rng) - 1) * std::ldexp(1.0, -21)) / L;          // 2-ulp reciprocal normalisation
            d[i] = {(float)(x * sc), (float)(y * sc), (float)(z * sc)};
            tmax[i] = finite_window ? (float)(S * (0.05 + 3 * U(rng))) : INFINITY;
            act[i] = U(rng) < 0.9 || i == 16;
        }
        rtd::Bundle B{};
        const int mode = (int)((tr / 120) % 3);
        if (mode == 0) {
            // central ray + spreads: ray 16 is the central ray
            B.mx = o[16].x; B.my = o[16].y; B.mz = o[16].z; B.cx = d[16].x; B.cy = d[16].y; B.cz = d[16].z;
            float h2 = 0.f, d2 = 0.f, th = -INFINITY;
            for (int i = 0; i < 32; ++i) if (act[i]) {
                const float ax = o[i].x - B.mx, ay = o[i].y - B.my, az = o[i].z - B.mz, bx = d[i].x - B.cx, by = d[i].y - B.cy, bz = d[i].z - B.cz;
                h2 = fmaxf(h2, fmaf(ax, ax, fmaf(ay, ay, az * az)));
                d2 = fmaxf(d2, fmaf(bx, bx, fmaf(by, by, bz * bz)));
                th = fmaxf(th, tmax[i]);
            }
            rtd::bundle_finish(B, h2, d2, th);
        } else if (mode == 1) {
            // camera rays (rt_device.cuh cand_build, bundle 0): one origin, box of the directions, no far end
            float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
            for (int i = 0; i < 32; ++i) { o[i] = o[16]; tmax[i] = INFINITY; }
            for (int i = 0; i < 32; ++i) if (act[i]) {
                lo[0] = fminf(lo[0], d[i].x); lo[1] = fminf(lo[1], d[i].y); lo[2] = fminf(lo[2], d[i].z);
                hi[0] = fmaxf(hi[0], d[i].x); hi[1] = fmaxf(hi[1], d[i].y); hi[2] = fmaxf(hi[2], d[i].z);
            }
            const float eye[3] = {o[16].x, o[16].y, o[16].z};
            rtd::bundle_from_dirbox(B, eye, lo, hi);
        } else {
            // shadow rays (cand_build, bundle 1 + l): origins = hit point + eps * normal in a box, aimed from the hit point at a light
            const bool directional = U(rng) < 0.25;
            auto lu = unit(); const double ld = S * (0.02 + 3 * U(rng));
            const float L[3] = {directional ? (float)lu[0] : (float)(m0[0] + ld * lu[0]), directional ? (float)lu[1] : (float)(m0[1] + ld * lu[1]),
                                directional ? (float)lu[2] : (float)(m0[2] + ld * lu[2])};
            float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY}, eps_max = 0.f;
            for (int i = 0; i < 32; ++i) {
                const float eps = fmaxf(1e-3f, 1e-4f * (float)(S * 3 * U(rng)));
                auto nn = unit();
                const V p{(float)(o[i].x - eps * nn[0]), (float)(o[i].y - eps * nn[1]), (float)(o[i].z - eps * nn[2])};   // the hit point
                if (directional) { d[i] = {L[0], L[1], L[2]}; tmax[i] = INFINITY; }
                else {                                             // light_vector + normalized (rt_trace.cuh) in FP32
                    V a{L[0] - p.x, L[1] - p.y, L[2] - p.z};
                    const float dist = sqrtf(fmaxf(dotf(a, a), 0.01f)), inv = 1.0f / dist;
                    V wi{a.x * inv, a.y * inv, a.z * inv};
                    const float l2 = sqrtf(dotf(wi, wi)), i2 = 1.0f / l2;
                    d[i] = {wi.x * i2, wi.y * i2, wi.z * i2};
                    tmax[i] = dist - eps;
                }
                if (act[i]) {
                    lo[0] = fminf(lo[0], o[i].x); lo[1] = fminf(lo[1], o[i].y); lo[2] = fminf(lo[2], o[i].z);
                    hi[0] = fmaxf(hi[0], o[i].x); hi[1] = fmaxf(hi[1], o[i].y); hi[2] = fmaxf(hi[2], o[i].z);
                    eps_max = fmaxf(eps_max, eps);
                }
            }
            rtd::bundle_from_box_light(B, lo, hi, eps_max, L, directional);
        }
        if (B.wide) continue;
        for (int k = 0; k < 24; ++k) {
            // a sphere near-tangent to ray `i` at parameter t
            const int i = (int)(U(rng) * 32) & 31;
            const double r = std::exp(std::log(1e-3) + U(rng) * std::log(1e6)) * (S > 100 ? S / 100 : 1.0);
            double t = (U(rng) < 0.2 ? -1.0 : 1.0) * S * 3 * U(rng);
            if (finite_window && U(rng) < 0.3) t = tmax[i] + (2 * U(rng) - 1) * r * 1.5;          // around the far end of the window
            if (U(rng) < 0.15) t = (2 * U(rng) - 1) * r * 1.5;                                     // around the origin
            auto u = unit();
            const double dd[3] = {d[i].x, d[i].y, d[i].z};
            const double ud = u[0] * dd[0] + u[1] * dd[1] + u[2] * dd[2];
            double px = u[0] - ud * dd[0], py = u[1] - ud * dd[1], pz = u[2] - ud * dd[2], pl = std::sqrt(px * px + py * py + pz * pz);
            const double gap = gaps[(int)(U(rng) * 11) % 11] * U(rng);
            const double dist = (k % 6 == 5) ? r * 10 * U(rng) : r * (1.0 + gap);
            rtd::Sphere4 g{(float)(o[i].x + t * dd[0] + px / pl * dist), (float)(o[i].y + t * dd[1] + py / pl * dist),
                           (float)(o[i].z + t * dd[2] + pz / pl * dist), (float)r};
            for (int segment = 0; segment < 2; ++segment) {
                ++cases;
                if (!rtd::bundle_rejects(B, g, B.t_hi, segment != 0)) continue;
                ++rejected;
                for (int j = 0; j < 32; ++j) if (act[j]) {
                    float hb; const float disc = sphere_disc(g, o[j], d[j], hb);
                    if (disc < 0.f) continue;                       // the exact test misses too
                    bool accepted = true;
                    if (segment) {                                  // sphere_t's window test on either root (tmin > 0)
                        const float s = sqrtf(disc), t1 = -hb - s, t2 = -hb + s;
                        accepted = (t1 >= 0.f && t1 <= tmax[j]) || (t2 >= 0.f && t2 <= tmax[j]);
                    }
                    if (accepted && ++bad <= 10)
                        std::printf("# violation: scale %g h %g delta %g segment %d lane %d disc %g r %g\n", S, hh, dl, segment, j, (double)disc, r);
                }
            }
        }
    }
    // ---- half-spaces: plane_t (rt_device.cuh) restated; planes through / near the bundle's origins, normals near-orthogonal to the rays
    long long pcases = 0, prej = 0;
    for (long long tr = 0; tr < trials / 4; ++tr) {
        const double S = scales[tr % 4], hh = hs[(tr / 4) % 5] * (S > 100 ? 10 : 1), dl = deltas[(tr / 20) % 6];
        V o[32], d[32]; float tmax[32];
        auto c0 = unit();
        const double m0[3] = {(2 * U(rng) - 1) * S, (2 * U(rng) - 1) * S, (2 * U(rng) - 1) * S};
        for (int i = 0; i < 32; ++i) {
            auto u = unit(); const double ro = (i == 16 ? 0.0 : hh * std::cbrt(U(rng)));
            o[i] = {(float)(m0[0] + ro * u[0]), (float)(m0[1] + ro * u[1]), (float)(m0[2] + ro * u[2])};
            auto w = unit(); const double rd = (i == 16 ? 0.0 : dl * U(rng));
            double x = c0[0] + rd * w[0], y = c0[1] + rd * w[1], z = c0[2] + rd * w[2], L = std::sqrt(x * x + y * y + z * z);
            d[i] = {(float)(x / L), (float)(y / L), (float)(z / L)};
            tmax[i] = INFINITY;
        }
        rtd::Bundle B{};
        B.mx = o[16].x; B.my = o[16].y; B.mz = o[16].z; B.cx = d[16].x; B.cy = d[16].y; B.cz = d[16].z;
        float h2 = 0.f, d2 = 0.f;
        for (int i = 0; i < 32; ++i) {
            const float ax = o[i].x - B.mx, ay = o[i].y - B.my, az = o[i].z - B.mz, bx = d[i].x - B.cx, by = d[i].y - B.cy, bz = d[i].z - B.cz;
            h2 = fmaxf(h2, fmaf(ax, ax, fmaf(ay, ay, az * az))); d2 = fmaxf(d2, fmaf(bx, bx, fmaf(by, by, bz * bz)));
        }
        rtd::bundle_finish(B, h2, d2, INFINITY);
        if (B.wide) continue;
        for (int k = 0; k < 16; ++k) {
            auto n = unit();
            if (k % 2) {                                             // nearly orthogonal to the central direction
                const double nc = n[0] * c0[0] + n[1] * c0[1] + n[2] * c0[2], a = (k % 4 == 1 ? 1.0 : 1e-3) * (2 * U(rng) - 1) * (dl + 1e-6) * 2;
                double x = n[0] - (nc - a) * c0[0], y = n[1] - (nc - a) * c0[1], z = n[2] - (nc - a) * c0[2], L = std::sqrt(x * x + y * y + z * z);
                n = {x / L, y / L, z / L};
            }
            const int i = (int)(U(rng) * 32) & 31;
            const double off = (k % 3 == 0 ? (2 * U(rng) - 1) * (hh + 1e-4) * 2 : (2 * U(rng) - 1) * S);    // plane near an origin, or anywhere
            rtd::Sphere4 g{(float)n[0], (float)n[1], (float)n[2], 0.f};
            g.w = (float)(g.x * (double)o[i].x + g.y * (double)o[i].y + g.z * (double)o[i].z + off);
            ++pcases;
            if (!rtd::bundle_rejects_plane(B, g)) continue;
            ++prej;
            for (int j = 0; j < 32; ++j) {                            // plane_t with tmin = 1e-4, tmax = inf
                const float nd = fmaf(g.x, d[j].x, fmaf(g.y, d[j].y, g.z * d[j].z));
                if (fabsf(nd) < 1e-12f) continue;
                const float no = fmaf(g.x, o[j].x, fmaf(g.y, o[j].y, g.z * o[j].z));
                const float tt = (g.w - no) / nd;
                if (!(tt < 1e-4f || tt > tmax[j]) && ++bad <= 10) std::printf("# plane violation: scale %g h %g delta %g lane %d t %g\n", S, hh, dl, j, (double)tt);
            }
        }
    }
    cases += pcases; rejected += prej;
    std::printf("# planes: %lld cases, %lld rejected\n", pcases, prej);
    std::printf("%lld %lld %lld\n", cases, rejected, bad);
    return bad ? 1 : 0;
}
