// Host-side check that the uniform grid (csrc/rt_grid.cuh) is CONSERVATIVE: for random sphere sets (lattices like the
// glass scene, random clouds, mixed radii, several scales) and random segments — origins inside and outside the grid,
// random / axis-parallel / near-tangent directions, finite and infinite windows — every sphere that the device's exact FP32
// test (sphere_disc / sphere_t of csrc/rt_device.cuh, restated below with the same fmaf sequence) accepts inside the window
// has its bit set in the words grid_walk returns, and the front-to-back walk with early exit finds the same closest hit (and
// the same any-hit answer) as the sequential scan.  Prints "<rays> <accepted sphere hits> <mean cells per ray x 100>
// <mean candidate bits per ray x 100> <violations>"; test infrastructure only (tests/test_grid.py builds and runs it).
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <random>
#include <vector>

#include "../../raytracing-project_b200/csrc/rt_grid.cuh"

struct V { float x, y, z; };
static float dotf(V a, V b) { return fmaf(a.x, b.x, fmaf(a.y, b.y, a.z * b.z)); }
static float sphere_disc(const float* g, V o, V d, float& hb) {
    V oc{o.x - g[0], o.y - g[1], o.z - g[2]};
    hb = dotf(oc, d);
    V l{fmaf(-hb, d.x, oc.x), fmaf(-hb, d.y, oc.y), fmaf(-hb, d.z, oc.z)};
    return fmaf(-l.x, l.x, fmaf(-l.y, l.y, fmaf(-l.z, l.z, g[3] * g[3])));
}
static bool sphere_t(const float* g, V o, V d, float tmin, float tmax, float& t) {
    float hb; const float disc = sphere_disc(g, o, d, hb);
    if (disc < 0.f) return false;
    const float s = sqrtf(disc);
    float tt = -hb - s;
    if (tt < tmin || tt > tmax) { tt = -hb + s; if (tt < tmin || tt > tmax) return false; }
    t = tt; return true;
}

int main(int argc, char** argv) {
    const int scenes = argc > 1 ? atoi(argv[1]) : 60, rays_per_scene = argc > 2 ? atoi(argv[2]) : 20000;
    std::mt19937_64 rng(20261105);
    std::uniform_real_distribution<double> U(0.0, 1.0);
    std::normal_distribution<double> N(0.0, 1.0);
    long long rays = 0, hits = 0, cells = 0, bits = 0, bad = 0;
    for (int sc = 0; sc < scenes; ++sc) {
        const double S = (sc % 4 == 0) ? 1.0 : (sc % 4 == 1) ? 10.0 : (sc % 4 == 2) ? 300.0 : 0.05;
        const double off[3] = {(2 * U(rng) - 1) * 5 * S, (2 * U(rng) - 1) * 5 * S, (2 * U(rng) - 1) * 5 * S};
        std::vector<float> sph; std::vector<int> bit;
        const int n = 24 + (int)(U(rng) * 100);
        if (sc % 3 == 0) {                                    // a lattice (+ a few big spheres), like the glass scene
            const int nx = 3 + (int)(U(rng) * 3), ny = 3 + (int)(U(rng) * 3);
            for (int i = 0; i < n; ++i) {
                const int ix = i % nx, iy = (i / nx) % ny, iz = i / (nx * ny);
                const double r = (i % 17 == 0) ? 0.8 * S : 0.33 * S;
                sph.insert(sph.end(), {(float)(off[0] + 0.8 * S * ix), (float)(off[1] + 0.8 * S * iy), (float)(off[2] - 0.9 * S * iz), (float)r});
            }
        } else {
            for (int i = 0; i < n; ++i) {
                const double r = S * (sc % 3 == 1 ? 0.05 + 0.4 * U(rng) : std::exp(N(rng)) * 0.2);
                sph.insert(sph.end(), {(float)(off[0] + S * 4 * (U(rng) - 0.5)), (float)(off[1] + S * 4 * (U(rng) - 0.5)), (float)(off[2] + S * 4 * (U(rng) - 0.5)), (float)r});
            }
        }
        if (sc % 5 == 1) { const double R = 1000 * S; sph[0] = (float)off[0]; sph[1] = (float)(off[1] - R - 2 * S); sph[2] = (float)off[2]; sph[3] = (float)R; }   // a "floor" sphere
        const int shift = sc % 4;
        for (int i = 0; i < n; ++i) bit.push_back(i + shift);                     // device indices need not start at 0
        rtd::GridHeader g; std::vector<uint32_t> masks;
        if (!rtd::grid_build(sph.data(), bit.data(), n, n + shift, g, masks)) { printf("grid_build failed\n"); return 2; }
        for (int r = 0; r < rays_per_scene; ++r) {
            V o, d;
            const int mode = r % 6;
            // origin: far outside, near the box, or on a sphere (a secondary / shadow ray)
            const int k = (int)(U(rng) * n);
            const float* s = sph.data() + 4 * k;
            auto unit = [&]() { double x = N(rng), y = N(rng), z = N(rng), L = std::sqrt(x * x + y * y + z * z); return V{(float)(x / L), (float)(y / L), (float)(z / L)}; };
            if (mode == 0) { V u = unit(); o = {(float)(off[0] + 30 * S * u.x), (float)(off[1] + 30 * S * u.y), (float)(off[2] + 30 * S * u.z)}; }
            else if (mode == 1) o = {(float)(off[0] + S * 6 * (U(rng) - 0.5)), (float)(off[1] + S * 6 * (U(rng) - 0.5)), (float)(off[2] + S * 6 * (U(rng) - 0.5))};
            else { V u = unit(); const double rr = std::fabs(s[3]) * (1.0 + (mode == 2 ? 1e-3 : mode == 3 ? -1e-3 : 1e-6)); o = {(float)(s[0] + rr * u.x), (float)(s[1] + rr * u.y), (float)(s[2] + rr * u.z)}; }
            // direction: random, axis-parallel, in a coordinate plane, or aimed (near-)tangent at another sphere
            const int dm = (r / 6) % 5;
            if (dm == 0) d = unit();
            else if (dm == 1) { d = {0.f, 0.f, 0.f}; const int a = (int)(U(rng) * 3); (a == 0 ? d.x : a == 1 ? d.y : d.z) = U(rng) < 0.5 ? 1.f : -1.f; }
            else if (dm == 2) { V u = unit(); d = {u.x, u.y, 0.f}; const float L = sqrtf(dotf(d, d)); if (L < 1e-3f) d = {1.f, 0.f, 0.f}; else d = {d.x / L, d.y / L, 0.f}; }
            else {
                const float* q = sph.data() + 4 * (int)(U(rng) * n);
                V u = unit();
                const double gap = (dm == 3 ? 1.0 + (U(rng) - 0.5) * 2e-5 : U(rng) * 1.2), rr = std::fabs(q[3]) * gap;
                double x = q[0] + rr * u.x - o.x, y = q[1] + rr * u.y - o.y, z = q[2] + rr * u.z - o.z, L = std::sqrt(x * x + y * y + z * z);
                if (L < 1e-9) { d = unit(); } else { const double sc2 = (1.0 + (2 * U(rng) - 1) * std::ldexp(1.0, -21)) / L; d = {(float)(x * sc2), (float)(y * sc2), (float)(z * sc2)}; }
            }
            const float tmin = (r % 2) ? 1e-4f : 1e-3f;
            const float tmax = (r % 3 == 0) ? INFINITY : (float)(S * (0.02 + 8 * U(rng)));
            rtd::GridWords gwd{0, 0, 0, 0};
            const float oo[3] = {o.x, o.y, o.z}, dd[3] = {d.x, d.y, d.z};
            const uint32_t* mp = masks.data();
            const int visited = rtd::grid_walk(g, [mp](int c) { return rtd::GridWords{mp[4 * c], mp[4 * c + 1], mp[4 * c + 2], mp[4 * c + 3]}; }, oo, dd, tmin, tmax, gwd);
            const uint32_t w[rtd::kGridMaxWords] = {gwd.x, gwd.y, gwd.z, gwd.w};
            ++rays; cells += visited > 0 ? visited : 0;
            // closest hit: the sequential scan of scene.cpp:10-24 (index order, inclusive shrinking window: the later index
            // wins an exact tie) against the front-to-back walk with early exit (rt_trace.cuh scan_body), and any hit
            if (visited >= 0) {
                int sb = -1; float st = tmax;
                for (int i = 0; i < n; ++i) { float t; if (sphere_t(sph.data() + 4 * i, o, d, tmin, st, t)) { sb = i; st = t; } }
                for (int any = 0; any < 2; ++any) {
                    int wb = -1; float wt = tmax; uint32_t tested[4] = {0, 0, 0, 0};
                    rtd::grid_visit(g, [mp](int c) { return rtd::GridWords{mp[4 * c], mp[4 * c + 1], mp[4 * c + 2], mp[4 * c + 3]}; }, oo, dd, tmin, tmax,
                        [&](const rtd::GridWords& m, float& t1) {
                            const uint32_t mw[4] = {m.x, m.y, m.z, m.w};
                            for (int w2 = 0; w2 < g.words; ++w2) {
                                uint32_t b2 = mw[w2] & ~tested[w2];
                                while (b2) {
                                    const int dev = 32 * w2 + __builtin_ctz(b2); b2 &= b2 - 1;
                                    const int i = dev - shift;
                                    float t;
                                    if (i < 0 || i >= n || !sphere_t(sph.data() + 4 * i, o, d, tmin, wt, t)) continue;
                                    if (t == wt && i < wb) continue;
                                    wb = i; wt = t;
                                }
                                tested[w2] |= mw[w2];
                            }
                            t1 = (any && wb >= 0) ? -INFINITY : fminf(t1, wt);
                        });
                    const bool ok = any ? ((wb >= 0) == (sb >= 0)) : (wb == sb && (sb < 0 || wt == st));
                    if (!ok && ++bad <= 5) fprintf(stderr, "CLOSEST scene %d ray %d any %d: scan %d %g walk %d %g\n", sc, r, any, sb, st, wb, wt);
                }
            }
            for (int ww = 0; ww < g.words; ++ww) bits += __builtin_popcount(w[ww]);
            for (int i = 0; i < n; ++i) {
                float t;
                if (!sphere_t(sph.data() + 4 * i, o, d, tmin, tmax, t)) continue;
                ++hits;
                if (visited < 0) continue;                                        // the caller takes every sphere
                if (!((w[bit[i] >> 5] >> (bit[i] & 31)) & 1u)) {
                    if (++bad <= 5) fprintf(stderr, "MISS scene %d ray %d sphere %d t %g visited %d\n", sc, r, i, t, visited);
                }
            }
        }
    }
    printf("%lld %lld %lld %lld %lld\n", rays, hits, 100 * cells / rays, 100 * bits / rays, bad);
    return bad ? 1 : 0;
}
