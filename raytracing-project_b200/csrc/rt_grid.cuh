// rt_grid.cuh — a uniform grid over the plain spheres / pokeballs of a scene, for rays that belong to no bundle.
//
// Scene::intersect / Scene::occluded are linear scans (scene.cpp:10-42).  Coherent rays (camera rays, their shadow rays) are
// served by the round-level bundle cull (rt_cull.cuh); the secondary rays of a reflecting / refracting scene are not
// coherent, and each of them paid one conservative mask test per sphere of the scene (synthetic-B: 68 tests per ray, 57 % of
// the wavefront kernel's instructions).  Here every cell of a uniform grid holds a candidate bit mask of the spheres whose
// (inflated) bounding box overlaps it; a ray walks the cells it crosses (3-D DDA, Amanatides & Woo 1987) and ORs their masks.
// The scan then finishes the candidates in JSON order with the reference's exact arithmetic, so results equal the linear
// scan as long as the mask is a SUPERSET of the spheres the exact FP32 test accepts:
//   * every sphere is binned into all cells its box [c - r - pad, c + r + pad] overlaps, pad = 2e-3 of the cell size plus
//     1e-5 of the coordinates — three orders of magnitude above the rounding of the walk (~1e-6 relative), so a cell the walk
//     skips or enters early because of rounding lies within pad of a visited cell;
//   * the grid box contains every binned box; a ray whose slab interval against it is empty misses them all;
//   * spheres much larger than the rest (the r = 1000 "floor") would blow the box up until every small sphere shares a cell:
//     they are not binned but kept in an ALWAYS mask (the cell after the last one) that every walk starts from;
//   * a NaN anywhere in the ray selects every sphere (the exact test treats NaN as a hit candidate).
// Pure arithmetic (no scene access beyond the mask table, no CUDA intrinsics) so that tests/host/grid_conservative.cpp can
// check the superset property on the host with the same code.
#pragma once
#include <math.h>
#include <stdint.h>

#ifdef __CUDACC__
#define RT_GRID_FN __host__ __device__ __forceinline__
#else
#define RT_GRID_FN inline
#endif
namespace rtd {

constexpr int kGridMaxWords = 4;          // scenes of up to 128 top-level objects; a cell is always 4 words (one 16-byte load)
struct GridHeader {
    float lo[3], hi[3];                   // the grid box (contains every binned box)
    float cell[3], inv_cell[3];
    int n[3];
    int words;                            // candidate words in use (0: the scene has no grid)
};
struct GridWords { uint32_t x, y, z, w; };                                 // uint4 on the device

// Walks the cells the segment o + t d, t in [tmin, tmax], crosses, FRONT TO BACK, and calls visit(words, t1) for the ALWAYS
// cell and then for every cell; `t1` is the far end of the segment and the visitor may lower it — to its closest hit so far
// (no sphere can be hit before the next cell's entry that was not a candidate of a visited cell: a hit point lies in a visited
// cell), or to -inf to stop.  Returns the number of cells visited (0: the segment misses the grid box; -1: NaN in the ray, no
// statement, no visit).  Everything lives in scalars (no array is indexed by the stepping axis: that would put the walk's
// state into local memory).  `cell(i)` returns the four words of cell i (shared-memory or read-only global load on the
// device, an array on the host).
template <class CellLoad, class Visit>
RT_GRID_FN int grid_visit(const GridHeader& g, CellLoad cell, const float o[3], const float d[3], float tmin, float tmax, Visit visit) {
    const float s = o[0] + o[1] + o[2] + d[0] + d[1] + d[2] + tmin;
    if (!(s == s) || !(tmax == tmax)) return -1;                       // NaN (or inf - inf): no statement about the ray
    float t0 = tmin, t1 = tmax;
    visit(cell(g.n[0] * g.n[1] * g.n[2]), t1);                          // the spheres that are not binned
    float inv[3];
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        if (fabsf(d[a]) < 1e-20f) {                                       // parallel to the slab: inside or never
            if (o[a] < g.lo[a] || o[a] > g.hi[a]) return 0;
            inv[a] = 0.f;
        } else {
            inv[a] = 1.0f / d[a];
            const float ta = (g.lo[a] - o[a]) * inv[a], tb = (g.hi[a] - o[a]) * inv[a];
            t0 = fmaxf(t0, fminf(ta, tb)); t1 = fminf(t1, fmaxf(ta, tb));
        }
    }
    // (the binned boxes lie at least a cell's pad inside the grid box: the rounding of this interval cannot lose a hit)
    if (!(t0 <= t1)) return 0;
    const float t_slab = t1;                                              // the visitor lowers t1 below the box's far side only
    int ix[3], step[3]; float tnext[3], tdelta[3];
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        const float p = fmaf(d[a], t0, o[a]);
        int i = (int)floorf((p - g.lo[a]) * g.inv_cell[a]);
        i = i < 0 ? 0 : (i >= g.n[a] ? g.n[a] - 1 : i);
        ix[a] = i;
        if (inv[a] == 0.f) { step[a] = 0; tnext[a] = INFINITY; tdelta[a] = 0.f; }
        else {
            step[a] = d[a] > 0.f ? 1 : -1;
            const float edge = fmaf((float)(i + (d[a] > 0.f ? 1 : 0)), g.cell[a], g.lo[a]);
            tnext[a] = (edge - o[a]) * inv[a];
            tdelta[a] = g.cell[a] * fabsf(inv[a]);
        }
    }
    int x = ix[0], y = ix[1], z = ix[2];
    float tx = tnext[0], ty = tnext[1], tz = tnext[2];
    const int nx = g.n[0], ny = g.n[1], nz = g.n[2];
    const int limit = nx + ny + nz + 3;                                   // (a walk crosses at most nx + ny + nz cells)
    int visited = 0;
    for (;;) {
        visit(cell((z * ny + y) * nx + x), t1);
        t1 = fminf(t1, t_slab);
        ++visited;
        // next cell along the axis whose boundary comes first
        const bool sx = tx <= ty && tx <= tz, sy = !sx && ty <= tz, sz = !sx && !sy;
        const float tn = sx ? tx : (sy ? ty : tz);
        if (!(tn <= t1) || visited > limit) break;                        // the segment ends inside this cell (inf: no step left)
        x += sx ? step[0] : 0; y += sy ? step[1] : 0; z += sz ? step[2] : 0;
        if ((unsigned)x >= (unsigned)nx || (unsigned)y >= (unsigned)ny || (unsigned)z >= (unsigned)nz) break;     // left the grid
        tx += sx ? tdelta[0] : 0.f; ty += sy ? tdelta[1] : 0.f; tz += sz ? tdelta[2] : 0.f;
    }
    return visited;
}

// The plain form: ORs the masks of every cell the segment crosses into `out`.
template <class CellLoad>
RT_GRID_FN int grid_walk(const GridHeader& g, CellLoad cell, const float o[3], const float d[3], float tmin, float tmax, GridWords& out) {
    return grid_visit(g, cell, o, d, tmin, tmax, [&out](const GridWords& m, float&) { out.x |= m.x; out.y |= m.y; out.z |= m.z; out.w |= m.w; });
}

}  // namespace rtd

#include <algorithm>
#include <vector>
namespace rtd {
// Host: build the grid over `n` spheres (cx, cy, cz, r) whose candidate bits are `bit[i]` (device object index).  Returns
// false (no grid) when there is nothing finite to bin.  target_cells ~ cells per sphere.
inline bool grid_build(const float* spheres4, const int* bit, int n, int n_bits, GridHeader& g, std::vector<uint32_t>& masks, double cells_per_sphere = 2.0) {
    g = GridHeader{};
    if (n <= 0 || n_bits > 32 * kGridMaxWords) return false;
    // spheres beyond 8 x the median radius are not binned (ALWAYS mask)
    std::vector<double> radii;
    for (int i = 0; i < n; ++i) {
        const float* s = spheres4 + 4 * i;
        if (!(std::isfinite(s[0]) && std::isfinite(s[1]) && std::isfinite(s[2]) && std::isfinite(s[3]))) return false;
        radii.push_back(std::fabs((double)s[3]));
    }
    std::nth_element(radii.begin(), radii.begin() + n / 2, radii.end());
    const double r_big = 8.0 * radii[(size_t)(n / 2)];
    auto binned = [&](int i) { return std::fabs((double)spheres4[4 * i + 3]) <= r_big; };
    double lo[3] = {1e300, 1e300, 1e300}, hi[3] = {-1e300, -1e300, -1e300};
    int n_binned = 0;
    for (int i = 0; i < n; ++i) {
        if (!binned(i)) continue;
        ++n_binned;
        const float* s = spheres4 + 4 * i;
        for (int a = 0; a < 3; ++a) { lo[a] = std::min(lo[a], (double)s[a] - std::fabs((double)s[3])); hi[a] = std::max(hi[a], (double)s[a] + std::fabs((double)s[3])); }
    }
    if (n_binned == 0) return false;
    double ext[3], vol = 1.0, maxabs = 0.0;
    for (int a = 0; a < 3; ++a) { ext[a] = std::max(hi[a] - lo[a], 1e-6); vol *= ext[a]; maxabs = std::max({maxabs, std::fabs(lo[a]), std::fabs(hi[a])}); }
    const double want = std::cbrt(vol / std::max(1.0, cells_per_sphere * n_binned));
    int total = 1;
    for (int a = 0; a < 3; ++a) { g.n[a] = (int)std::min(32.0, std::max(1.0, std::ceil(ext[a] / want))); total *= g.n[a]; }
    // the grid box: the spheres' box plus a margin of 1 % of a cell + 1e-4 of the coordinates on every side
    for (int a = 0; a < 3; ++a) {
        const double c0 = ext[a] / g.n[a], margin = 0.01 * c0 + 1e-4 * maxabs + 1e-6;
        const double l = lo[a] - margin, h = hi[a] + margin;
        g.lo[a] = (float)l; g.hi[a] = (float)h;
        g.cell[a] = (float)((h - l) / g.n[a]); g.inv_cell[a] = (float)(g.n[a] / (h - l));
    }
    g.words = (n_bits + 31) / 32;
    masks.assign((size_t)(total + 1) * kGridMaxWords, 0u);     // + the ALWAYS cell; a cell is always 4 words
    for (int i = 0; i < n; ++i) {
        const float* s = spheres4 + 4 * i;
        if (!binned(i)) { masks[(size_t)total * kGridMaxWords + (bit[i] >> 5)] |= 1u << (bit[i] & 31); continue; }
        int c0[3], c1[3];
        for (int a = 0; a < 3; ++a) {
            const double pad = 2e-3 * g.cell[a] + 1e-5 * maxabs + 1e-7, r = std::fabs((double)s[3]);
            c0[a] = std::max(0, std::min(g.n[a] - 1, (int)std::floor(((double)s[a] - r - pad - g.lo[a]) / g.cell[a])));
            c1[a] = std::max(0, std::min(g.n[a] - 1, (int)std::floor(((double)s[a] + r + pad - g.lo[a]) / g.cell[a])));
        }
        for (int z = c0[2]; z <= c1[2]; ++z) for (int y = c0[1]; y <= c1[1]; ++y) for (int x = c0[0]; x <= c1[0]; ++x)
            masks[(size_t)((z * g.n[1] + y) * g.n[0] + x) * kGridMaxWords + (bit[i] >> 5)] |= 1u << (bit[i] & 31);
    }
    return true;
}
}  // namespace rtd
