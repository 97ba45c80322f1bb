// rt_trace.cuh — Scene::intersect / Scene::occluded, shade_lambert_phong and the bounded reflection / refraction
// recursion as device functions (FP32).
//
// There is ONE scan body (closest hit and any hit), ONE light term and ONE recursion step; the per-thread kernel, the
// wavefront kernel, paper mode and the known-answer kernel all inline the same code.  The scans are plain per-thread code
// (callable from divergent code); what they take from the round-level culling (rt_cull.cuh, rt_device.cuh cand_*) is the
// index of their bundle's candidate words — `cb`, or -1 for rays that belong to no bundle (secondary rays,
// known-answer rays), which then scan every run with the per-lane conservative mask.  Only shade<LISTS = true> and the
// kernels' round prologue contain CTA barriers (they build the masks) and must be reached by all threads of the CTA.
#pragma once
#include "rt_device.cuh"

namespace rtd {

// ---- generic object (wrapped / affine / program) ---------------------------------------------------
// Returns true and the hit parameter when object `i` is hit inside the caller's window.  `inclusive`
// windows are [tmin,tmax] (sphere/plane/pokeball, geometry.cpp:26-28,99), CSG uses t<tmax
// (csg.cpp:176), transform-wrapped objects tmin<t<tmax (transform.cpp:38).  For programs the entry
// code and raw entry time are returned for the deferred normal.  lb: the bundle whose candidate words cover the CSG leaves, or -1.
template <bool C, bool K>
__device__ __forceinline__ bool generic_hit(int i, V3 o, V3 d, float tmin, float tmax, const int lb, float& t_out,
                                            uint32_t& code_out, float& tent_out, unsigned& tests) {
    const int4 m = SV<C>::template obj_meta_k<K>(i);
    const int kind = m.x & 0xff, flags = m.x >> 8;
    if (flags & RT_FLAG_NEVER) return false;
    const bool wrapped = flags & RT_FLAG_WRAPPED;
    const float lo = wrapped ? 0.f : tmin, hi = wrapped ? kInf : tmax;       // transform.cpp:29: child on [0,inf]
    float t;
    if (kind == RT_PRIM_PROGRAM) {
        // Conservative bounding sphere of the program's leaves (scene compiler; w < 0: none).  A ray whose line
        // misses it misses every leaf, so CSG::interval fails (csg.cpp:158); if all of it lies behind the
        // origin or beyond the window the first inside run cannot pass csg.cpp:175-176 / transform.cpp:38.
        const float4 bs = SV<C>::template obj_geom_k<K>(i);
        if (bs.w >= 0.f) {
            ++tests;
            float hb;
            const float disc = sphere_disc(bs, o, d, hb);
            if (disc < 0.f) return false;
            const float sq = sqrtf(disc);
            if (-hb + sq < 0.f || -hb - sq > tmax) return false;
        }
        const ProgResult pr = run_program<C>(m.z, m.w, o, d, lb);
        const Ivl v = pr.v; tests += pr.tests;
        if (!v.ok) return false;
        t = fmaxf(v.t0, lo);                                                 // csg.cpp:175-176
        if (!(t < v.t1 && t < hi)) return false;
        code_out = v.c0; tent_out = v.t0;
    } else {
        ++tests;
        const float4 g = SV<C>::template obj_geom_k<K>(i);
        if (kind == RT_PRIM_HALFSPACE) {
            if (!plane_t(g, o, d, lo, hi, t)) return false;
        } else if (flags & RT_FLAG_AFFINE) {
            V3 ol, dl; float L = to_local<C>(m.w, o, d, ol, dl);
            float tl;
            if (!sphere_t(g, ol, dl, lo, hi, tl)) return false;              // local window is [0,inf]: scale-free
            t = tl / L;
        } else {
            if (!sphere_t(g, o, d, lo, hi, t)) return false;
        }
        code_out = 0; tent_out = t;
    }
    if (wrapped && !(t > tmin && t < tmax)) return false;                    // transform.cpp:38
    t_out = t;
    return true;
}

// ---- Scene::intersect (scene.cpp:10-24) and Scene::occluded (scene.cpp:33-42): ONE scan per kind of ray -----------------
// Both queries walk the object list in JSON order; `any` (Scene::occluded) stops at the first accepted hit, otherwise the
// window's far end shrinks to the best hit so far (tmax == best_t until the first hit, so the two are the same loop).
// Every candidate is finished in JSON order with exactly the reference's root selection and window test, so the result
// equals a one-by-one scan.
struct ScanResult { int best; float t; uint32_t code; float tent; };

// (1) A ray of a round's bundle (camera rays, their shadow rays): the candidates are the set bits of the round's mask
// `cand` (rt_device.cuh cand_build) — objects in device order, then the CSG leaves.  A handful of bits per query.
#ifndef RT_SCAN_TWO_PHASE
#define RT_SCAN_TWO_PHASE 0          // A/B: plain primitives of a word first (classified by bit planes, no meta load, no call in the loop), generic objects after
#endif
template <bool C, bool K>
__device__ __forceinline__ void scan_masked(V3 o, V3 d, float tmin, float tmax, const bool any, const int cb, ScanResult& res, unsigned& tests) {
    int best = -1; float best_t = tmax; uint32_t best_code = 0; float best_tent = 0.f;
    const int W = c_hdr.cull_leaf_word;
    const uint32_t* cand = cand_words(cb);
#if RT_SCAN_TWO_PHASE
    // Per word: the plain spheres / pokeballs / half-spaces first — classified by two bit planes of the culling table, no
    // meta record is loaded and the loop contains no call — then the generic objects.  The order inside a word changes, the
    // result does not: plain primitives accept t <= best_t (the later index wins an exact tie, as in scene.cpp:15-21) and
    // are still visited in index order; a generic object visited after a plain one of higher index must not take over an
    // exact tie (CSG programs and wrapped primitives use t < tmax and never do, transform.cpp:38 / csg.cpp:176).
    if (!K) {
        const int Wc = c_hdr.cull_words;
        for (int w = 0; w < W; ++w) {
            const uint32_t bits = cand[w];
            if (!bits) continue;
            const uint32_t hs = __ldg(c_hdr.cull_plane_bits + Wc + w), gen = __ldg(c_hdr.cull_plane_bits + 2 * Wc + w);
            uint32_t pb = bits & ~gen;
            while (pb) {
                const int k = __ffs(pb) - 1; pb &= pb - 1;
                const int i = (w << 5) + k;
                ++tests;
                const float4 g = SV<C>::template obj_geom_k<K>(i);
                float t;
                if (!(((hs >> k) & 1u) ? plane_t(g, o, d, tmin, best_t, t) : sphere_t(g, o, d, tmin, best_t, t))) continue;
                best = i; best_t = t;
                if (any) goto done;
            }
            uint32_t gb = bits & gen;
            while (gb) {
                const int i = (w << 5) + __ffs(gb) - 1; gb &= gb - 1;
                float t, tent; uint32_t code;
                if (!generic_hit<C, K>(i, o, d, tmin, best_t, cb, t, code, tent, tests)) continue;
                if (t == best_t && i < best) continue;                       // (an exact tie stays with the later index)
                best = i; best_t = t; best_code = code; best_tent = tent;
                if (any) goto done;
            }
        }
        goto done;
    }
#endif
    for (int w = 0; w < W; ++w) {
        uint32_t bits = cand[w];
        while (bits) {
            const int i = (w << 5) + __ffs(bits) - 1; bits &= bits - 1;
            const int4 m = SV<C>::template obj_meta_k<K>(i);
            const int kind = m.x & 0xff;
            float t;
            if ((m.x >> 8) == 0 && kind != RT_PRIM_PROGRAM) {
                ++tests;
                const float4 g = SV<C>::template obj_geom_k<K>(i);
                if (!(kind == RT_PRIM_HALFSPACE ? plane_t(g, o, d, tmin, best_t, t) : sphere_t(g, o, d, tmin, best_t, t))) continue;
                // clustered runs are reordered spatially: exact ties go to the later JSON index as in scene.cpp:15-21
                if (K && t == best_t && best >= 0 && SV<C>::obj_orig(i) < SV<C>::obj_orig(best)) continue;
                best = i; best_t = t;
            } else {
                float tent; uint32_t code;
                if (!generic_hit<C, K>(i, o, d, tmin, best_t, cb, t, code, tent, tests)) continue;
                best = i; best_t = t; best_code = code; best_tent = tent;
            }
            if (any) goto done;
        }
    }
done:
    res.best = best; res.t = best_t; res.code = best_code; res.tent = best_tent;
}

// (2) A ray that belongs to no bundle (secondary rays, known-answer rays, rounds whose bundle is too wide): run by run,
// the candidates of a run from the per-lane conservative mask (sphere_mask).  Scenes of a few dozen primitives walk the
// runs backwards for any-hit (the order cannot be observed, scene.cpp:33-42 returns a bool, and later objects are usually
// the foreground detail that casts the shadows).
// G: the kernel serves scenes with a grid (rt_grid.cuh; the wavefront kernel only: the other kernels carry none of that code)
template <bool C, bool K, bool G = false>
__device__ __forceinline__ void scan_body(V3 o, V3 d, float tmin, float tmax, const bool any, ScanResult& res, unsigned& tests) {
    int best = -1; float best_t = tmax; uint32_t best_code = 0; float best_tent = 0.f;
    const int n_runs = c_hdr.n_runs;
    // Scenes with a grid (rt_grid.cuh): the candidates of every unclustered run of plain spheres / pokeballs
    // come from ONE walk through the cells the ray crosses instead of one conservative test per sphere and run.
    GridWords gw = {0u, 0u, 0u, 0u};
    bool grid = false;
    if (G && c_hdr.grid.words) {
        const float oo[3] = {o.x, o.y, o.z}, dd[3] = {d.x, d.y, d.z};
        const int visited = grid_walk(c_hdr.grid, [](int c) { return SV<C>::grid_cell(c); }, oo, dd, tmin, tmax, gw);
        grid = visited >= 0;                                                 // (NaN in the ray: the per-sphere masks decide)
        tests += (unsigned)max(visited, 0);
    }
    for (int rr = 0; rr < n_runs; ++rr) {
        const int4 run = SV<C>::run((any && !K) ? n_runs - 1 - rr : rr);
        const int first = run.y, end = run.y + run.z;
        if (K && (run.x == RUN_SPHERE || run.x == RUN_POKEBALL) && run.w > 0) {
            // Clustered run (spatially reordered, 16 primitives per cluster): candidate mask over the cluster bounds, then
            // over the members of the clusters whose bound overlaps the window.  A sphere's candidate root does not depend on
            // the order of the scan; exact ties go to the later JSON index as in scene.cpp:15-21.
            const int ctab = run.w - 1, ncl = (run.z + 15) >> 4;
            tests += ncl;
            for (int cb = 0; cb < ncl; cb += 32) {
                uint32_t cmask = sphere_mask<C>(SV<C>::cluster_ptr(ctab + cb), min(32, ncl - cb), o, d);
                while (cmask) {
                    const int c = cb + __ffs(cmask) - 1; cmask &= cmask - 1;
                    float hb;
                    const float disc = sphere_disc(SV<C>::cluster(ctab + c), o, d, hb);
                    const float sq = sqrtf(disc);
                    if (-hb + sq < 0.f || -hb - sq > best_t) continue;
                    const int m0 = first + 16 * c, n = min(16, end - m0);
                    tests += n;
                    uint32_t mask = sphere_mask<C>(SV<C>::obj_ptr(m0), n, o, d);
                    while (mask) {
                        const int i = m0 + __ffs(mask) - 1; mask &= mask - 1;
                        float t;
                        if (!sphere_t(SV<C>::template obj_geom_k<K>(i), o, d, tmin, best_t, t)) continue;
                        if (t == best_t && best >= 0 && SV<C>::obj_orig(i) < SV<C>::obj_orig(best)) continue;
                        best = i; best_t = t;
                        if (any) goto done;
                    }
                }
            }
        } else if (run.x == RUN_SPHERE || run.x == RUN_POKEBALL) {
            for (int c0 = first; c0 < end; c0 += 32) {                       // candidates first, finished in JSON order
                const int n = min(32, end - c0);
                uint32_t mask;
                if (G && grid) {                                             // bits [c0, c0 + n) of the walk's words
                    const int w = c0 >> 5, sh = c0 & 31;
                    const uint32_t lo = w == 0 ? gw.x : w == 1 ? gw.y : w == 2 ? gw.z : gw.w;
                    const uint32_t hi = (sh + n > 32) ? (w == 0 ? gw.y : w == 1 ? gw.z : w == 2 ? gw.w : 0u) : 0u;
                    mask = __funnelshift_r(lo, hi, sh) & (n >= 32 ? 0xffffffffu : ((1u << n) - 1u));
                    tests += __popc(mask);
                } else {
                    tests += n;
                    mask = sphere_mask<C>(SV<C>::obj_ptr(c0), n, o, d);
                }
                while (mask) {
                    const int i = c0 + __ffs(mask) - 1; mask &= mask - 1;
                    float t;
                    if (sphere_t(SV<C>::template obj_geom_k<K>(i), o, d, tmin, best_t, t)) { best = i; best_t = t; if (any) goto done; }
                }
            }
        } else if (run.x == RUN_HALFSPACE) {
            for (int i = first; i < end; ++i) {
                float t; ++tests;
                if (plane_t(SV<C>::template obj_geom_k<K>(i), o, d, tmin, best_t, t)) { best = i; best_t = t; if (any) goto done; }
            }
        } else {
            // generic objects (CSG programs, wrapped / affine primitives): one world bound per object (R = inf: unknown),
            // scanned with the candidate mask before the object itself is looked at
            const int ctab = run.w - 1;
            tests += run.z;
            for (int c0 = first; c0 < end; c0 += 32) {
                const int n = min(32, end - c0);
                uint32_t mask = sphere_mask<C>(SV<C>::cluster_ptr(ctab + c0 - first), n, o, d);
                while (mask) {
                    const int i = c0 + __ffs(mask) - 1; mask &= mask - 1;
                    float t, tent; uint32_t code;
                    if (generic_hit<C, K>(i, o, d, tmin, best_t, -1, t, code, tent, tests)) {
                        best = i; best_t = t; best_code = code; best_tent = tent;
                        if (any) goto done;
                    }
                }
            }
        }
    }
done:
    res.best = best; res.t = best_t; res.code = best_code; res.tent = best_tent;
}

// Scenes with clustered runs (K: ~10^3 primitives) keep ONE out-of-line copy of scan (2) for both queries: all warps of an
// SM spend most of their time in this one function, so its code stays resident in the instruction cache (synthetic-A 2K
// 29.0 -> 23.7 ms when this was introduced).  Scenes of a few dozen primitives inline the two specialisations (the call
// cost them more than it saved).  The test count is returned by value (a reference would pin the caller's counter to
// local memory).
struct ScanOut { ScanResult r; unsigned tests; };
template <bool C, bool K, bool G = false>
__device__ __noinline__ ScanOut scan_shared(V3 o, V3 d, float tmin, float tmax, bool any) {
    ScanOut out; out.tests = 0;
    scan_body<C, K, G>(o, d, tmin, tmax, any, out.r, out.tests);
    return out;
}
#ifndef RT_SCAN_SHARED
#define RT_SCAN_SHARED 0          // A/B: 1 = the out-of-line scan for every scene
#endif
// A/B build switch: 1 = ONE out-of-line copy of the masked scan serves closest hit and any hit (the callee gets a fresh
// register budget and the callers' loop nests lose a level of convergence barriers; the call costs a spill of the caller's state)
#ifndef RT_MASKED_CALL
#define RT_MASKED_CALL 0
#endif
template <bool C, bool K>
__device__ __noinline__ ScanOut scan_masked_shared(V3 o, V3 d, float tmin, float tmax, bool any, int cb) {
    ScanOut out; out.tests = 0;
    scan_masked<C, K>(o, d, tmin, tmax, any, cb, out.r, out.tests);
    return out;
}
// M: the kernel was built with round-level masks (rt_device.cuh); kernels without them contain none of that code.
template <bool C, bool K, bool M, bool ANY, bool G = false>
__device__ __forceinline__ void scan(V3 o, V3 d, float tmin, float tmax, const int cb, ScanResult& res, Counters& cnt) {
    if (M && cb >= 0) {
        if (RT_MASKED_CALL) { const ScanOut so = scan_masked_shared<C, K>(o, d, tmin, tmax, ANY, cb); res = so.r; cnt.tests += so.tests; }
        else scan_masked<C, K>(o, d, tmin, tmax, ANY, cb, res, cnt.tests);
    }
    else if (K || RT_SCAN_SHARED) { const ScanOut so = scan_shared<C, K, G>(o, d, tmin, tmax, ANY); res = so.r; cnt.tests += so.tests; }
    else scan_body<C, K, G>(o, d, tmin, tmax, ANY, res, cnt.tests);
}

template <bool C, bool K, bool M, bool G = false>
__device__ __forceinline__ bool closest_hit(V3 o, V3 d, float tmin, float tmax, const int cb, Hit& hit, Counters& cnt) {
    ScanResult sr;
    scan<C, K, M, false, G>(o, d, tmin, tmax, cb, sr, cnt);
    const int best = sr.best;
    if (best < 0) return false;
    // ---- finalise the winner: normal, front_face, material, id code ----
    const int4 m = SV<C>::template obj_meta_k<K>(best);
    const int kind = m.x & 0xff, flags = m.x >> 8;
    hit.t = sr.t;
    hit.code = SV<C>::obj_orig(best) * 8;
    V3 outward; bool front; int mat;
    if (kind == RT_PRIM_PROGRAM) {
        const uint32_t code = sr.code;
        const int li = code & 0xFFFFF, region = (code >> 20) & 7, nmode = (code >> 24) & 3;
        const int4 lm = SV<C>::leaf_meta(li);
        mat = lm.y + region;
        front = (code & kFrontBit) != 0;
        V3 n;
        if (nmode == 1) n = mk(0.f, 0.f, 0.f);                               // csg.cpp:119
        else {
            if (nmode == 2) { float4 f = SV<C>::fallback(code >> 26); outward = mk(f.x, f.y, f.z); }
            else if ((lm.x & 0xff) == RT_PRIM_HALFSPACE) { float4 g = SV<C>::leaf_geom(li); outward = mk(g.x, g.y, g.z); }
            else { int rg; outward = sphere_outward<C>(SV<C>::leaf_geom(li), lm, o, d, sr.tent, rg); }
            n = dot(d, outward) < 0.f ? outward : neg(outward);              // geometry.h:42-45
        }
        if (flags & RT_FLAG_WRAPPED) front = dot(d, n) < 0.f;               // transform.cpp:41
        hit.n = n; hit.mat = mat; hit.front = front;
        return true;
    }
    const float4 g = SV<C>::template obj_geom_k<K>(best);
    int region = 0;
    if (kind == RT_PRIM_HALFSPACE) outward = mk(g.x, g.y, g.z);
    else outward = sphere_outward<C>(g, m, o, d, sr.t, region);
    front = dot(d, outward) < 0.f;
    V3 n = front ? outward : neg(outward);
    if (flags & RT_FLAG_XFORMED) front = dot(d, n) < 0.f;                   // transform.cpp:41
    mat = m.y;
    if (kind == RT_PRIM_POKEBALL) { mat += region; if (!(flags & RT_FLAG_XFORMED)) hit.code += region + 1; }   // code: only a bare top-level pokeball reports its region
    hit.n = n; hit.mat = mat; hit.front = front;
    return true;
}

template <bool C, bool K, bool M, bool G = false>
__device__ __forceinline__ bool occluded(V3 o, V3 d, float tmin, float tmax, const int cb, Counters& cnt) {
    ScanResult sr;
    scan<C, K, M, true, G>(o, d, tmin, tmax, cb, sr, cnt);
    return sr.best >= 0;
}

// ---- shade_lambert_phong (shading.cpp:31-138) ------------------------------------------------------------------
struct RGBf { float r, g, b; };
__device__ __forceinline__ RGBf rgb(float r, float g, float b) { RGBf c; c.r = r; c.g = g; c.b = b; return c; }
// "screen" blend (shading.cpp:6-12): 1 - (1 - a)(1 - b) = a + b - a b, as one add and one explicit FMA instead of three adds
// and a multiply (the blend runs ~11 times per sample and light: it was 5 % of the snorlax kernel's instructions)
#ifndef RT_SCREEN_FMA
#define RT_SCREEN_FMA 1
#endif
__device__ __forceinline__ float screen1(float a, float b) { return RT_SCREEN_FMA ? fmaf(-a, b, a + b) : 1.f - (1.f - a) * (1.f - b); }
__device__ __forceinline__ RGBf screen(RGBf a, RGBf b) { return rgb(screen1(a.r, b.r), screen1(a.g, b.g), screen1(a.b, b.b)); }

// Direction to light `lp` and its distance (shading.cpp:80-92; a directional light, lp.w != 0, is at infinity: shading.cpp:47).
__device__ __forceinline__ float light_vector(float4 lp, V3 p, V3 n, V3& wi, float& dist) {
    if (lp.w != 0.f) { wi = mk(lp.x, lp.y, lp.z); dist = kInf; return fmaxf(0.f, dot(n, wi)); }
    V3 L = mk(lp.x - p.x, lp.y - p.y, lp.z - p.z);
    float d2 = fmaxf(dot(L, L), 0.01f);
#if RT_RSQRT
    const float inv = rsqrtf(d2);                                            // one MUFU.RSQ serves the length and its reciprocal
    dist = d2 * inv;
#else
    dist = sqrtf(d2);
    float inv = 1.f / dist;
#endif
    wi = mk(L.x * inv, L.y * inv, L.z * inv);
    return fmaxf(0.f, dot(n, wi));
}
// One light's contribution, screen-blended into E (shading.cpp:56-73 directional, :108-129 point); the light vector is
// recomputed (15 instructions) so that no shading state has to stay live across the shadow scans.
template <bool C>
__device__ __forceinline__ void light_shade(int l, const float4 m0, float ks, float shininess, float kd, V3 n, V3 p, V3 wo, RGBf& E) {
    const float4 lp = SV<C>::light(l, 0), li = SV<C>::light(l, 1);
    V3 wi; float dist;
    const float ndotl = light_vector(lp, p, n, wi, dist);
    const bool directional = lp.w != 0.f;
    float eff = fmaxf(0.5f, dist);
    float fall = directional ? 1.f : 2.0f / (eff * eff);                     // falloff * 2.0 (shading.cpp:108-112); none for a directional light
    RGBf I = rgb(li.x * fall, li.y * fall, li.z * fall);
    float kdn = directional ? kd * ndotl : kd * ndotl * 1.5f;                // shading.cpp:115 / :61
    RGBf Ed = rgb(m0.x * I.r * kdn, m0.y * I.g * kdn, m0.z * I.b * kdn);
    RGBf Es = rgb(0.f, 0.f, 0.f);
    if (ks > 0.f) {                                                          // shading.cpp:118-126
        float two = 2.f * dot(n, wi);
        V3 rr = normalized(mk(fmaf(two, n.x, -wi.x), fmaf(two, n.y, -wi.y), fmaf(two, n.z, -wi.z)));
        float rdotv = fmaxf(0.f, dot(rr, wo));
        // exp2(s * log2(x)) on the MUFU pipe (relative error ~ s * 2^-21, far below 1/255).  x == 0 (every light whose mirror
        // direction points away from the viewer) is std::pow(+0, s): 0 for s > 0, 1 for s == 0, inf for s < 0, NaN for NaN —
        // spelled out, because the full powf behind this branch used to run for most warps
        const float pow0 = shininess > 0.f ? 0.f : (shininess == 0.f ? 1.f : (shininess < 0.f ? kInf : shininess));
        float spec = (rdotv > 0.f ? __powf(rdotv, shininess) : pow0) * ks;
        Es = rgb(I.r * spec, I.g * spec, I.b * spec);
    }
    E = screen(E, screen(Ed, Es));
}

// shade_lambert_phong for a thread with `lit` (a hit with a material); the others get (0,0,0).  Two passes per chunk of 32
// lights: all shadow queries first (one visibility bit each), the contributions of the visible lights after, so that no
// shading state is live across the scans.
// LISTS (uniform over the callers): how the shadow scans get their candidates.
//   kListsNone  per-lane conservative masks only; the call may come from divergent code (secondary rays, known-answer rays);
//   kListsCta   the call is made by ALL threads of the CTA for the camera rays of one round: the CTA reduces the box of the
//               shadow-ray origins and culls the scene against the bundle towards every light (cand_build, two barriers);
//   kListsWarp  the call is made by all 32 lanes of a WARP for its 32 camera rays (4 neighbouring pixels): the warp reduces
//               the box of its own shadow-ray origins into uniform registers and, light by light, culls the scene against the
//               bundle towards that light, one entry per lane, into its own mask words — no barrier, no shared atomics, and
//               the warps of a CTA need not walk in step.
enum { kListsNone = 0, kListsCta = 1, kListsWarp = 2 };
#ifndef RT_SHADOW_RENORM
#define RT_SHADOW_RENORM 0
#endif
template <bool C, bool K, bool M, bool G = false>
__device__ __forceinline__ RGBf shade(bool lit, const int LISTS, const Hit& h, V3 p, V3 wo, Counters& cnt) {
    const float eps = fmaxf(1e-3f, 1e-4f * h.t);                             // shading.cpp:43
    const V3 so = madd(p, h.n, eps);
    const int nl = c_hdr.n_lights;
    const int lists = M ? LISTS : kListsNone;
    float blo[3] = {0.f, 0.f, 0.f}, bhi[3] = {0.f, 0.f, 0.f}, beps = 0.f; bool bok = false;      // kListsWarp: the warp's origin box
    if (lists == kListsCta) {
        CandShared* S = cand_shared();
        cand_box_add(lit, so, S->org_lo, S->org_hi, &S->nan_flags, 2);
        const float e = warp_max_nan(lit ? eps : 0.f);
        if ((threadIdx.x & 31) == 0) { if (e == e) atomicMax(&S->eps_max, __float_as_int(e)); else atomicOr(&S->nan_flags, 2); }
        __syncthreads();
        cnt.tests += cand_build(1, min(nl, kMaxListLights));
        if (threadIdx.x < 3) { S->org_lo[threadIdx.x] = kOrdPosInf; S->org_hi[threadIdx.x] = kOrdNegInf; }     // for the next round
        if (threadIdx.x == 3) { S->eps_max = 0; atomicAnd(&S->nan_flags, ~2); }
    } else if (lists == kListsWarp) {
        blo[0] = warp_min_nan(lit ? so.x : kInf); blo[1] = warp_min_nan(lit ? so.y : kInf); blo[2] = warp_min_nan(lit ? so.z : kInf);
        bhi[0] = warp_max_nan(lit ? so.x : -kInf); bhi[1] = warp_max_nan(lit ? so.y : -kInf); bhi[2] = warp_max_nan(lit ? so.z : -kInf);
        beps = warp_max_nan(lit ? eps : 0.f);
        // (a NaN anywhere fails these comparisons: the warp then scans without masks)
        bok = blo[0] <= bhi[0] && blo[1] <= bhi[1] && blo[2] <= bhi[2] && beps >= 0.f;
    }
    RGBf E = rgb(0.f, 0.f, 0.f);
    if (lists != kListsWarp && !lit) return E;
    const int my_bundle = 1 + (int)(threadIdx.x >> 5);                      // kListsWarp: the warp's own mask slot
    uint32_t* my_words = lists == kListsWarp ? cand_words(my_bundle) : nullptr;
    // One running light index l: a chunk is the lights [l & ~31, ...) that share the 32 visibility bits, so two registers
    // (l, vis) stay alive across the scans; nothing of the shading (E, materials, wo's products) is live in pass 1.
    int l = 0;
    do {
        uint32_t vis = 0;
        const int chunk_end = min(nl, (l | 31) + 1);
        for (; l < chunk_end; ++l) {
            V3 wi; float dist;
            const float4 lp = SV<C>::light(l, 0);
            const float ndotl = light_vector(lp, p, h.n, wi, dist);
            const float tmaxS = dist - eps;
            const bool need = lit && !(ndotl <= 0.f) && !(tmaxS <= eps);     // shading.cpp:51,94 / :98
            int cb = -1;                                                     // bundle whose mask serves this light (-1: none)
            if (lists == kListsWarp) {
                if (!__any_sync(kFullWarp, need)) continue;                 // (warp-uniform)
                if (bok) {
                    Bundle B;
                    const float L[3] = {lp.x, lp.y, lp.z};
                    bundle_from_box_light(B, blo, bhi, beps, L, lp.w != 0.f);
                    const int W = c_hdr.cull_words, lane = threadIdx.x & 31;
                    for (int w = 0; w < W; ++w) {
                        const float4 g = __ldg(c_hdr.cull_geom + 32 * w + lane);
                        const bool plane = (__ldg(c_hdr.cull_plane_bits + w) >> lane) & 1u;
                        Sphere4 q; q.x = g.x; q.y = g.y; q.z = g.z; q.w = g.w;
                        const bool c1 = plane ? (B.wide || !bundle_rejects_plane(B, q)) : (g.w >= 0.f && (B.wide || !bundle_rejects(B, q, B.t_hi, w < c_hdr.cull_leaf_word)));
                        ++cnt.tests;
                        const uint32_t bits = __ballot_sync(kFullWarp, c1);
                        if (lane == 0) my_words[w] = bits;
                    }
                    __syncwarp();
                    cb = my_bundle;
                }
            } else {
                if (!need) continue;
                if (lists == kListsCta) cb = cand_id(1 + l);
            }
            if (need) {
                ++cnt.shadow;
                // (the Ray ctor re-normalises its direction, core.h:278: in double that is the identity up to 1e-16; wi = L * rsqrt(|L|^2)
                // is unit to 2 ulp already, and a second FP32 normalisation brings it no closer to the exact direction)
                if (!occluded<C, K, M, G>(so, RT_SHADOW_RENORM ? normalized(wi) : wi, eps, tmaxS, cb, cnt)) vis |= 1u << (l & 31);
            }
            if (lists == kListsWarp) __syncwarp();                          // the words are rewritten for the next light
        }
        if (!lit) continue;
        const int l0 = (l - 1) & ~31;                                        // first light of the chunk just queried
        if (l0 == 0) { const float4 m1 = material(h.mat, 1); E = rgb(m1.x, m1.y, m1.z); }   // ambient (x) scene ambient, pre-multiplied
        const float4 m0 = material(h.mat, 0), m2 = material(h.mat, 2);
        while (vis) {
            const int k = __ffs(vis) - 1; vis &= vis - 1;
            light_shade<C>(l0 + k, m0, m0.w, m2.y, m2.w, h.n, p, wo, E);
        }
    } while (l < nl);
    if (nl <= 0 && lit) { const float4 m1 = material(h.mat, 1); E = rgb(m1.x, m1.y, m1.z); }       // no lights: ambient only
    return rgb(fminf(1.5f, E.r), fminf(1.5f, E.g), fminf(1.5f, E.b));        // shading.cpp:133-135
}

// ---- Tracer::trace_recursive (tracer.cpp:22-104), one ray at a time ---------------------------------------------------
// A node with both kr > 0 and kt > 0 is bilinear in its two children under the screen blend, so the recursion cannot be
// flattened into a throughput product; each pending node keeps its running total and, when a refraction is still owed,
// the refracted ray: 3 float4 per depth (total.rgb k_pending | kt_owed ro.xyz | rd.xyz -) in `frames` — thread-local
// memory for the per-thread kernel, global memory for the wavefront kernel (any thread can then take any unfinished
// sample from the CTA's queue).
//
// ray_step traces the ray (o, d) of a thread with `alive` at depth sp: closest hit, shading incl. shadow rays, then the
// push / pop of the tree.  Returns true when the thread's tree is finished (`result` = Tracer::trace's colour), false when
// (o, d, sp) is its next ray.  FLAT: no ray of the scene can spawn a child (no reflective / refractive material, or
// recursion == 1): no frames are touched and every tree finishes with its first ray.  LISTS: the ray is a camera ray of
// a round — `cand` is the round's mask of the camera bundle, ALL threads of the CTA make the call (see shade).
template <bool C, bool K, bool M, bool FLAT, bool G = false>
__device__ __forceinline__ bool ray_step(bool alive, const int LISTS, const int cb, V3& o, V3& d, int& sp, float4* __restrict__ frames,
                                         RGBf& result, Hit& primary, bool& primary_hit, Counters& cnt) {
    const int limit = c_hdr.recursion;
    Hit h; h.mat = -1; h.t = 0.f; h.n = mk(0.f, 1.f, 0.f); h.front = true; h.code = -1;
    bool hit = false;
    if (alive) {
        hit = closest_hit<C, K, M, G>(o, d, 1e-4f, kInf, cb, h, cnt);      // tracer.cpp:29
        if (FLAT || LISTS || sp == 0) { ++cnt.primary; primary = h; primary_hit = hit; } else ++cnt.secondary;
    }
    const V3 p = madd(o, d, h.t);
    const V3 wo = normalized(neg(d));
    const bool lit = hit && h.mat >= 0;
    if (!LISTS && !alive) return true;
    const RGBf total = shade<C, K, M, G>(lit, LISTS, h, p, wo, cnt);        // tracer.cpp:31
    if (!alive) return true;
    if (!hit) result = rgb(c_hdr.bg[0], c_hdr.bg[1], c_hdr.bg[2]);          // tracer.cpp:72
    else if (!lit) result = rgb(1.f, 0.f, 1.f);                              // shading.cpp:33
    else {
        result = total;
        if (!FLAT && sp < limit - 1) {
            const float4 m1 = material(h.mat, 1), m2 = material(h.mat, 2);
            const float kr = m1.w, kt = m2.x, ior = m2.z;
            const V3 I = normalized(d);
            bool want_r = kr > 0.f, want_t = false;
            V3 to = mk(0.f, 0.f, 0.f), td = mk(0.f, 0.f, 0.f);
            if (kt > 0.f) {                                                  // tracer.cpp:51-68
                float eta = h.front ? (c_hdr.medium_index / ior) : (ior / c_hdr.medium_index);
                float cos_i = -dot(I, h.n);
                float sin_t2 = eta * eta * fmaxf(0.f, 1.f - cos_i * cos_i);
                if (!(sin_t2 >= 1.f)) {                                      // TIR guard, tracer.cpp:100-104
                    float cos_t = sqrtf(1.f - sin_t2);
                    float kk = eta * cos_i - cos_t;
                    td = normalized(mk(fmaf(kk, h.n.x, eta * I.x), fmaf(kk, h.n.y, eta * I.y), fmaf(kk, h.n.z, eta * I.z)));
                    to = h.front ? madd(p, h.n, -1e-6f) : madd(p, h.n, 1e-6f);
                    want_t = true;
                }
            }
            if (want_r || want_t) {
                float4* f = frames + 3 * sp;
                const float owed = want_t && want_r ? kt : 0.f;
                if (want_r) {                                                // tracer.cpp:38-48
                    float two = 2.f * dot(I, h.n);
                    V3 rd = normalized(mk(fmaf(-two, h.n.x, I.x), fmaf(-two, h.n.y, I.y), fmaf(-two, h.n.z, I.z)));
                    V3 ro = h.front ? madd(p, h.n, 1e-6f) : madd(p, h.n, -1e-6f);
                    f[0] = make_float4(total.r, total.g, total.b, kr);
                    f[1] = make_float4(owed, to.x, to.y, to.z);
                    if (want_t) f[2] = make_float4(td.x, td.y, td.z, 0.f);
                    o = ro; d = rd;
                } else {
                    f[0] = make_float4(total.r, total.g, total.b, kt);
                    f[1] = make_float4(0.f, 0.f, 0.f, 0.f);
                    o = to; d = td;
                }
                ++sp;
                return false;                                                // descend
            }
        }
    }
    if (FLAT) return true;
    // ---- return `result` to the pending parents ----
    for (;;) {
        if (sp == 0) return true;
        float4* f = frames + 3 * (sp - 1);
        const float4 f0 = f[0], f1 = f[1];
        const float k = f0.w;
        const RGBf tot = screen(rgb(f0.x, f0.y, f0.z), rgb(result.r * k, result.g * k, result.b * k));   // tracer.cpp:47,66
        if (f1.x > 0.f) {                      // reflection done, refraction owed: trace it at depth sp (frame stays)
            const float4 f2 = f[2];
            f[0] = make_float4(tot.r, tot.g, tot.b, f1.x);
            f[1] = make_float4(0.f, 0.f, 0.f, 0.f);
            o = mk(f1.y, f1.z, f1.w); d = mk(f2.x, f2.y, f2.z);
            return false;
        }
        result = tot;
        --sp;
    }
}

// Tracer::trace (tracer.cpp:12-20) for one sample per thread.  `active` false: the thread has no sample.  D = frames
// available = deepest recursion served.  LISTS: the sample is a camera ray of a round whose camera-bundle mask is `cand`;
// ALL threads of the CTA make the call (the first ray and its shadow rays use the round's masks, deeper rays none).
template <bool C, bool K, bool M, bool FLAT, int D>
__device__ __forceinline__ RGBf trace(bool active, const int LISTS, const int cb, V3 o, V3 d, Hit& primary, bool& primary_hit, Counters& cnt) {
    primary_hit = false;
    RGBf result = rgb(0.f, 0.f, 0.f);
    if (c_hdr.recursion <= 0) {                                              // tracer.cpp:26 at depth 0: black, ids still defined
        if (active) { ++cnt.primary; primary_hit = closest_hit<C, K, M>(o, d, 1e-4f, kInf, cb, primary, cnt); }
        return result;
    }
    if (FLAT) {
        int sp = 0;
        ray_step<C, K, M, true>(active, LISTS, cb, o, d, sp, nullptr, result, primary, primary_hit, cnt);
        return result;
    }
    float4 frames[3 * D];
    int sp = 0;                     // == depth of the ray being traced
    int lists = LISTS;              // the camera ray: all threads; deeper rays: the threads whose tree goes on, no masks
    for (;;) {
        const bool done = ray_step<C, K, M, false>(active, lists, lists ? cb : -1, o, d, sp, frames, result, primary, primary_hit, cnt);
        if (!active || done) break;
        lists = kListsNone;
    }
    return result;
}

}  // namespace rtd
