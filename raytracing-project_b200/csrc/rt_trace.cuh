// rt_trace.cuh — Scene::intersect / Scene::occluded, shade_lambert_phong and the bounded
// reflection/refraction recursion as device functions (FP32).
#pragma once
#include "rt_device.cuh"

namespace rtd {

// ---- generic object (wrapped / affine / program) ---------------------------------------------------
// Returns true and the hit parameter when object `i` is hit inside the caller's window.  `inclusive`
// windows are [tmin,tmax] (sphere/plane/pokeball, geometry.cpp:26-28,99), CSG uses t<tmax
// (csg.cpp:176), transform-wrapped objects tmin<t<tmax (transform.cpp:38).  For programs the entry
// code and raw entry time are returned for the deferred normal.
template <bool C, bool K>
__device__ bool generic_hit(int i, V3 o, V3 d, float tmin, float tmax, float& t_out, uint32_t& code_out, float& tent_out, unsigned& tests) {
    const int4 m = SV<C>::obj_meta(i);
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
        const ProgResult pr = run_program<C>(m.z, m.w, o, d);
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

// ---- Scene::intersect (scene.cpp:10-24) and Scene::occluded (scene.cpp:33-42): ONE scan -------------------------------
// Both queries walk the object list in JSON order; `any` (Scene::occluded) stops at the first accepted hit, otherwise the
// window's far end shrinks to the best hit so far (tmax == best_t until the first hit, so the two are the same loop).
// One out-of-line copy serves the primary / secondary rays and the shadow rays: all warps of an SM spend most of their
// time in this one function, so its code stays resident in the instruction cache (two inlined copies, one in trace and
// one in shade, did not: `no_instruction` was the top stall of the generic kernel).  The test count is returned by
// value (a reference would pin the caller's counter to local memory).
struct ScanResult { int best; float t; uint32_t code; float tent; unsigned tests; };

#ifndef RT_SCAN_INLINE
#define RT_SCAN_INLINE 0
#endif
#if RT_SCAN_INLINE
#define RT_SCAN_FN __device__ __forceinline__
#else
#define RT_SCAN_FN __device__ __noinline__
#endif

template <bool C, bool K>
RT_SCAN_FN ScanResult scan_scene(V3 o, V3 d, float tmin, float tmax, bool any) {
    int best = -1; float best_t = tmax; uint32_t best_code = 0; float best_tent = 0.f; unsigned tests = 0;
    const int n_runs = c_hdr.n_runs;
    for (int r = 0; r < n_runs; ++r) {
        const int4 run = SV<C>::run(r);
        const int first = run.y, end = run.y + run.z;
        if (K && (run.x == RUN_SPHERE || run.x == RUN_POKEBALL) && run.w > 0) {
            // Clustered run (spatially reordered, 16 primitives per cluster): candidate mask over the cluster bounds,
            // then over the members of the clusters whose bound overlaps the window.  A sphere's candidate root does
            // not depend on the order of the scan; exact ties go to the later JSON index as in scene.cpp:15-21.
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
                tests += n;
                uint32_t mask = sphere_mask<C>(SV<C>::obj_ptr(c0), n, o, d);
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
            // generic objects (CSG programs, wrapped / affine primitives); a long run has one world bound per object,
            // scanned with the candidate mask first
            const bool bounded = K && run.w > 0;
            const int ctab = run.w - 1;
            if (bounded) tests += run.z;
            for (int c0 = first; c0 < end; c0 += 32) {
                const int n = min(32, end - c0);
                uint32_t mask = bounded ? sphere_mask<C>(SV<C>::cluster_ptr(ctab + c0 - first), n, o, d) : (n == 32 ? 0xffffffffu : (1u << n) - 1u);
                while (mask) {
                    const int i = c0 + __ffs(mask) - 1; mask &= mask - 1;
                    float t, tent; uint32_t code;
                    if (generic_hit<C, K>(i, o, d, tmin, best_t, t, code, tent, tests)) {
                        best = i; best_t = t; best_code = code; best_tent = tent;
                        if (any) goto done;
                    }
                }
            }
        }
    }
done:
    ScanResult res; res.best = best; res.t = best_t; res.code = best_code; res.tent = best_tent; res.tests = tests;
    return res;
}

// Scenes without clustered runs / per-object bounds (K = false: a few dozen primitives) keep the two scans inlined and
// specialised: their hot code fits the instruction cache anyway and the call costs more than it saves (snorlax 4K 12.6 ->
// 14.3 ms, pokeballs 0.54 -> 0.69 ms with the out-of-line scan; the 1 100-primitive scene 29.0 -> 23.7 ms).
// One run of the object list against the best hit so far.
template <bool C, bool K>
__device__ __forceinline__ void run_closest(const int4 run, V3 o, V3 d, float tmin, int& best, float& best_t, uint32_t& best_code,
                                            float& best_tent, Counters& cnt) {
    {
        const int first = run.y, end = run.y + run.z;
        if (K && (run.x == RUN_SPHERE || run.x == RUN_POKEBALL) && run.w > 0) {
            // Clustered run (spatially reordered, 16 primitives per cluster): candidate mask over the cluster bounds,
            // then over the members of the clusters whose bound overlaps the window.  A sphere's candidate root does
            // not depend on the order of the scan; exact ties go to the later JSON index as in scene.cpp:15-21.
            const int ctab = run.w - 1, ncl = (run.z + 15) >> 4;
            cnt.tests += ncl;
            for (int cb = 0; cb < ncl; cb += 32) {
                uint32_t cmask = sphere_mask<C>(SV<C>::cluster_ptr(ctab + cb), min(32, ncl - cb), o, d);
                while (cmask) {
                    const int c = cb + __ffs(cmask) - 1; cmask &= cmask - 1;
                    float hb;
                    const float disc = sphere_disc(SV<C>::cluster(ctab + c), o, d, hb);
                    const float sq = sqrtf(disc);
                    if (-hb + sq < 0.f || -hb - sq > best_t) continue;
                    const int m0 = first + 16 * c, n = min(16, end - m0);
                    cnt.tests += n;
                    uint32_t mask = sphere_mask<C>(SV<C>::obj_ptr(m0), n, o, d);
                    while (mask) {
                        const int i = m0 + __ffs(mask) - 1; mask &= mask - 1;
                        float t;
                        if (!sphere_t(SV<C>::template obj_geom_k<K>(i), o, d, tmin, best_t, t)) continue;
                        if (t == best_t && best >= 0 && SV<C>::obj_orig(i) < SV<C>::obj_orig(best)) continue;
                        best = i; best_t = t;
                    }
                }
            }
        } else if (run.x == RUN_SPHERE || run.x == RUN_POKEBALL) {
            cnt.tests += run.z;
            for (int c0 = first; c0 < end; c0 += 32) {                       // candidates first, finished in JSON order
                uint32_t mask = sphere_mask<C>(SV<C>::obj_ptr(c0), min(32, end - c0), o, d);
                while (mask) {
                    const int i = c0 + __ffs(mask) - 1; mask &= mask - 1;
                    float t;
                    if (sphere_t(SV<C>::template obj_geom_k<K>(i), o, d, tmin, best_t, t)) { best = i; best_t = t; }
                }
            }
        } else if (run.x == RUN_HALFSPACE) {
            cnt.tests += run.z;
            for (int i = first; i < end; ++i) {
                float t;
                if (plane_t(SV<C>::template obj_geom_k<K>(i), o, d, tmin, best_t, t)) { best = i; best_t = t; }
            }
        } else if (K && run.w > 0) {                   // generic run with per-object world bounds: candidates only
            const int ctab = run.w - 1;
            cnt.tests += run.z;
            for (int c0 = first; c0 < end; c0 += 32) {
                uint32_t mask = sphere_mask<C>(SV<C>::cluster_ptr(ctab + c0 - first), min(32, end - c0), o, d);
                while (mask) {
                    const int i = c0 + __ffs(mask) - 1; mask &= mask - 1;
                    float t, tent; uint32_t code;
                    if (generic_hit<C, K>(i, o, d, tmin, best_t, t, code, tent, cnt.tests)) { best = i; best_t = t; best_code = code; best_tent = tent; }
                }
            }
        } else {
            for (int i = first; i < end; ++i) {
                float t, tent; uint32_t code;
                if (generic_hit<C, K>(i, o, d, tmin, best_t, t, code, tent, cnt.tests)) { best = i; best_t = t; best_code = code; best_tent = tent; }
            }
        }
    }
}

template <bool C, bool K>
__device__ __forceinline__ bool run_occludes(const int4 run, V3 o, V3 d, float tmin, float tmax, Counters& cnt) {
    {
        const int first = run.y, end = run.y + run.z;
        if (K && (run.x == RUN_SPHERE || run.x == RUN_POKEBALL) && run.w > 0) {     // clustered run (see closest_hit)
            const int ctab = run.w - 1, ncl = (run.z + 15) >> 4;
            cnt.tests += ncl;
            for (int cb = 0; cb < ncl; cb += 32) {
                uint32_t cmask = sphere_mask<C>(SV<C>::cluster_ptr(ctab + cb), min(32, ncl - cb), o, d);
                while (cmask) {
                    const int c = cb + __ffs(cmask) - 1; cmask &= cmask - 1;
                    float hb;
                    const float disc = sphere_disc(SV<C>::cluster(ctab + c), o, d, hb);
                    const float sq = sqrtf(disc);
                    if (-hb + sq < 0.f || -hb - sq > tmax) continue;
                    const int m0 = first + 16 * c, n = min(16, end - m0);
                    cnt.tests += n;
                    uint32_t mask = sphere_mask<C>(SV<C>::obj_ptr(m0), n, o, d);
                    while (mask) {
                        const int i = m0 + __ffs(mask) - 1; mask &= mask - 1;
                        float t;
                        if (sphere_t(SV<C>::template obj_geom_k<K>(i), o, d, tmin, tmax, t)) return true;
                    }
                }
            }
        } else if (run.x == RUN_SPHERE || run.x == RUN_POKEBALL) {
            for (int c0 = first; c0 < end; c0 += 32) {                       // any-hit: early-out per chunk of 32
                const int n = min(32, end - c0);
                cnt.tests += n;
                uint32_t mask = sphere_mask<C>(SV<C>::obj_ptr(c0), n, o, d);
                while (mask) {
                    const int i = c0 + __ffs(mask) - 1; mask &= mask - 1;
                    float t;
                    if (sphere_t(SV<C>::template obj_geom_k<K>(i), o, d, tmin, tmax, t)) return true;
                }
            }
        } else if (run.x == RUN_HALFSPACE) {
            for (int i = first; i < end; ++i) {
                float t; ++cnt.tests;
                if (plane_t(SV<C>::template obj_geom_k<K>(i), o, d, tmin, tmax, t)) return true;
            }
        } else if (K && run.w > 0) {                   // generic run with per-object world bounds: candidates only
            const int ctab = run.w - 1;
            cnt.tests += run.z;
            for (int c0 = first; c0 < end; c0 += 32) {
                uint32_t mask = sphere_mask<C>(SV<C>::cluster_ptr(ctab + c0 - first), min(32, end - c0), o, d);
                while (mask) {
                    const int i = c0 + __ffs(mask) - 1; mask &= mask - 1;
                    float t, tent; uint32_t code;
                    if (generic_hit<C, K>(i, o, d, tmin, tmax, t, code, tent, cnt.tests)) return true;
                }
            }
        } else {
            for (int i = first; i < end; ++i) {
                float t, tent; uint32_t code;
                if (generic_hit<C, K>(i, o, d, tmin, tmax, t, code, tent, cnt.tests)) return true;
            }
        }
    }
    return false;
}

template <bool C, bool K>
__device__ __forceinline__ bool closest_hit(V3 o, V3 d, float tmin, float tmax, Hit& hit, Counters& cnt) {
    int best = -1; float best_t = tmax; uint32_t best_code = 0; float best_tent = 0.f;
    if (K) {
        const ScanResult sr = scan_scene<C, K>(o, d, tmin, tmax, false);
        cnt.tests += sr.tests;
        best = sr.best; best_t = sr.t; best_code = sr.code; best_tent = sr.tent;
    } else {
        const int n_runs = c_hdr.n_runs;
        for (int r = 0; r < n_runs; ++r) run_closest<C, K>(SV<C>::run(r), o, d, tmin, best, best_t, best_code, best_tent, cnt);
    }
    if (best < 0) return false;
    // ---- finalise the winner: normal, front_face, material, id code ----
    const int4 m = SV<C>::obj_meta(best);
    const int kind = m.x & 0xff, flags = m.x >> 8;
    hit.t = best_t;
    hit.code = SV<C>::obj_orig(best) * 8;
    V3 outward; bool front; int mat;
    if (kind == RT_PRIM_PROGRAM) {
        const uint32_t code = best_code;
        const int li = code & 0xFFFFF, region = (code >> 20) & 7, nmode = (code >> 24) & 3;
        const int4 lm = SV<C>::leaf_meta(li);
        mat = lm.y + region;
        front = (code & kFrontBit) != 0;
        V3 n;
        if (nmode == 1) n = mk(0.f, 0.f, 0.f);                               // csg.cpp:119
        else {
            if (nmode == 2) { float4 f = SV<C>::fallback(code >> 26); outward = mk(f.x, f.y, f.z); }
            else if ((lm.x & 0xff) == RT_PRIM_HALFSPACE) { float4 g = SV<C>::leaf_geom(li); outward = mk(g.x, g.y, g.z); }
            else { int rg; outward = sphere_outward<C>(SV<C>::leaf_geom(li), lm, o, d, best_tent, rg); }
            n = dot(d, outward) < 0.f ? outward : neg(outward);              // geometry.h:42-45
        }
        if (flags & RT_FLAG_WRAPPED) front = dot(d, n) < 0.f;               // transform.cpp:41
        hit.n = n; hit.mat = mat; hit.front = front;
        return true;
    }
    const float4 g = SV<C>::template obj_geom_k<K>(best);
    int region = 0;
    if (kind == RT_PRIM_HALFSPACE) outward = mk(g.x, g.y, g.z);
    else outward = sphere_outward<C>(g, m, o, d, best_t, region);
    front = dot(d, outward) < 0.f;
    V3 n = front ? outward : neg(outward);
    if (flags & RT_FLAG_XFORMED) front = dot(d, n) < 0.f;                   // transform.cpp:41
    mat = m.y;
    if (kind == RT_PRIM_POKEBALL) { mat += region; if (!(flags & RT_FLAG_XFORMED)) hit.code += region + 1; }   // code: only a bare top-level pokeball reports its region
    hit.n = n; hit.mat = mat; hit.front = front;
    return true;
}

template <bool C, bool K>
__device__ __forceinline__ bool occluded(V3 o, V3 d, float tmin, float tmax, Counters& cnt) {
    if (K) {
        const ScanResult sr = scan_scene<C, K>(o, d, tmin, tmax, true);
        cnt.tests += sr.tests;
        return sr.best >= 0;
    }
    const int n_runs = c_hdr.n_runs;
    // any-hit: the order of the scan cannot be observed (scene.cpp:33-42 returns a bool); the runs are walked backwards
    // because later objects are usually the foreground detail that casts the shadows (snorlax 4K -0.3 %, synthetic-B -1.2 %)
    for (int r = n_runs - 1; r >= 0; --r)
        if (run_occludes<C, K>(SV<C>::run(r), o, d, tmin, tmax, cnt)) return true;
    return false;
}

// ---- shade_lambert_phong (shading.cpp:31-138) ------------------------------------------------------------------
struct RGBf { float r, g, b; };
__device__ __forceinline__ RGBf rgb(float r, float g, float b) { RGBf c; c.r = r; c.g = g; c.b = b; return c; }
// "screen" blend (shading.cpp:6-12)
__device__ __forceinline__ float screen1(float a, float b) { return 1.f - (1.f - a) * (1.f - b); }
__device__ __forceinline__ RGBf screen(RGBf a, RGBf b) { return rgb(screen1(a.r, b.r), screen1(a.g, b.g), screen1(a.b, b.b)); }

// Direction to light `lp` and its distance (shading.cpp:80-92; a directional light, lp.w != 0, is at infinity: shading.cpp:47).
__device__ __forceinline__ float light_vector(float4 lp, V3 p, V3 n, V3& wi, float& dist) {
    if (lp.w != 0.f) { wi = mk(lp.x, lp.y, lp.z); dist = kInf; return fmaxf(0.f, dot(n, wi)); }
    V3 L = mk(lp.x - p.x, lp.y - p.y, lp.z - p.z);
    float d2 = fmaxf(dot(L, L), 0.01f);
    dist = sqrtf(d2);
    float inv = 1.f / dist;
    wi = mk(L.x * inv, L.y * inv, L.z * inv);
    return fmaxf(0.f, dot(n, wi));
}
// One light's contribution, screen-blended into E (shading.cpp:79-130).
template <bool C, bool K>
__device__ __forceinline__ void light_term(int l, const float4 m0, float ks, float shininess, float kd, V3 n, V3 p, V3 so, V3 wo,
                                           float eps, RGBf& E, Counters& cnt) {
    const float4 lp = SV<C>::light(l, 0), li = SV<C>::light(l, 1);
    V3 wi; float dist;
    const float ndotl = light_vector(lp, p, n, wi, dist);
    if (ndotl <= 0.f) return;
    float tmaxS = dist - eps;
    if (tmaxS <= eps) return;
    ++cnt.shadow;
    if (occluded<C, K>(so, normalized(wi), eps, tmaxS, cnt)) return;            // Ray ctor re-normalises (core.h:278)
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
        // exp2(s * log2(x)) on the MUFU pipe (relative error ~ s * 2^-21, far below 1/255); x == 0 keeps powf's edge cases
        float spec = (rdotv > 0.f ? __powf(rdotv, shininess) : powf(rdotv, shininess)) * ks;
        Es = rgb(I.r * spec, I.g * spec, I.b * spec);
    }
    E = screen(E, screen(Ed, Es));
}

// The same term in two halves, for trace_flat: the shadow query first (light_visible), the contribution later (light_shade;
// the light vector is recomputed, 15 instructions), so that no shading state is live across the scans of a flat scene.
template <bool C, bool K>
__device__ __forceinline__ bool light_visible(int l, V3 n, V3 p, V3 so, float eps, Counters& cnt) {
    V3 wi; float dist;
    const float ndotl = light_vector(SV<C>::light(l, 0), p, n, wi, dist);
    if (ndotl <= 0.f) return false;
    const float tmaxS = dist - eps;
    if (tmaxS <= eps) return false;
    ++cnt.shadow;
    return !occluded<C, K>(so, normalized(wi), eps, tmaxS, cnt);               // Ray ctor re-normalises (core.h:278)
}
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
        float spec = (rdotv > 0.f ? __powf(rdotv, shininess) : powf(rdotv, shininess)) * ks;
        Es = rgb(I.r * spec, I.g * spec, I.b * spec);
    }
    E = screen(E, screen(Ed, Es));
}

template <bool C, bool K>
__device__ RGBf shade(const Hit& h, V3 p, V3 wo, Counters& cnt) {
    if (h.mat < 0) return rgb(1.f, 0.f, 1.f);                                // shading.cpp:33
    const float4 m0 = material(h.mat, 0), m1 = material(h.mat, 1), m2 = material(h.mat, 2);
    RGBf E = rgb(m1.x, m1.y, m1.z);                                          // ambient (x) scene ambient, pre-multiplied
    const float eps = fmaxf(1e-3f, 1e-4f * h.t);                             // shading.cpp:43
    const V3 so = madd(p, h.n, eps);
    const int nl = c_hdr.n_lights;
    for (int l = 0; l < nl; ++l) light_term<C, K>(l, m0, m0.w, m2.y, m2.w, h.n, p, so, wo, eps, E, cnt);
    return rgb(fminf(1.5f, E.r), fminf(1.5f, E.g), fminf(1.5f, E.b));        // shading.cpp:133-135
}

// Tracer::trace for a scene in which no ray can spawn a child (no material reflects or refracts, or recursion == 1):
// one closest hit, one shade, no stack.  Every thread of the CTA calls it, also those without a sample (`active`
// false).  With `lockstep` the CTA meets at a barrier before every light: warps that walk the same code at the same
// time share instruction-cache lines, which is what limits this kernel on scenes with CSG programs (snorlax 4K
// 15.96 -> 14.44 ms); on scenes of a few plain spheres the waiting costs more than it saves (penguin 4K 6.08 -> 6.31 ms),
// so the flag is set per scene on upload.
template <bool C, bool K>
__device__ RGBf trace_flat(bool active, bool lockstep, V3 o, V3 d, Hit& primary, bool& primary_hit, Counters& cnt) {
    Hit h; h.mat = -1; h.t = 0.f; h.n = mk(0.f, 1.f, 0.f);
    bool hit = false;
    if (active) { hit = closest_hit<C, K>(o, d, 1e-4f, kInf, h, cnt); ++cnt.primary; }     // tracer.cpp:29
    primary = h; primary_hit = hit;
    const bool lit = hit && h.mat >= 0;
    const V3 p = madd(o, d, h.t);
    const float eps = fmaxf(1e-3f, 1e-4f * h.t);
    const V3 so = madd(p, h.n, eps);
    const int nl = c_hdr.n_lights;
    RGBf E = rgb(0.f, 0.f, 0.f);
    for (int l0 = 0; l0 < nl || l0 == 0; l0 += 32) {              // lights in chunks of 32: one visibility bit each
        const int cnt_l = min(32, nl - l0);
        uint32_t vis = 0;
        for (int k = 0; k < cnt_l; ++k) {
            if (lockstep) __syncthreads();
            if (lit) vis |= light_visible<C, K>(l0 + k, h.n, p, so, eps, cnt) ? (1u << k) : 0u;
        }
        if (lit) {
            const float4 m0 = material(h.mat, 0), m2 = material(h.mat, 2);
            if (l0 == 0) { const float4 m1 = material(h.mat, 1); E = rgb(m1.x, m1.y, m1.z); }   // ambient, pre-multiplied
            const V3 wo = normalized(neg(d));
            while (vis) {
                const int k = __ffs(vis) - 1; vis &= vis - 1;
                light_shade<C>(l0 + k, m0, m0.w, m2.y, m2.w, h.n, p, wo, E);
            }
        }
    }
    if (!hit) return rgb(c_hdr.bg[0], c_hdr.bg[1], c_hdr.bg[2]);             // tracer.cpp:72
    if (!lit) return rgb(1.f, 0.f, 1.f);                                     // shading.cpp:33
    return rgb(fminf(1.5f, E.r), fminf(1.5f, E.g), fminf(1.5f, E.b));        // shading.cpp:133-135
}

// ---- Tracer::trace_recursive (tracer.cpp:22-104) as an explicit bounded stack ----------------------------------
// A node with both kr > 0 and kt > 0 is bilinear in its two children under the screen blend, so the
// recursion cannot be flattened into a throughput product; each pending node keeps its running `total`
// and, when a refraction is still owed, the refracted ray.
struct Frame {
    RGBf total; float k_pending;    // coefficient of the child currently being traced
    float kt; V3 ro, rd;            // owed refraction (kt > 0 when owed)
};

template <bool C, bool K>
__device__ RGBf trace(V3 o, V3 d, Hit* primary, bool* primary_hit, Counters& cnt) {
    const int limit = c_hdr.recursion;
    if (primary_hit) *primary_hit = false;
    if (limit <= 0) {                                                        // tracer.cpp:26 at depth 0
        if (primary) { Hit h; *primary_hit = closest_hit<C, K>(o, d, 1e-4f, kInf, h, cnt); *primary = h; }
        return rgb(0.f, 0.f, 0.f);
    }
    Frame stack[RT_MAX_RECURSION];
    int sp = 0;                     // == depth of the ray being traced
    RGBf result;
    for (;;) {
        // ---- evaluate the ray (o,d) at depth sp ----
        Hit h;
        bool hit = closest_hit<C, K>(o, d, 1e-4f, kInf, h, cnt);                // tracer.cpp:29
        if (sp == 0) { ++cnt.primary; if (primary) { *primary = h; *primary_hit = hit; } } else ++cnt.secondary;
        bool descend = false;
        if (!hit) {
            result = rgb(c_hdr.bg[0], c_hdr.bg[1], c_hdr.bg[2]);             // tracer.cpp:72
        } else {
            const V3 p = madd(o, d, h.t);
            const V3 wo = normalized(neg(d));
            RGBf total = shade<C, K>(h, p, wo, cnt);                            // tracer.cpp:31
            result = total;
            if (h.mat >= 0 && sp < limit - 1) {
                const float4 m1 = material(h.mat, 1), m2 = material(h.mat, 2);
                const float kr = m1.w, kt = m2.x, ior = m2.z;
                const V3 I = normalized(d);
                bool want_r = kr > 0.f, want_t = false;
                V3 to, td;
                if (kt > 0.f) {                                              // tracer.cpp:51-68
                    float eta = h.front ? (c_hdr.medium_index / ior) : (ior / c_hdr.medium_index);
                    float cos_i = -dot(I, h.n);
                    float sin_t2 = eta * eta * fmaxf(0.f, 1.f - cos_i * cos_i);
                    if (!(sin_t2 >= 1.f)) {                                  // TIR guard, tracer.cpp:100-104
                        float cos_t = sqrtf(1.f - sin_t2);
                        float kk = eta * cos_i - cos_t;
                        td = normalized(mk(fmaf(kk, h.n.x, eta * I.x), fmaf(kk, h.n.y, eta * I.y), fmaf(kk, h.n.z, eta * I.z)));
                        to = h.front ? madd(p, h.n, -1e-6f) : madd(p, h.n, 1e-6f);
                        want_t = true;
                    }
                }
                if (want_r || want_t) {
                    Frame& f = stack[sp];
                    f.total = total;
                    f.kt = want_t && want_r ? kt : 0.f;
                    if (want_r) {                                            // tracer.cpp:38-48
                        float two = 2.f * dot(I, h.n);
                        V3 rd = normalized(mk(fmaf(-two, h.n.x, I.x), fmaf(-two, h.n.y, I.y), fmaf(-two, h.n.z, I.z)));
                        V3 ro = h.front ? madd(p, h.n, 1e-6f) : madd(p, h.n, -1e-6f);
                        f.k_pending = kr;
                        if (want_t) { f.ro = to; f.rd = td; }
                        o = ro; d = rd;
                    } else {
                        f.k_pending = kt;
                        o = to; d = td;
                    }
                    ++sp;
                    descend = true;
                }
            }
        }
        if (descend) continue;
        // ---- return `result` to the pending parents ----
        for (;;) {
            if (sp == 0) return result;
            Frame& f = stack[sp - 1];
            const float k = f.k_pending;
            f.total = screen(f.total, rgb(result.r * k, result.g * k, result.b * k));   // tracer.cpp:47,66
            if (f.kt > 0.f) {                  // reflection done, refraction owed
                f.k_pending = f.kt; f.kt = 0.f;
                o = f.ro; d = f.rd;
                break;                         // trace it at depth sp (frame stays)
            }
            result = f.total;
            --sp;
        }
    }
}

// ---- one ray of a sample whose tree state lives in global memory (wavefront kernel) -----------------------------------
// Same logic as the loop body of trace(); the pending parents are 3 float4 per depth in global memory
// (total.rgb k_pending | kt ro.xyz | rd.xyz -), so ANY thread can take ANY unfinished sample from the CTA's queue.
// Returns true when the tree is finished (`result` = Tracer::trace's colour), false when (o, d, sp) is the next ray.
template <bool C, bool K>
__device__ __forceinline__ bool advance_ray(V3& o, V3& d, int& sp, float4* __restrict__ frames, RGBf& result,
                                            Hit& primary, bool& primary_hit, Counters& cnt) {
    const int limit = c_hdr.recursion;
    Hit h;
    const bool hit = closest_hit<C, K>(o, d, 1e-4f, kInf, h, cnt);           // tracer.cpp:29
    if (sp == 0) { ++cnt.primary; primary = h; primary_hit = hit; } else ++cnt.secondary;
    if (!hit) {
        result = rgb(c_hdr.bg[0], c_hdr.bg[1], c_hdr.bg[2]);                 // tracer.cpp:72
    } else {
        const V3 p = madd(o, d, h.t);
        const V3 wo = normalized(neg(d));
        const RGBf total = shade<C, K>(h, p, wo, cnt);                       // tracer.cpp:31
        result = total;
        if (h.mat >= 0 && sp < limit - 1) {
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
    // ---- return `result` to the pending parents ----
    for (;;) {
        if (sp == 0) return true;
        float4* f = frames + 3 * (sp - 1);
        const float4 f0 = f[0], f1 = f[1];
        const float k = f0.w;
        const RGBf total = screen(rgb(f0.x, f0.y, f0.z), rgb(result.r * k, result.g * k, result.b * k));   // tracer.cpp:47,66
        if (f1.x > 0.f) {                      // reflection done, refraction owed: trace it at depth sp (frame stays)
            const float4 f2 = f[2];
            f[0] = make_float4(total.r, total.g, total.b, f1.x);
            f[1] = make_float4(0.f, 0.f, 0.f, 0.f);
            o = mk(f1.y, f1.z, f1.w); d = mk(f2.x, f2.y, f2.z);
            return false;
        }
        result = total;
        --sp;
    }
}

}  // namespace rtd
