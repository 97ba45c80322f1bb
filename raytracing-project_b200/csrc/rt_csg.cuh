// rt_csg.cuh — CSG::interval (csg.cpp:61-163) on interval ends kept in registers.  Pure arithmetic on (t, code) pairs: no
// scene access, no CUDA intrinsics, so the same code compiles for the device (included by rt_device.cuh) and, for the
// host-side equivalence test of its shortcuts (tests/test_csg_shortcuts.py), with a plain C++ compiler.
#pragma once
#include <math.h>
#include <stdint.h>

#include "../../include/rt_b200.h"

// A/B build switch: 0 = no ordered-case shortcut, the sweep inlined at every call site (the previous behaviour)
#ifndef RT_CSG_SHORTCUT
#define RT_CSG_SHORTCUT 1
#endif
#ifdef __CUDACC__
#define RT_CSG_FN __device__ __forceinline__
#if RT_CSG_SHORTCUT
#define RT_CSG_COLD __device__ __noinline__
#else
#define RT_CSG_COLD __device__ __forceinline__
#endif
#else
#define RT_CSG_FN inline
#define RT_CSG_COLD inline
#endif

namespace rtd {
#ifndef __CUDACC__
constexpr float kInf = __builtin_huge_valf();
#endif

// ---- CSG programs (csg.cpp:61-163) ------------------------------------------------------------------
// An interval end is (t, code): code = leaf | region<<20 | front<<23 | nmode<<24 | fallback<<26,
// nmode 0 = geometric normal of the leaf at t, 1 = zero normal (csg.cpp:119), 2 = fallbacks[fallback].
// Normals are evaluated lazily, only for the end that becomes the visible hit.
struct Ivl { float t0, t1; uint32_t c0, c1; int ok; };   // ok: 0 / 1 (an int: a bool member costs byte packing at every call boundary)
constexpr uint32_t kFrontBit = 1u << 23;
constexpr uint32_t kNModeMask = 3u << 24;

RT_CSG_FN bool csg_in(int op, bool a, bool b) {
    return op == RT_CSG_UNION ? (a || b) : op == RT_CSG_INTERSECTION ? (a && b) : (a && !b);
}
// event_less (csg.cpp:83-88); k = type | who<<1 (type 0 Enter, 1 Exit; who 0 A, 1 B)
RT_CSG_FN bool ev_less(float ta, int ka, float tb, int kb) {
    if (fabsf(ta - tb) > 1e-6f) return ta < tb;
    if ((ka & 1) != (kb & 1)) return (ka & 1) == 0;
    return (ka >> 1) < (kb >> 1);
}

// The synthetic entry CSG::interval records when the combined solid already contains the ray origin
// (csg.cpp:92-96,114-122): t = 0, zero normal, front_face, material of the child's entry hit.
RT_CSG_FN void origin_rewrite(Ivl& v) {
    if (v.t0 < 1e-6f && v.t1 > 1e-6f) {
        v.t0 = 0.f;
        v.c0 = (v.c0 & ~(kNModeMask | (63u << 26))) | kFrontBit | (1u << 24);
    }
}
// CSG::interval (csg.cpp:61-163) on two child intervals, literally: the <= 4 finite ends are sorted with the reference's
// tolerance comparator by libstdc++'s insertion sort, then swept for the first inside run.  Every case goes through here
// unless csg_combine below proves a shortcut; out of line (one copy, arguments by value) because the shortcuts take
// nearly all calls.
RT_CSG_COLD Ivl csg_sweep(int op, Ivl A, Ivl B) {
    Ivl R; R.ok = false; R.t0 = 0.f; R.t1 = kInf; R.c0 = R.c1 = 0;
    float et[4]; int ek[4]; int n = 0;
    // finite ends only (csg.cpp:76-81), initial order A-enter, A-exit, B-enter, B-exit
    {
        const float ts[4] = {A.t0, A.t1, B.t0, B.t1};
        const bool ok[4] = {A.ok != 0, A.ok != 0, B.ok != 0, B.ok != 0};
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            bool use = ok[i] && isfinite(ts[i]);
            // compact without dynamic indexing
#pragma unroll
            for (int k = 0; k < 4; ++k) if (use && k == n) { et[k] = ts[i]; ek[k] = i; }
            n += use ? 1 : 0;
        }
    }
    // libstdc++ insertion sort (std::sort on <= 16 elements), literal: comparator is not a strict weak order
#pragma unroll
    for (int i = 1; i < 4; ++i) {
        if (i < n) {
            float vt = et[i]; int vk = ek[i];
            if (ev_less(vt, vk, et[0], ek[0])) {
#pragma unroll
                for (int k = i; k > 0; --k) { et[k] = et[k - 1]; ek[k] = ek[k - 1]; }
                et[0] = vt; ek[0] = vk;
            } else {
                bool moving = true;
#pragma unroll
                for (int k = i; k > 0; --k) {
                    if (moving && ev_less(vt, vk, et[k - 1], ek[k - 1])) { et[k] = et[k - 1]; ek[k] = ek[k - 1]; }
                    else if (moving) { et[k] = vt; ek[k] = vk; moving = false; }
                }
                // (the unguarded loop always stops at k >= 1 because !less(val, first))
            }
        }
    }
    bool inA = A.ok && (A.t0 < 1e-6f) && (A.t1 > 1e-6f);       // csg.cpp:92-96
    bool inB = B.ok && (B.t0 < 1e-6f) && (B.t1 > 1e-6f);
    bool inR = csg_in(op, inA, inB);
    bool haveEnter = false, done = false;
    float tEnt = 0.f, tExt = kInf; uint32_t cEnt = 0, cExt = 0;
    if (inR) {                                                 // csg.cpp:114-122: entry at t=0, zero normal, front
        haveEnter = true;
        uint32_t src = inA ? A.c0 : B.c0;
        cEnt = (src & ~(kNModeMask | (63u << 26))) | kFrontBit | (1u << 24);
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {                              // csg.cpp:125-156
        if (i < n && !done) {
            const int k = ek[i];
            const bool before = inR;
            const bool enter = (k & 1) == 0;
            if ((k >> 1) == 0) inA = enter; else inB = enter;
            const bool after = csg_in(op, inA, inB);
            const uint32_t code = k == 0 ? A.c0 : k == 1 ? A.c1 : k == 2 ? B.c0 : B.c1;
            // A \ B: a boundary contributed by B has its normal negated and re-oriented (csg.cpp:140-142,148-150):
            // the oriented normal is unchanged, front_face becomes false.
            if (!before && after) {
                haveEnter = true; tEnt = et[i];
                cEnt = (op == RT_CSG_DIFFERENCE && k == 3) ? (code & ~kFrontBit) : code;
            } else if (before && !after) {
                tExt = et[i];
                cExt = (op == RT_CSG_DIFFERENCE && k == 2) ? (code & ~kFrontBit) : code;
                done = true;
            }
            inR = after;
        }
    }
    if (!haveEnter || !isfinite(tExt)) return R;              // csg.cpp:158
    R.ok = true; R.t0 = tEnt; R.t1 = tExt; R.c0 = cEnt; R.c1 = cExt;
    return R;
}

// csg_combine = csg_sweep with two shortcuts whose results are identical to the sweep's:
//  * one child missed and the other is a finite ordered interval;
//  * ORDERED case: both children hit, all four ends finite, t0 <= t1 in each child and no end of A within the comparator's
//    tolerance (1e-6) of an end of B.  Then ev_less orders the four events like the real numbers (a child's own tie
//    stays enter-before-exit), so the insertion sort returns the unique sorted order, and the sweep's state around each
//    event follows from comparisons alone: when the sweep reaches an event of A at time x it has processed exactly B's
//    events with t < x, so B's flag is  (b1 < x) ? false : (b0 < x) ? true : <B contains the origin>  (and vice versa).
//    The run starts at t = 0 if the combination contains the origin (csg.cpp:114-122), else at the earliest event that
//    switches it on, and ends at the earliest event that switches it off; events of one kind are never within tolerance
//    of each other here, so "earliest" is a plain minimum.
RT_CSG_FN Ivl csg_combine(int op, const Ivl& A, const Ivl& B) {
    {
        Ivl R; R.ok = false; R.t0 = 0.f; R.t1 = kInf; R.c0 = R.c1 = 0;
        if (!A.ok && !B.ok) return R;
        if (A.ok != B.ok) {
            // One child missed.  With a finite, ordered interval X = [t0,t1] of the other child the sweep
            // reduces to: Intersection -> empty; Difference with only B -> empty; otherwise X itself, except that
            // an X containing the ray origin gets the reference's synthetic entry at t=0 (csg.cpp:114-122).
            const Ivl& X = A.ok ? A : B;
            if (isfinite(X.t0) && isfinite(X.t1) && X.t0 <= X.t1) {
                if (op == RT_CSG_INTERSECTION || (op == RT_CSG_DIFFERENCE && !A.ok)) return R;
                R = X;
                origin_rewrite(R);
                return R;
            }
        }
    }
    const float a0 = A.t0, a1 = A.t1, b0 = B.t0, b1 = B.t1;
    const bool ordered = A.ok && B.ok && fabsf(a0) < kInf && fabsf(a1) < kInf && fabsf(b0) < kInf && fabsf(b1) < kInf &&
                         a0 <= a1 && b0 <= b1 && fabsf(a0 - b0) > 1e-6f && fabsf(a0 - b1) > 1e-6f &&
                         fabsf(a1 - b0) > 1e-6f && fabsf(a1 - b1) > 1e-6f;
    if (RT_CSG_SHORTCUT && ordered) {
        const bool isU = op == RT_CSG_UNION, nb = op == RT_CSG_DIFFERENCE;     // A \ B = A and (not B)
        const bool iA = a0 < 1e-6f && a1 > 1e-6f, iB = b0 < 1e-6f && b1 > 1e-6f;   // csg.cpp:92-96
        // the other child's flag when the sweep reaches each event (B's already xor-ed with `nb`)
        const bool bAt0 = ((b1 < a0) ? false : (b0 < a0) ? true : iB) != nb;
        const bool bAt1 = ((b1 < a1) ? false : (b0 < a1) ? true : iB) != nb;
        const bool aAt0 = (a1 < b0) ? false : (a0 < b0) ? true : iA;
        const bool aAt1 = (a1 < b1) ? false : (a0 < b1) ? true : iA;
        // state before / after each event: f(a, b) = isU ? a || b : a && b
        const bool upA0 = isU ? (!iA && !bAt0) : (!iA && bAt0);                // A enters: false -> true
        const bool dnA1 = isU ? !bAt1 : bAt1;                                  // A exits:  true -> false
        const bool bOld0 = iB != nb, bNew0 = !nb;                              // B enters: its flag goes iB -> true
        const bool bef0 = isU ? (aAt0 || bOld0) : (aAt0 && bOld0), aft0 = isU ? (aAt0 || bNew0) : (aAt0 && bNew0);
        const bool bOld1 = !nb, bNew1 = nb;                                    // B exits: true -> false
        const bool bef1 = isU ? (aAt1 || bOld1) : (aAt1 && bOld1), aft1 = isU ? (aAt1 || bNew1) : (aAt1 && bNew1);
        const bool upB0 = !bef0 && aft0, dnB0 = bef0 && !aft0, upB1 = !bef1 && aft1, dnB1 = bef1 && !aft1;
        const bool inR = isU ? (iA || (iB != nb)) : (iA && (iB != nb));
        Ivl R; R.ok = false; R.t0 = 0.f; R.t1 = kInf; R.c0 = R.c1 = 0;
        // exit: earliest switch-off among A's exit and B's exit (or B's entry under difference, csg.cpp:148-150)
        const bool dnB = dnB0 || dnB1;
        const float tB = dnB0 ? b0 : b1;
        const uint32_t cB = dnB0 ? (B.c0 & ~kFrontBit) : B.c1;
        if (!dnA1 && !dnB) return R;                                           // csg.cpp:158: no exit
        const bool exitA = dnA1 && (!dnB || a1 < tB);
        const float tExt = exitA ? a1 : tB;
        const uint32_t cExt = exitA ? A.c1 : cB;
        float tEnt; uint32_t cEnt;
        if (inR) {                                                             // csg.cpp:114-122
            tEnt = 0.f;
            cEnt = ((iA ? A.c0 : B.c0) & ~(kNModeMask | (63u << 26))) | kFrontBit | (1u << 24);
        } else {
            const bool upB = upB0 || upB1;
            if (!upA0 && !upB) return R;                                       // csg.cpp:158: no entry
            const float uB = upB0 ? b0 : b1;
            const uint32_t kB = upB0 ? B.c0 : (B.c1 & ~kFrontBit);             // csg.cpp:140-142
            const bool entA = upA0 && (!upB || a0 < uB);
            tEnt = entA ? a0 : uB;
            cEnt = entA ? A.c0 : kB;
        }
        R.ok = true; R.t0 = tEnt; R.t1 = tExt; R.c0 = cEnt; R.c1 = cExt;
        return R;
    }
    return csg_sweep(op, A, B);
}

// RT_OP_XFIX: what a transform node does to its child's interval besides moving geometry
// (transform.cpp:57-86, 140-172, 225-255): ends re-projected from their hit point (non-finite -> 0),
// zero normals replaced by normalized()'s fallback under Scaling/Rotation, front_face re-derived.
RT_CSG_FN uint32_t xfix_code(uint32_t c, uint32_t arg) {
    uint32_t nmode = (c >> 24) & 3u;
    if (nmode == 1u && arg != 0xFFFFFu) { c = (c & ~(kNModeMask | (63u << 26))) | (2u << 24) | (arg << 26); nmode = 2u; }
    return nmode == 1u ? (c & ~kFrontBit) : (c | kFrontBit);
}
RT_CSG_FN void xfix(Ivl& v, uint32_t arg) {
    if (!v.ok) return;
    if (!isfinite(v.t0)) v.t0 = 0.f;
    if (!isfinite(v.t1)) v.t1 = 0.f;
    v.c0 = xfix_code(v.c0, arg);
    v.c1 = xfix_code(v.c1, arg);
}

}  // namespace rtd
