// Host-side equivalence check of csg_combine's shortcuts against the literal sweep (csrc/rt_csg.cuh):
// exhaustive over a grid of adversarial interval ends (ties within the comparator's tolerance, ends at / around the ray
// origin, negative and inverted intervals, +-inf, NaN) for the three operations, plus random intervals.
// Prints "<cases> <mismatches>"; test infrastructure only (tests/test_csg_shortcuts.py builds and runs it).
#include <cstdio>
#include <cstring>
#include <random>
#include <vector>

#include "../../raytracing-project_b200/csrc/rt_csg.cuh"

using rtd::Ivl;

static bool same(const Ivl& x, const Ivl& y) {
    if (x.ok != y.ok) return false;
    if (!x.ok) return true;
    return std::memcmp(&x.t0, &y.t0, 4) == 0 && std::memcmp(&x.t1, &y.t1, 4) == 0 && x.c0 == y.c0 && x.c1 == y.c1;
}

int main() {
    const float inf = rtd::kInf, nan = __builtin_nanf("");
    const std::vector<float> grid = {-inf, -3.f, -1.f, -1e-6f, -5e-7f, -0.f, 0.f, 4e-7f, 9.9e-7f, 1e-6f, 1.1e-6f, 2e-6f, 0.5f,
                                     0.5f + 6e-8f, 0.5000005f, 0.500001f, 0.5000011f, 0.500002f, 1.f, 1.f + 1e-6f, 2.f, 7.5f, inf, nan};
    long long cases = 0, bad = 0, fast = 0;
    auto check = [&](int op, const Ivl& A, const Ivl& B) {
        const Ivl want = rtd::csg_sweep(op, A, B), got = rtd::csg_combine(op, A, B);
        ++cases;
        if (!same(want, got)) {
            if (++bad <= 10)
                std::printf("# op %d A(%d %.9g %.9g %x %x) B(%d %.9g %.9g %x %x): sweep (%d %.9g %.9g %x %x) combine (%d %.9g %.9g %x %x)\n", op,
                            A.ok, A.t0, A.t1, A.c0, A.c1, B.ok, B.t0, B.t1, B.c0, B.c1, want.ok, want.t0, want.t1, want.c0, want.c1,
                            got.ok, got.t0, got.t1, got.c0, got.c1);
        }
    };
    // codes: distinct leaves, front bits, a zero-normal mode and a fallback, so every select is observable
    const uint32_t ca0 = 1u | rtd::kFrontBit, ca1 = 1u | (3u << 20), cb0 = 2u | rtd::kFrontBit | (2u << 24) | (5u << 26), cb1 = 2u | rtd::kFrontBit | (1u << 24);
    for (int op = 0; op < 3; ++op)
        for (float a0 : grid) for (float a1 : grid) for (float b0 : grid) for (float b1 : grid)
            for (int oks = 0; oks < 4; ++oks) {
                Ivl A{a0, a1, ca0, ca1, (oks & 1) != 0}, B{b0, b1, cb0, cb1, (oks & 2) != 0};
                check(op, A, B);
            }
    std::mt19937 rng(20261104);
    std::uniform_real_distribution<float> U(-2.f, 6.f), W(0.f, 3.f);
    std::uniform_int_distribution<int> J(-3, 3);
    for (int i = 0; i < 3000000; ++i) {
        Ivl A{0, 0, (uint32_t)rng() & 0x0FFFFFFFu, (uint32_t)rng() & 0x0FFFFFFFu, true}, B{0, 0, (uint32_t)rng() & 0x0FFFFFFFu, (uint32_t)rng() & 0x0FFFFFFFu, true};
        A.t0 = U(rng); A.t1 = A.t0 + W(rng); B.t0 = U(rng); B.t1 = B.t0 + W(rng);
        if (i % 3 == 0) B.t0 = A.t1 + J(rng) * 4e-7f;            // near-ties across the children
        if (i % 5 == 0) B.t1 = A.t0 + J(rng) * 4e-7f;
        if (i % 7 == 0) A.t0 = J(rng) * 5e-7f;                   // ends around the origin test
        if (i % 11 == 0) std::swap(B.t0, B.t1);                   // inverted child
        const bool ordered = A.t0 <= A.t1 && B.t0 <= B.t1 && fabsf(A.t0 - B.t0) > 1e-6f && fabsf(A.t0 - B.t1) > 1e-6f &&
                             fabsf(A.t1 - B.t0) > 1e-6f && fabsf(A.t1 - B.t1) > 1e-6f;
        fast += ordered;
        check(i % 3, A, B);
    }
    std::printf("%lld %lld %lld\n", cases, bad, fast);
    return bad ? 1 : 0;
}
