// mt19937.cuh — the reference's jitter stream (tracer.cpp:284-285,293) as a block-parallel generator.
//
// The reference consumes ONE std::mt19937(12345) sequentially over the whole frame, 4 raw
// 32-bit outputs per sample (libstdc++ uniform_real_distribution<double> draws two per double).
// Sample s of loop-pixel p therefore uses raw outputs 4*(8p+s) .. +3, whatever the scene.
// MT19937's state advances in blocks of 624 outputs ("twists"); a render tile is 8 twists
// = 4992 outputs = 1248 samples = 156 pixels, and the state at every tile boundary is kept in a
// checkpoint table in HBM (2496 B per tile) that is generated once per process and extended on
// demand — the stream does not depend on the scene or on the resolution.
//
// In-place MT19937 recurrence (Matsumoto & Nishimura 1998): for k = 0..623
//     s[k] = s[(k+397) % 624] ^ g(s[k], s[(k+1) % 624]),  g(u,v) = (y >> 1) ^ (y & 1 ? 0x9908b0df : 0),
//     y = (u & 0x80000000) | (v & 0x7fffffff)
// where indices already visited hold NEW values.  Substituting those gives every new word as a
// function of OLD words only, so a whole twist is one data-parallel step (one barrier, double buffer).
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define MT_HD __host__ __device__ __forceinline__
#else
#define MT_HD inline
#endif

namespace mt {

constexpr int kN = 624, kM = 397;
constexpr int kTwistsPerTile = 8;
constexpr int kOutputsPerTile = kN * kTwistsPerTile;     // 4992
constexpr int kSamplesPerTile = kOutputsPerTile / 4;     // 1248
constexpr int kPixelsPerTile = kSamplesPerTile / 8;      // 156  (== RT_TILE_PIXELS)

MT_HD uint32_t g(uint32_t u, uint32_t v) {
    uint32_t y = (u & 0x80000000u) | (v & 0x7fffffffu);
    return (y >> 1) ^ ((y & 1u) ? 0x9908b0dfu : 0u);
}
// new word k of the next state, from the old state only
MT_HD uint32_t next_word(const uint32_t* old, int k) {
    if (k < 227) return old[k + 397] ^ g(old[k], old[k + 1]);
    if (k < 454) return old[k + 170] ^ g(old[k - 227], old[k - 226]) ^ g(old[k], old[k + 1]);
    if (k < 623) return old[k - 57] ^ g(old[k - 454], old[k - 453]) ^ g(old[k - 227], old[k - 226]) ^ g(old[k], old[k + 1]);
    // k == 623: s[623] = new[396] ^ g(old[623], new[0])
    uint32_t new0 = old[397] ^ g(old[0], old[1]);
    uint32_t new169 = old[566] ^ g(old[169], old[170]);
    uint32_t new396 = new169 ^ g(old[396], old[397]);
    return new396 ^ g(old[623], new0);
}
// One twist by the first 227 threads of a CTA, one barrier: thread t produces new[t], new[t + 227] and new[t + 454] — each
// from the one before it, which it still holds in a register — so every word costs ONE g() instead of the 1 / 2 / 3 of
// next_word (the render kernels' tile prologue: 1.9 -> 1 g per word).  Thread 169 also owns word 623.  `emit(k, w)` stores.
template <class Emit>
MT_HD void twist_by_thread(const uint32_t* old, int t, Emit emit) {
    if (t >= 227) return;
    const uint32_t n0 = old[t + 397] ^ g(old[t], old[t + 1]);
    const uint32_t n1 = n0 ^ g(old[t + 227], old[t + 228]);
    emit(t, n0); emit(t + 227, n1);
    if (t < 169) emit(t + 454, n1 ^ g(old[t + 454], old[t + 455]));
    if (t == 169) emit(623, n1 ^ g(old[623], old[397] ^ g(old[0], old[1])));      // s[623] = new[396] ^ g(old[623], new[0])
}
MT_HD uint32_t temper(uint32_t y) {
    y ^= y >> 11; y ^= (y << 7) & 0x9d2c5680u; y ^= (y << 15) & 0xefc60000u; y ^= y >> 18;
    return y;
}
// std::mt19937 seeding (init_genrand)
inline void seed_state(uint32_t seed, uint32_t* s) {
    s[0] = seed;
    for (int i = 1; i < kN; ++i) s[i] = 1812433253u * (s[i - 1] ^ (s[i - 1] >> 30)) + (uint32_t)i;
}
// ---- jump ahead (see csrc/mt_jump_gen.cpp) -----------------------------------------------------------------------
// kJumpPoly[m] = t^(624 * 2^m) mod phi: the state 2^m twists after state s is XOR_{i : bit i set} T^i s, and T^i s is
// the window x[i .. i+623] of the word sequence that starts with s.  All coefficients are below 19937, so the first 33
// blocks of the sequence suffice.  (Word 0 of a state only contributes its top bit; the low 31 bits of the jumped
// state's word 0 are therefore meaningless, exactly like those of a seeded state, and nothing reads them.)
constexpr int kJumpPolys = 32;
constexpr int kJumpBlocks = 33;
extern const uint32_t kJumpPoly[kJumpPolys][kN];     // generated at build time (build/mt_jump_table.cpp)
MT_HD uint32_t jump_word(const uint32_t* x /* kJumpBlocks * kN words */, const uint32_t* poly /* kN words */, int w) {
    uint32_t acc = 0;
    for (int pw = 0; pw < kN; ++pw) {
        uint32_t bits = poly[pw];
        while (bits) {
#if defined(__CUDA_ARCH__)
            const int b = __ffs((int)bits) - 1;
#else
            const int b = __builtin_ctz(bits);
#endif
            bits &= bits - 1;
            acc ^= x[32 * pw + b + w];
        }
    }
    return acc;
}
// libstdc++ generate_canonical<double,53> + uniform_real_distribution<double>(-0.5, 0.5):
// lo drawn first; (lo + hi*2^32) rounded to double, / 2^64, clamped below 1, then * 1.0 + (-0.5)
MT_HD double jitter(uint32_t lo, uint32_t hi) {
    double sum = (double)lo + (double)hi * 4294967296.0;
    double r = sum * (1.0 / 18446744073709551616.0);
    if (r >= 1.0) r = 0.99999999999999988897769753748;   // nextafter(1.0, 0.0)
    return r - 0.5;
}

}  // namespace mt
