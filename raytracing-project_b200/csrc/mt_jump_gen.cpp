// mt_jump_gen.cpp — build-time generator of the MT19937 jump polynomials (host tool, run by the Makefile).
//
// The reference draws its jitter from ONE std::mt19937(12345) over the whole frame (tracer.cpp:284-293), so the state at
// a tile boundary is "the seed state advanced by 8 * tile twists".  Walking there twist by twist is sequential (108 ms for
// a 3840x2160 frame, 3.5 s for 16384^2).  MT19937 is linear over GF(2): with T = the one-word step of the recurrence
// x[n+624] = x[n+397] ^ g(x[n], x[n+1]) on the 19937-bit state and phi its characteristic polynomial,
//     T^J s = (t^J mod phi)(T) s = XOR over the set coefficients i of T^i s,
// and T^i s is just the window x[i .. i+623] of the word sequence (Haramoto, Matsumoto, Nishimura, Panneton, L'Ecuyer,
// "Efficient jump ahead for F2-linear random number generators", 2008).  This tool computes
//   * phi by Berlekamp-Massey on the low bit of x[1], x[2], ... (phi is irreducible, so every non-zero output bit
//     sequence has it as minimal polynomial),
//   * p_m = t^(624 * 2^m) mod phi for m = 0 .. 31 by repeated squaring (jump by 2^m twists),
// and prints them as a C array (624 words each, bit i of the polynomial = bit (i & 31) of word (i >> 5)).
// Output depends on nothing but the MT19937 recurrence; csrc/mt19937.cuh applies the polynomials (host restatement for
// the CPU tests, device kernel for the checkpoint table).
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <vector>

#include "mt19937.cuh"

namespace {

constexpr int kDeg = 19937;
constexpr int kWords64 = (2 * kDeg + 64 + 63) / 64;      // room for a square

using Poly = std::vector<uint64_t>;                       // bit i = coefficient of t^i

inline int get(const Poly& p, int i) { return (int)((p[(size_t)i >> 6] >> (i & 63)) & 1u); }
inline void flip(Poly& p, int i) { p[(size_t)i >> 6] ^= (uint64_t)1 << (i & 63); }

// p ^= q << s   (q has `nq` significant words)
void xor_shifted(Poly& p, const Poly& q, int nq, int s) {
    const int ws = s >> 6, bs = s & 63;
    for (int k = 0; k < nq; ++k) {
        const uint64_t v = q[(size_t)k];
        if (!v) continue;
        p[(size_t)(k + ws)] ^= v << bs;
        if (bs) p[(size_t)(k + ws + 1)] ^= v >> (64 - bs);
    }
}

// Berlekamp-Massey over GF(2) on `n` bits; returns the connection polynomial C (c_0 = 1) and its length L:
// s[j] = XOR_{i=1..L} c_i s[j-i].
int berlekamp_massey(const std::vector<uint8_t>& s, int n, Poly& C) {
    const int words = (n + 64) / 64 + 2;
    Poly B(words, 0), Tm(words, 0), rev(words, 0);         // rev bit i = s[j - i] at step j
    C.assign(words, 0);
    C[0] = 1; B[0] = 1;
    int L = 0, m = 1;
    for (int j = 0; j < n; ++j) {
        // rev <<= 1; rev bit 0 = s[j]
        for (int k = words - 1; k > 0; --k) rev[(size_t)k] = (rev[(size_t)k] << 1) | (rev[(size_t)k - 1] >> 63);
        rev[0] = (rev[0] << 1) | (uint64_t)s[(size_t)j];
        uint64_t acc = 0;
        const int lw = (L >> 6) + 1;
        for (int k = 0; k < lw; ++k) acc ^= C[(size_t)k] & rev[(size_t)k];
        const int d = __builtin_parityll(acc);
        if (!d) { ++m; continue; }
        if (2 * L <= j) {
            Tm = C;
            xor_shifted(C, B, words - (m >> 6) - 2, m);
            L = j + 1 - L; B = Tm; m = 1;
        } else {
            xor_shifted(C, B, words - (m >> 6) - 2, m);
            ++m;
        }
    }
    return L;
}

// r = a^2 mod phi   (phi monic of degree kDeg)
void square_mod(const Poly& a, const Poly& phi, Poly& r) {
    Poly sq((size_t)kWords64 + 2, 0);
    for (int i = 0; i < kDeg; ++i) if (get(a, i)) flip(sq, 2 * i);
    const int phi_words = (kDeg + 64) / 64;
    for (int i = 2 * kDeg - 2; i >= kDeg; --i) if (get(sq, i)) xor_shifted(sq, phi, phi_words, i - kDeg);
    r.assign((size_t)kWords64 + 2, 0);
    for (int k = 0; k < (kDeg + 63) / 64; ++k) r[(size_t)k] = sq[(size_t)k];
}

}  // namespace

int main(int argc, char** argv) {
    const int n_poly = 32;
    // word sequence x[0 ..] from an arbitrary non-zero state
    const int n_bits = 2 * kDeg + 200;
    const int n_blocks = (n_bits + 1 + mt::kN - 1) / mt::kN + 1;
    std::vector<uint32_t> x((size_t)(n_blocks + 1) * mt::kN);
    mt::seed_state(5489u, x.data());
    for (int b = 0; b < n_blocks; ++b)
        for (int k = 0; k < mt::kN; ++k) x[(size_t)(b + 1) * mt::kN + k] = mt::next_word(x.data() + (size_t)b * mt::kN, k);
    std::vector<uint8_t> bits((size_t)n_bits);
    for (int j = 0; j < n_bits; ++j) bits[(size_t)j] = (uint8_t)(x[(size_t)j + 1] & 1u);   // from x[1]: the low bits of x[0] are not state
    Poly C;
    const int L = berlekamp_massey(bits, n_bits, C);
    if (L != kDeg) { std::fprintf(stderr, "mt_jump_gen: linear complexity %d, expected %d\n", L, kDeg); return 1; }
    // characteristic polynomial phi(t) = t^L C(1/t): coefficient of t^(L-i) is c_i
    Poly phi((size_t)kWords64 + 2, 0);
    for (int i = 0; i <= L; ++i) if (get(C, i)) flip(phi, L - i);
    // p_0 = t^624 (no reduction needed), p_{m+1} = p_m^2 mod phi
    Poly p((size_t)kWords64 + 2, 0), q;
    flip(p, mt::kN);
    FILE* out = argc > 1 ? std::fopen(argv[1], "w") : stdout;
    if (!out) { std::fprintf(stderr, "mt_jump_gen: cannot write %s\n", argv[1]); return 1; }
    std::fprintf(out, "// generated by csrc/mt_jump_gen.cpp: t^(624 * 2^m) mod phi_MT19937, m = 0 .. %d\n", n_poly - 1);
    std::fprintf(out, "#include <stdint.h>\n#include \"mt19937.cuh\"\nnamespace mt {\nconst uint32_t kJumpPoly[%d][kN] = {\n", n_poly);
    for (int m = 0; m < n_poly; ++m) {
        std::fprintf(out, "{");
        for (int w = 0; w < mt::kN; ++w) {
            const uint32_t v = (uint32_t)(p[(size_t)w >> 1] >> ((w & 1) * 32));
            std::fprintf(out, "0x%08xu,%s", v, (w % 12 == 11) ? "\n" : "");
        }
        std::fprintf(out, "},\n");
        square_mod(p, phi, q);
        p.swap(q);
    }
    std::fprintf(out, "};\n}  // namespace mt\n");
    if (out != stdout) std::fclose(out);
    return 0;
}
