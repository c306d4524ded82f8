// hostmath.h — host-side (setup-time) number theory of the backend: parameter recipe and the
// twiddle tables the device kernels consume.  Product code (independent of oracle/).
//
// What it reproduces from the reference (citations relative to the reference tree):
//   FirstPrime / NextPrime / PreviousPrime      math/nbtheory-impl.h:334-379
//   RootOfUnity (minimum primitive m-th root)   math/nbtheory-impl.h:183-231
//   ChineseRemainderTransformFTTNat::PreCompute math/hal/intnat/transformnat-impl.h:708-763
//   NativeIntegerT::PrepModMulConst             math/hal/intnat/ubintnat.h:1447-1455
#pragma once
#include <stdint.h>
#include <vector>

namespace fhe {
namespace host {

typedef unsigned __int128 u128;

inline unsigned msb(uint64_t x) { return x ? 64u - (unsigned)__builtin_clzll(x) : 0u; }
inline uint64_t mulmod(uint64_t a, uint64_t b, uint64_t q) { return (uint64_t)(((u128)a * b) % q); }

inline uint64_t powmod(uint64_t b, uint64_t e, uint64_t q) {
    uint64_t r = 1 % q;
    b %= q;
    for (; e; e >>= 1) {
        if (e & 1) r = mulmod(r, b, q);
        b = mulmod(b, b, q);
    }
    return r;
}

inline uint64_t invmod(uint64_t a, uint64_t q) {
    __int128 t = 0, nt = 1, r = q, nr = a % q;
    while (nr != 0) {
        __int128 k = r / nr, tmp;
        tmp = t - k * nt; t = nt; nt = tmp;
        tmp = r - k * nr; r = nr; nr = tmp;
    }
    if (r != 1) return 0;                         // not invertible
    if (t < 0) t += q;
    return (uint64_t)t;
}

// floor(w * 2^64 / q): the Shoup companion of a constant multiplicand
inline uint64_t shoup_companion(uint64_t w, uint64_t q) { return (uint64_t)(((u128)w << 64) / q); }

// Deterministic Miller-Rabin, exact for all 64-bit inputs with the first twelve prime bases.
inline bool is_prime(uint64_t p) {
    static const uint64_t bases[12] = {2, 3, 5, 7, 11, 13, 17, 19, 23, 29, 31, 37};
    if (p < 2) return false;
    for (uint64_t b : bases) {
        if (p == b) return true;
        if (p % b == 0) return false;
    }
    uint64_t d = p - 1;
    unsigned s = 0;
    while (!(d & 1)) { d >>= 1; s++; }
    for (uint64_t b : bases) {
        uint64_t x = powmod(b, d, p);
        if (x == 1 || x == p - 1) continue;
        bool witness = true;
        for (unsigned r = 1; r < s && witness; r++) {
            x = mulmod(x, x, p);
            if (x == p - 1) witness = false;
        }
        if (witness) return false;
    }
    return true;
}

// smallest prime > 2^nbits that is 1 mod m
inline uint64_t first_prime(unsigned nbits, uint64_t m) {
    uint64_t q = 1ull << nbits, r = q % m, c = q + 1;
    if (r) c += m - r;
    while (!is_prime(c)) c += m;
    return c;
}
inline uint64_t next_prime(uint64_t q, uint64_t m) {
    uint64_t c = q + m;
    while (!is_prime(c)) c += m;
    return c;
}
inline uint64_t previous_prime(uint64_t q, uint64_t m) {
    uint64_t c = q - m;
    while (c > m && !is_prime(c)) c -= m;
    return c;
}

// Minimum primitive m-th root of unity mod prime q (m a power of two).  The minimum over all
// primitive roots is independent of the generator the reference draws at random.
inline uint64_t root_of_unity(uint64_t m, uint64_t q) {
    if (m < 2 || (m & (m - 1)) || (q - 1) % m) return 0;
    uint64_t e = (q - 1) / m, r = 0;
    for (uint64_t g = 2; g < q; g++) {
        r = powmod(g, e, q);
        if (powmod(r, m / 2, q) == q - 1) break;   // order exactly m
        r = 0;
    }
    if (!r) return 0;
    uint64_t r2 = mulmod(r, r, q), x = r, best = r;
    for (uint64_t k = 1; k < m; k += 2) {          // all odd powers = all primitive m-th roots
        if (x < best) best = x;
        x = mulmod(x, r2, q);
    }
    return best;
}

inline uint32_t bitrev(uint32_t x, unsigned bits) {
    uint32_t r = 0;
    for (unsigned i = 0; i < bits; i++) { r = (r << 1) | (x & 1); x >>= 1; }
    return r;
}

// Interleaved {w, companion} tables in bit-reversed order, plus the two N^-1 records the inverse
// transform's last stage needs.  tw[2*i], tw[2*i+1] = W[i], companion.
struct TowerTables {
    std::vector<uint64_t> twF, twI;               // 2*N words each
    uint64_t invc[4];                             // {N^-1, comp, WI[1]*N^-1 mod q, comp}
};

inline void precompute(uint64_t psi, uint32_t N, uint64_t q, TowerTables& t) {
    unsigned logN = msb(N) - 1;
    t.twF.assign(2 * (size_t)N, 0);
    t.twI.assign(2 * (size_t)N, 0);
    uint64_t psiinv = invmod(psi, q), x = 1 % q, xi = 1 % q;
    for (uint32_t i = 0; i < N; i++) {
        uint32_t r = bitrev(i, logN);
        t.twF[2 * (size_t)r] = x;  t.twF[2 * (size_t)r + 1] = shoup_companion(x, q);
        t.twI[2 * (size_t)r] = xi; t.twI[2 * (size_t)r + 1] = shoup_companion(xi, q);
        x = mulmod(x, psi, q);
        xi = mulmod(xi, psiinv, q);
    }
    uint64_t ninv = invmod((uint64_t)N % q, q);
    uint64_t w1 = N > 1 ? mulmod(t.twI[2], ninv, q) : ninv;
    t.invc[0] = ninv; t.invc[1] = shoup_companion(ninv, q);
    t.invc[2] = w1;   t.invc[3] = shoup_companion(w1, q);
}

}  // namespace host
}  // namespace fhe
