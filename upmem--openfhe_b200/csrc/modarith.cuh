// modarith.cuh — 64-bit modular arithmetic on the 32-bit integer datapath of sm_100a.
//
// Device-side counterparts of the scalar primitives the reference's CPU path is built from
// (citations relative to the reference tree):
//   addmod / submod      NativeIntegerT::ModAddFastEq / ModSubFastEq   ubintnat.h:760-767, 934-938
//   mulmod_barrett       NativeIntegerT::ModMulFastEq(b, q, mu)        ubintnat.h:1398-1413
//   shoup_lazy           NativeIntegerT::ModMulFastConstEq             ubintnat.h:1491-1497
// Every function returns (or, for the *_lazy forms, is later corrected to) the canonical residue
// in [0,q), so results are bit-identical to the reference for in-range inputs.
#pragma once
#include <stdint.h>

typedef unsigned long long u64;
typedef unsigned int u32;

namespace fhe {

// Per-tower constants kept in HBM next to the data (one 32-byte record per modulus).
struct __align__(32) TowerMod {
    u64 q;        // modulus, 4 <= q < 2^62
    u64 mu;       // floor(2^(2*msb(q)+3) / q)          (ComputeMu, ubintnat.h:650-656)
    u32 n;        // msb(q) - 2                          (first Barrett shift)
    u32 s;        // n + 7                               (second Barrett shift)
    u64 pad;
};

__device__ __forceinline__ u64 addmod(u64 a, u64 b, u64 q) {
    u64 r = a + b;
    return r >= q ? r - q : r;
}

__device__ __forceinline__ u64 submod(u64 a, u64 b, u64 q) {
    u64 r = a - b;
    return a < b ? r + q : r;
}

// Generalized Barrett with alpha = n + 3, the same quotient estimate as the reference, so the
// single conditional subtraction is sufficient for a, b < q.
__device__ __forceinline__ u64 mulmod_barrett(u64 a, u64 b, const TowerMod& m) {
    u64 lo = a * b;
    u64 hi = __umul64hi(a, b);
    u64 t  = (lo >> m.n) | (hi << (64 - m.n));           // RShiftD(prod, n), one word
    u64 qt;
    if (m.s >= 64) {                                     // 58..62-bit moduli: only the high word matters
        qt = __umul64hi(t, m.mu) >> (m.s - 64);
    } else {
        u64 l2 = t * m.mu, h2 = __umul64hi(t, m.mu);
        qt = (l2 >> m.s) | (h2 << (64 - m.s));
    }
    u64 r = lo - qt * m.q;
    return r >= m.q ? r - m.q : r;
}

// x - m if that does not go negative, else x.  Needs x, m < 2^63.
__device__ __forceinline__ u64 csub(u64 x, u64 m) {
    u64 t = x - m;
    return ((long long)t < 0) ? x : t;
}

// Shoup multiplication by a constant w with companion wp = floor(w * 2^64 / q), WITHOUT the final
// correction: result in [0, 2q) for any 64-bit x (Harvey).  nq = 2^64 - q lets the subtraction
// ride on the multiply-add chain.
__device__ __forceinline__ u64 shoup_lazy(u64 x, u64 w, u64 wp, u64 nq) {
    u64 qh = __umul64hi(x, wp);
    return x * w + qh * nq;
}

// Fully reduced Shoup product (what ModMulFastConstEq returns).
__device__ __forceinline__ u64 shoup(u64 x, u64 w, u64 wp, u64 q) {
    return csub(shoup_lazy(x, w, wp, 0ull - q), q);
}

// Harvey butterflies.  q2 = 2q, nq = -q mod 2^64.
// Forward (Cooley-Tukey): X, Y in [0,4q) -> X, Y in [0,4q).
__device__ __forceinline__ void fwd_bf(u64& X, u64& Y, u64 w, u64 wp, u64 q2, u64 nq) {
    u64 x = csub(X, q2);
    u64 t = shoup_lazy(Y, w, wp, nq);
    X = x + t;
    Y = x - t + q2;
}
// Inverse (Gentleman-Sande): X, Y in [0,2q) -> X, Y in [0,2q).
__device__ __forceinline__ void inv_bf(u64& X, u64& Y, u64 w, u64 wp, u64 q2, u64 nq) {
    u64 s = X + Y;
    u64 d = X - Y + q2;
    X = csub(s, q2);
    Y = shoup_lazy(d, w, wp, nq);
}

}  // namespace fhe
