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
    u32 rsh;      // range reduction: shift that brings 16q below 2^24 ...
    float rinv;   // ... and a rounded-DOWN reciprocal of (q >> rsh) (+1 if rsh > 0)
};

// Host-side fill of the range-reduction constants (used by runtime.cu / elementwise.cu).
inline void towermod_set_reduce(TowerMod& m) {
    unsigned bits = m.q ? 64u - (unsigned)__builtin_clzll(m.q) : 0u;
    m.rsh = bits + 4 > 24 ? bits + 4 - 24 : 0;
    double d = (double)((m.q >> m.rsh) + (m.rsh ? 1 : 0));
    float r = (float)(1.0 / d);
    while ((double)r * d > 1.0) r = __builtin_nextafterf(r, 0.0f);
    m.rinv = r;
}

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
    // low 64 bits of x*w + qh*nq with every 32x32 partial product accumulated by the multiplier itself
    // (2 IMAD.WIDE + 4 IMAD chained through the high word, no separate adds)
    u32 x0 = (u32)x, x1 = (u32)(x >> 32), w0 = (u32)w, w1 = (u32)(w >> 32);
    u32 h0 = (u32)qh, h1 = (u32)(qh >> 32), n0 = (u32)nq, n1 = (u32)(nq >> 32);
    u64 acc;
    u32 lo, hi;
    asm("mul.wide.u32 %0, %1, %2;" : "=l"(acc) : "r"(x0), "r"(w0));
    asm("mad.wide.u32 %0, %1, %2, %0;" : "+l"(acc) : "r"(h0), "r"(n0));
    asm("mov.b64 {%0, %1}, %2;" : "=r"(lo), "=r"(hi) : "l"(acc));
    asm("mad.lo.u32 %0, %1, %2, %0;" : "+r"(hi) : "r"(x0), "r"(w1));
    asm("mad.lo.u32 %0, %1, %2, %0;" : "+r"(hi) : "r"(x1), "r"(w0));
    asm("mad.lo.u32 %0, %1, %2, %0;" : "+r"(hi) : "r"(h0), "r"(n1));
    asm("mad.lo.u32 %0, %1, %2, %0;" : "+r"(hi) : "r"(h1), "r"(n0));
    asm("mov.b64 %0, {%1, %2};" : "=l"(acc) : "r"(lo), "r"(hi));
    return acc;
}

// The same with the quotient estimated from three of the four 32x32 partial products of x * wp (the low x low one
// is dropped, the two middle ones contribute only their high words): the estimate is at most 2 short, so the
// result is in [0, 4q) for any 64-bit x.  5 half-rate multiplies + 4 full-rate ones instead of 6 + 4.
__device__ __forceinline__ u64 shoup_lazy_approx(u64 x, u64 w, u64 wp, u64 nq) {
    u32 x0 = (u32)x, x1 = (u32)(x >> 32), w0 = (u32)w, w1 = (u32)(w >> 32);
    u32 p0 = (u32)wp, p1 = (u32)(wp >> 32), n0 = (u32)nq, n1 = (u32)(nq >> 32);
    u32 t1, t2;
    asm("mul.hi.u32 %0, %1, %2;" : "=r"(t1) : "r"(x1), "r"(p0));
    asm("mul.hi.u32 %0, %1, %2;" : "=r"(t2) : "r"(x0), "r"(p1));
    u64 qh = (u64)t1 + (u64)t2;
    asm("mad.wide.u32 %0, %1, %2, %0;" : "+l"(qh) : "r"(x1), "r"(p1));
    u32 h0 = (u32)qh, h1 = (u32)(qh >> 32);
    u64 acc;
    u32 lo, hi;
    asm("mul.wide.u32 %0, %1, %2;" : "=l"(acc) : "r"(x0), "r"(w0));
    asm("mad.wide.u32 %0, %1, %2, %0;" : "+l"(acc) : "r"(h0), "r"(n0));
    asm("mov.b64 {%0, %1}, %2;" : "=r"(lo), "=r"(hi) : "l"(acc));
    asm("mad.lo.u32 %0, %1, %2, %0;" : "+r"(hi) : "r"(x0), "r"(w1));
    asm("mad.lo.u32 %0, %1, %2, %0;" : "+r"(hi) : "r"(x1), "r"(w0));
    asm("mad.lo.u32 %0, %1, %2, %0;" : "+r"(hi) : "r"(h0), "r"(n1));
    asm("mad.lo.u32 %0, %1, %2, %0;" : "+r"(hi) : "r"(h1), "r"(n0));
    asm("mov.b64 %0, {%1, %2};" : "=l"(acc) : "r"(lo), "r"(hi));
    return acc;
}

// Butterfly multiplier of the NTT kernels: SHOUP_G = bound (in units of q) of its result.
#ifndef FHE_NTT_EXACT_MULHI
constexpr int SHOUP_G = 4;
__device__ __forceinline__ u64 shoup_bf(u64 x, u64 w, u64 wp, u64 nq) { return shoup_lazy_approx(x, w, wp, nq); }
#else
constexpr int SHOUP_G = 2;
__device__ __forceinline__ u64 shoup_bf(u64 x, u64 w, u64 wp, u64 nq) { return shoup_lazy(x, w, wp, nq); }
#endif

// Fully reduced Shoup product (what ModMulFastConstEq returns).
__device__ __forceinline__ u64 shoup(u64 x, u64 w, u64 wp, u64 q) {
    return csub(shoup_lazy(x, w, wp, 0ull - q), q);
}

// ---- lazy butterflies with compile-time range tracking ---------------------------------------
// Every modulus on the NTT path is below 2^60 (the reference's MAX_MODULUS_SIZE, basicint.h:44-45), so a
// 64-bit word holds values up to 16q.  Values are kept in [0, B*q) with B known at compile time and
// are only pulled back (conditional subtractions of 8q, 4q, 2q, q) where B would pass 16; the last
// store of a transform reduces to [0,q), so results equal the reference's fully reduced ones.
struct QMul { u64 q1, q2, q4, q8; };            // q, 2q, 4q, 8q
__device__ __forceinline__ QMul make_qmul(u64 q) { return QMul{q, 2 * q, 4 * q, 8 * q}; }

template <int M>
__device__ __forceinline__ u64 qm(const QMul& Q) {
    static_assert(M == 1 || M == 2 || M == 4 || M == 8, "multiple of q not kept in registers");
    return M == 1 ? Q.q1 : M == 2 ? Q.q2 : M == 4 ? Q.q4 : Q.q8;
}

// [0, FROM*q) -> [0, TO*q), FROM <= 16, by conditional subtractions of the power-of-two multiples.
template <int FROM, int TO>
__device__ __forceinline__ u64 reduce_range(u64 x, const QMul& Q) {
    if constexpr (FROM <= TO) {
        return x;
    } else {
        constexpr int M = FROM > 8 ? 8 : FROM > 4 ? 4 : FROM > 2 ? 2 : 1;   // largest kept multiple below FROM
        constexpr int NEXT = (FROM - M) > M ? (FROM - M) : M;
        return reduce_range<NEXT, TO>(csub(x, qm<M>(Q)), Q);
    }
}

// v in [0,16q) -> [0,q) with one quotient estimate instead of four conditional subtractions:
// k = floor(float(v >> rsh) * rinv) never exceeds floor(v/q) and is at most one short of it (the
// operands are below 2^24, hence exact in fp32; rinv and the product are rounded towards zero), so
// v - k*q is in [0,2q).  kq = shared table {0, q, 2q, .., 15q}.  Conversions run on the XU pipe, the
// table read on the LSU: neither touches the integer-multiply pipe the butterflies are bound by.
__device__ __forceinline__ u64 reduce16(u64 v, const u64* kq, u32 rsh, float rinv, u64 q) {
    u32 k = __float2uint_rz(__fmul_rz(__uint2float_rz((u32)(v >> rsh)), rinv));
    return csub(v - kq[k], q);
}

// Forward (Cooley-Tukey): X in [0,Bq), Y anything below 2^64 -> both in [0,(B+SHOUP_G)q).  No reduction of X.
// qg = SHOUP_G * q.
__device__ __forceinline__ void fwd_bf(u64& X, u64& Y, u64 w, u64 wp, u64 qg, u64 nq) {
    u64 t = shoup_bf(Y, w, wp, nq);              // [0, SHOUP_G*q)
    Y = X - t + qg;
    X = X + t;
}
// Inverse (Gentleman-Sande): X in [0,BXq), Y in [0,BYq) -> X in [0,(BX+BY)q), Y in [0,SHOUP_G*q).
// byq = BY*q keeps the difference non-negative.
__device__ __forceinline__ void inv_bf(u64& X, u64& Y, u64 w, u64 wp, u64 byq, u64 nq) {
    u64 d = X - Y + byq;
    X = X + Y;
    Y = shoup_bf(d, w, wp, nq);
}

}  // namespace fhe
