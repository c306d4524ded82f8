// ntt_steps.cuh — register-resident radix-2^S butterfly steps shared by the NTT kernels (ntt.cu, ntt_ws.cu).
//
// Same index pattern as the reference's CPU kernels (citations relative to the reference tree):
//   NumberTheoreticTransformNat::ForwardTransformToBitReverseInPlace   transformnat-impl.h:300-354
//   NumberTheoreticTransformNat::InverseTransformFromBitReverseInPlace transformnat-impl.h:492-552
// Lazy reduction with compile-time range tracking: values stay below 16q (q < 2^60) and the butterfly multiplier
// returns [0,4q) (modarith.cuh); the last store of a transform reduces to [0,q), so outputs equal the reference's.
#pragma once
#include "modarith.cuh"

#include <type_traits>

namespace fhe {

__device__ __forceinline__ ulonglong2 ldtw(const ulonglong2* __restrict__ p) { return __ldg(p); }

template <int I, int N, class F>
__device__ __forceinline__ void static_for(F&& f) {
    if constexpr (I < N) {
        f(std::integral_constant<int, I>{});
        static_for<I + 1, N>(f);
    }
}

constexpr int NTT_TILE_LOG_DEFAULT[2] = {11, 10};   // {strided, contig}: measured best on B200 (profiles/)
constexpr int NTT_FUSED_DEFAULT = 1;                    // single-pass kernel for N <= 2^14 where it wins (ntt.cu): FHE_B200_NTT_FUSED overrides
constexpr int NTT_PF_DEFAULT = 0;                       // persistent prefetching strided pass (ntt.cu): FHE_B200_NTT_PF overrides
constexpr int MAXB = 16;
constexpr int INV_NORM = 4;  // uniform bound the inverse restores between register steps and between the two passes     // values stay below MAXB*q < 2^64 (q < 2^60)

// ---- forward: E = bound (in units of q) of every coefficient entering a step of S stages.
// A stage adds SHOUP_G to the bound of both outputs whatever the bound of the multiplied input, so only the
// elements that are NOT multiplied in a stage have to be below 16 - SHOUP_G when it starts; where the running
// bound would pass that, those elements take one conditional subtraction of 8q first.  Simulated at compile time.
template <int S, int E>
struct FwdStep {
    static constexpr int before(int j) {                 // bound entering stage j, before its optional reduction
        int b = E;
        for (int k = 0; k < j; k++) {
            if (b + SHOUP_G > MAXB) b = 8;
            b += SHOUP_G;
        }
        return b;
    }
    static constexpr bool reduce(int j) { return before(j) + SHOUP_G > MAXB; }
    static constexpr int OUT = before(S);
    static_assert(E <= MAXB && OUT <= MAXB, "forward step overflows the 16q headroom");
};

// One butterfly stage J (of S) on 2^S register-resident coefficients.  The coefficients are the
// sub-NTT whose twiddle prefix is P: stage J uses table[(P << J) + (e >> (S-J))].
template <int S, int J, bool REDUCE>
__device__ __forceinline__ void fwd_stage(u64 (&x)[1 << S], const ulonglong2* __restrict__ tw, u32 P,
                                          u64 qg, u64 q8, u64 nq) {
    constexpr int half = 1 << (S - 1 - J);
#pragma unroll
    for (int g = 0; g < (1 << J); g++) {
        ulonglong2 w = ldtw(tw + ((P << J) + g));
#pragma unroll
        for (int i = 0; i < half; i++) {
            if constexpr (REDUCE) x[g * 2 * half + i] = csub(x[g * 2 * half + i], q8);
            fwd_bf(x[g * 2 * half + i], x[g * 2 * half + i + half], w.x, w.y, qg, nq);
        }
    }
}

template <int S, int E>
__device__ __forceinline__ void fwd_step(u64 (&x)[1 << S], const ulonglong2* __restrict__ tw, u32 P,
                                         const QMul& Q, u64 nq) {
    if constexpr (S > 0) {
        static_for<0, S>([&](auto j) {
            constexpr int J = decltype(j)::value;
            fwd_stage<S, J, FwdStep<S, E>::reduce(J)>(x, tw, P, qm<SHOUP_G>(Q), Q.q8, nq);
        });
    }
}

// ---- inverse: the sum output doubles the bound, the multiplied output comes back in [0,2q), so the
// bounds differ per register.  The schedule of conditional subtractions is simulated at compile time.
template <int S>
struct InvPlan {
    int cur[S + 1][1 << S];   // bound of register e before stage k (k-th executed: J = S-1-k), before `tgt`
    int tgt[S + 1][1 << S];   // bound register e is reduced to before stage k (== cur: nothing to do)
    int out;                  // uniform bound after the closing normalisation (row S of cur -> tgt)
};

template <int S>
constexpr InvPlan<S> make_inv_plan(int E, int NORM) {
    InvPlan<S> p{};
    int b[1 << S] = {};
    for (int e = 0; e < (1 << S); e++) b[e] = E;
    for (int k = 0; k < S; k++) {
        int half = 1 << k;                                  // executed order: J = S-1 first (half 1)
        for (int e = 0; e < (1 << S); e++) { p.cur[k][e] = b[e]; p.tgt[k][e] = b[e]; }
        for (int e = 0; e < (1 << S); e++) {
            if (e & half) continue;
            int bx = b[e], by = b[e + half];
            while (bx + by > MAXB || by > 8) {              // by*q must be one of q,2q,4q,8q
                if (bx >= by) bx = bx > 8 ? 8 : bx / 2; else by = by > 8 ? 8 : by / 2;
            }
            p.tgt[k][e] = bx; p.tgt[k][e + half] = by;
            b[e] = bx + by; b[e + half] = SHOUP_G;
        }
    }
    for (int e = 0; e < (1 << S); e++) {
        p.cur[S][e] = b[e];
        p.tgt[S][e] = b[e] > NORM ? NORM : b[e];
    }
    p.out = NORM;
    return p;
}

template <int S, int E, int NORM>
struct InvStep {
    static constexpr InvPlan<S> P = make_inv_plan<S>(E, NORM);
};

// executed stage K (J = S-1-K) of an inverse step
template <int S, int E, int NORM, int K>
__device__ __forceinline__ void inv_stage(u64 (&x)[1 << S], const ulonglong2* __restrict__ tw, u32 P,
                                          const QMul& Q, u64 nq) {
    constexpr int J = S - 1 - K, half = 1 << K;
    using PL = InvStep<S, E, NORM>;
    static_for<0, (1 << J)>([&](auto gi) {
        constexpr int g = decltype(gi)::value;
        ulonglong2 w = ldtw(tw + ((P << J) + g));
        static_for<0, half>([&](auto ii) {
            constexpr int a = g * 2 * half + decltype(ii)::value, b = a + half;
            constexpr int BY = PL::P.tgt[K][b];
            x[a] = reduce_range<PL::P.cur[K][a], PL::P.tgt[K][a]>(x[a], Q);
            x[b] = reduce_range<PL::P.cur[K][b], BY>(x[b], Q);
            inv_bf(x[a], x[b], w.x, w.y, qm<(BY < 1 ? 1 : BY)>(Q), nq);
        });
    });
}

// all S stages, then normalisation of every register to [0, NORM*q)
template <int S, int E, int NORM>
__device__ __forceinline__ void inv_step(u64 (&x)[1 << S], const ulonglong2* __restrict__ tw, u32 P,
                                         const QMul& Q, u64 nq) {
    using PL = InvStep<S, E, NORM>;
    static_for<0, S>([&](auto k) { inv_stage<S, E, NORM, decltype(k)::value>(x, tw, P, Q, nq); });
    static_for<0, (1 << S)>([&](auto ei) {
        constexpr int e = decltype(ei)::value;
        x[e] = reduce_range<PL::P.cur[S][e], PL::P.tgt[S][e]>(x[e], Q);
    });
}

// Inverse step that contains the very last stage of the whole transform (global stage m = 1):
// folds N^-1 into it like transformnat-impl.h:497-524 folds it into the first one, and reduces
// both outputs to [0,q).  sc = {N^-1, companion}, w1 = {WI[1]*N^-1 mod q, companion}.
template <int S, int E>
__device__ __forceinline__ void inv_step_last(u64 (&x)[1 << S], const ulonglong2* __restrict__ tw,
                                              ulonglong2 sc, ulonglong2 w1, const QMul& Q, u64 nq) {
    using PL = InvStep<S, E, MAXB>;
    static_for<0, S - 1>([&](auto k) { inv_stage<S, E, MAXB, decltype(k)::value>(x, tw, 1u, Q, nq); });
    constexpr int K = S - 1, half = 1 << K;
    static_for<0, half>([&](auto ii) {
        constexpr int a = decltype(ii)::value, b = a + half;
        constexpr int BY = PL::P.tgt[K][b];
        x[a] = reduce_range<PL::P.cur[K][a], PL::P.tgt[K][a]>(x[a], Q);
        x[b] = reduce_range<PL::P.cur[K][b], BY>(x[b], Q);
        u64 s = x[a] + x[b];
        u64 d = x[a] - x[b] + qm<(BY < 1 ? 1 : BY)>(Q);
        x[a] = csub(shoup_lazy(s, sc.x, sc.y, nq), Q.q1);
        x[b] = csub(shoup_lazy(d, w1.x, w1.y, nq), Q.q1);
    });
}

}  // namespace fhe
