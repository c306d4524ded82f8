// basis.cu — RNS basis conversion and scaling kernels for sm_100a.
//
// Device replacement for (citations relative to the reference tree, lattice/hal/default/dcrtpoly-impl.h):
//   DCRTPolyImpl::ApproxSwitchCRTBasis   :1034-1082  (HAVE_INT128 branch; Mul128 / BarrettUint128ModUint64,
//                                                     utils/utilities-int.h:47-110)
//   the element-wise tails of ApproxModDown :1166-1174, DropLastElementAndScale :753-767, ModReduce :803-811
//   NativeVectorT::SwitchModulus           math/hal/intnat/mubintvecnat.cpp:112-135
//
// basis_switch_kernel: one thread per coefficient index.  For each source tower i it forms
// y_i = x_i * [(Q/q_i)^-1]_{q_i} mod q_i (Shoup, canonical) and accumulates y_i * [Q/q_i]_{p_j} for up to ten target
// towers at a time - carry-free on 30-bit halves, three 32x32 multiplies per product (Acc30 / mac_group below; the
// 128-bit multiply-accumulate form mac128 serves tables with entries of 2^60 and more) - then reduces each sum modulo p_j
// (reduce128).  Every value stored is the canonical residue the CPU path produces.  The [Q/q_i]_{p_j} slice, the Shoup
// constants and the moduli sit in shared memory.  Integer-pipe bound: 3 IMAD.WIDE per (i, j) pair per coefficient; HBM
// traffic is 8 B per source tower and 8 B per target tower per coefficient.
// tower_affine_kernel / rescale_spread_kernel: one pass, 128-bit loads and stores, HBM-bound.
#include <stdlib.h>
#include <algorithm>
#include <atomic>
#include "modarith.cuh"
#include "kernels.h"

namespace fhe {

// acc += a * b on a 128-bit accumulator (the general form, any 64-bit factors).  Written with unsigned __int128 so that
// ptxas emits the four 32x32 partial products once (4 IMAD.WIDE.U32 with carry chains); the mad.lo.cc / madc.hi PTX pair
// recomputes the low product (about 6 wide multiplies per MAC, measured with tools/sass_mix.py).
__device__ __forceinline__ void mac128(u64& lo, u64& hi, u64 a, u64 b) {
    unsigned __int128 acc = ((unsigned __int128)hi << 64) | lo;
    acc += (unsigned __int128)a * b;
    lo = (u64)acc;
    hi = (u64)(acc >> 64);
}

// (hi * 2^64 + lo) mod p, canonical - the value BarrettUint128ModUint64 (utils/utilities-int.h:60-110) returns, by a
// cheaper route than its three 64x64 products against floor(2^128/p): with c = 2^64 mod p the sum is congruent to
// hi * c + lo, and both parts take a one-word quotient estimate (Shoup's for hi * c with c' = floor(c * 2^64 / p); the
// same with the constant 1 for lo, floor(2^64 / p)), each short by at most one, so
// hi * c + lo - (mulhi(hi, c') + mulhi(lo, pinv)) * p is in [0, 4p) and only its low 64 bits are ever formed.
// 2 high products + 2 low ones instead of ~5 full ones; no loop.  Needs p < 2^62.
struct TargetMod { u64 p, c64, c64p, pinv, post, postp; };
__device__ __forceinline__ u64 reduce128(u64 lo, u64 hi, const TargetMod& t) {
    u64 qh = __umul64hi(hi, t.c64p) + __umul64hi(lo, t.pinv);
    u64 r = lo + hi * t.c64 - qh * t.p;
    u64 p2 = t.p << 1;
    if (r >= p2) r -= p2;
    if (r >= t.p) r -= t.p;
    return r;
}

// NativeVectorT::SwitchModulus on one value (mubintvecnat.cpp:112-135); up = nm > om, diff = nm - om or nm - om % nm
__device__ __forceinline__ u64 switch_modulus(u64 v, u64 half, u64 nm, u64 diff, bool up) {
    if (v > half) v += diff;
    if (!up && v >= nm) {
        v -= nm;
        if (v >= nm) v %= nm;
    }
    return v;
}

#ifndef FHE_BS_THREADS
#define FHE_BS_THREADS 128
#endif
constexpr int BS_THREADS = FHE_BS_THREADS;
constexpr int BS_JT_MAX = 10;                  // target towers accumulated per thread (8 registers each in the split form)

// Carry-free form of the same sum for operands below 2^60 (every modulus on this path is, basicint.h:44-45; the host
// checks the tables and picks the mac128 form otherwise).  Both factors are cut at bit 30: y = y0 + y1 * 2^30,
// h = h0 + h1 * 2^30, and the product is taken Karatsuba-wise from THREE 32x32 multiplies, y0*h0, y1*h1 and
// (y0+y1)*(h0+h1) (the sums are below 2^31, the product below 2^62).  Up to four source towers are summed through the
// multiplier's own 64-bit addend (IMAD.WIDE chains that start from zero: 4 * 2^62 fits), then the middle term is
// recovered with two subtractions and the three partial sums join the running limbs a0 (weight 1), a1 (2^30), a2 (2^60)
// with plain 64-bit adds - no 128-bit carry chain anywhere.  The limbs hold 8 terms between normalisations
// (a1 takes less than 2^61 per term: 8 * 2 * (2^30-1)^2 + 2^34 + 2^30 < 2^64); normalise() pushes the bits above 30 of
// each limb into the next one, a3 collects everything from bit 90 up (the whole sum is below 2^128).
constexpr u32 M30 = (1u << 30) - 1;
#ifndef FHE_MINB
#define FHE_MINB 4
#endif
struct Acc30 {
    u64 a0, a1, a2, a3;
    __device__ __forceinline__ void clear() { a0 = a1 = a2 = a3 = 0; }
    __device__ __forceinline__ void normalise() {
        a1 += a0 >> 30; a0 &= M30;
        a2 += a1 >> 30; a1 &= M30;
        a3 += a2 >> 30; a2 &= M30;
    }
    __device__ __forceinline__ void value(u64& lo, u64& hi) {      // after normalise()
        lo = a0 | (a1 << 30) | (a2 << 60);
        hi = (a2 >> 4) | (a3 << 26);
    }
};
// value below 2^60 -> its low and high 30 bits, in 32-bit operations only (a 64-bit shift followed by a truncation makes
// ptxas carry a dead high word through every product that uses the result)
__device__ __forceinline__ void split30(u64 y, u32& y0, u32& y1) {
    u32 l, h;
    asm("mov.b64 {%0, %1}, %2;" : "=r"(l), "=r"(h) : "l"(y));
    y0 = l & M30;
    y1 = __funnelshift_r(l, h, 30);
}
// table entry in shared memory, split form: {h0, h1, h0 + h1, -}
__device__ __forceinline__ uint4 pack30(u64 h) {
    u32 h0 = (u32)h & M30, h1 = (u32)(h >> 30);
    return make_uint4(h0, h1, h0 + h1, 0u);
}
// acc[jj] += sum_g y[g] * h[g][jj] for G <= 4 source towers and JT target towers; hr = [G][JT] entries
template <int G, int JT>
__device__ __forceinline__ void mac_group(Acc30 (&acc)[JT], const u64 (&y)[G], const uint4* __restrict__ hr) {
    static_assert(G >= 1 && G <= 4, "the middle products of more than four terms overflow 64 bits");
    u32 y0[G], y1[G], ys[G];
#pragma unroll
    for (int g = 0; g < G; g++) {
        split30(y[g], y0[g], y1[g]);
        ys[g] = y0[g] + y1[g];
    }
#pragma unroll
    for (int jj = 0; jj < JT; jj++) {
        u64 t0 = 0, t2 = 0, tm = 0;
#pragma unroll
        for (int g = 0; g < G; g++) {
            const uint4 h = hr[g * JT + jj];
            t0 += (u64)y0[g] * h.x;
            t2 += (u64)y1[g] * h.y;
            tm += (u64)ys[g] * h.z;
        }
        acc[jj].a0 += t0;
        acc[jj].a1 += tm - t0 - t2;
        acc[jj].a2 += t2;
    }
}
// All sq source towers: groups of four, then two, then one.  load(i) fetches the coefficient of tower i, mul(i, x) turns
// it into the multiplier (the Shoup product for the basis switch, the coefficient itself for ScaleAndRound).  The loads
// of the next group are issued before the multiplies of the current one.
template <int JT, class L, class F>
__device__ __forceinline__ void mac_towers(Acc30 (&acc)[JT], u32 sq, const uint4* __restrict__ sH, L&& load, F&& mul) {
    u64 xv[4];
#pragma unroll
    for (int g = 0; g < 4; g++) xv[g] = (u32)g < sq ? load(g) : 0ull;
    u32 i = 0;
    for (; i + 4 <= sq; i += 4) {
        u64 y[4];
#pragma unroll
        for (int g = 0; g < 4; g++) y[g] = mul(i + g, xv[g]);
#pragma unroll
        for (int g = 0; g < 4; g++) xv[g] = i + 4 + g < sq ? load(i + 4 + g) : 0ull;
        mac_group<4, JT>(acc, y, sH + (size_t)i * JT);
        if (i & 4) {
#pragma unroll
            for (int jj = 0; jj < JT; jj++) acc[jj].normalise();
        }
    }
    if (i + 2 <= sq) {
        const u64 y[2] = {mul(i, xv[0]), mul(i + 1, xv[1])};
        mac_group<2, JT>(acc, y, sH + (size_t)i * JT);
        i += 2;
        xv[0] = xv[2];
    }
    if (i < sq) {
        const u64 y[1] = {mul(i, xv[0])};
        mac_group<1, JT>(acc, y, sH + (size_t)i * JT);
    }
#pragma unroll
    for (int jj = 0; jj < JT; jj++) acc[jj].normalise();
}

template <int JT, bool EXACT, bool SPLIT>
__global__ void __launch_bounds__(BS_THREADS, SPLIT ? FHE_MINB : 2 * FHE_MINB)
basis_switch_kernel(u64* __restrict__ dst, const u64* __restrict__ src, size_t items, u32 logN, BasisDev bd,
                    u32 srcT, u32 src0, u32 dstT, u32 dst0) {
    extern __shared__ __align__(16) u64 sm[];
    constexpr int HW = SPLIT ? 2 : 1;                      // words per table entry
    u64* sH = sm;                                          // [sq][JT] (split form: uint4 {h0, h1, h0+h1, -})
    u64* sQ = sH + (size_t)bd.sq * JT * HW;                // [sq]
    u64* sC = sQ + bd.sq;                                  // [sq][2] {c, companion}
    double* sI = reinterpret_cast<double*>(sC + 2 * bd.sq); // [sq] 1/q_i (EXACT only)
    TargetMod* sT = reinterpret_cast<TargetMod*>(sI + bd.sq); // [JT]
    const u32 j0 = blockIdx.y * JT;
    if (threadIdx.x < JT && j0 + threadIdx.x < bd.sp) {
        const BasisTarget g = bd.tgt[j0 + threadIdx.x];
        sT[threadIdx.x] = TargetMod{g.p, g.c64, g.c64p, g.pinv, g.post, g.postp};
    }
    for (u32 k = threadIdx.x; k < bd.sq * JT; k += BS_THREADS) {
        u32 i = k / JT, jj = k % JT;
        u64 h = (j0 + jj < bd.sp) ? bd.hat[(size_t)i * bd.sp + j0 + jj] : 0ull;
        if (SPLIT) reinterpret_cast<uint4*>(sH)[k] = pack30(h);
        else sH[k] = h;
    }
    for (u32 i = threadIdx.x; i < bd.sq; i += BS_THREADS) {
        sQ[i] = bd.q[i];
        sC[2 * i] = bd.pre[i].x;
        sC[2 * i + 1] = bd.pre[i].y;
        if (EXACT) sI[i] = bd.qinv[i];
    }
    __syncthreads();
    // a CTA walks over coefficient tiles: the tables above are staged once for all of them
    for (size_t e = (size_t)blockIdx.x * BS_THREADS + threadIdx.x; e < items; e += (size_t)gridDim.x * BS_THREADS) {
        const size_t b = e >> logN;
        const u32 n = (u32)e & ((1u << logN) - 1);
        const u64* sp = src + (((b * srcT) + src0) << logN) + n;
        u64 lo[JT], hi[JT];
        double nu = 0.5;
        // multiplier of source tower i: y_i = x_i * [(Q/q_i)^-1]_{q_i} mod q_i, canonical (:1056-1058)
        auto load = [&](u32 i) -> u64 { return sp[(size_t)i << logN]; };
        auto term = [&](u32 i, u64 x) -> u64 {
            u64 y = shoup(x, sC[2 * i], sC[2 * i + 1], sQ[i]);
            if (EXACT) nu = __dadd_rn(nu, __dmul_rn(__ull2double_rn(y), sI[i]));
            return y;
        };
        if constexpr (SPLIT) {
            Acc30 acc[JT];
#pragma unroll
            for (int jj = 0; jj < JT; jj++) acc[jj].clear();
            mac_towers<JT>(acc, bd.sq, reinterpret_cast<const uint4*>(sH), load, term);
#pragma unroll
            for (int jj = 0; jj < JT; jj++) acc[jj].value(lo[jj], hi[jj]);
        } else {
#pragma unroll
            for (int jj = 0; jj < JT; jj++) { lo[jj] = 0; hi[jj] = 0; }
            u64 x = sp[0];
            for (u32 i = 0; i < bd.sq; i++) {
                u64 xn = (i + 1 < bd.sq) ? sp[(size_t)(i + 1) << logN] : 0ull;     // prefetch the next tower's coefficient
                u64 y = shoup(x, sC[2 * i], sC[2 * i + 1], sQ[i]);
                if (EXACT) nu = __dadd_rn(nu, __dmul_rn(__ull2double_rn(y), sI[i]));
#pragma unroll
                for (int jj = 0; jj < JT; jj++) mac128(lo[jj], hi[jj], y, sH[i * JT + jj]);
                x = xn;
            }
        }
        const u32 alpha = EXACT ? (u32)nu : 0u;
#pragma unroll
        for (int jj = 0; jj < JT; jj++) {
            if (j0 + jj < bd.sp) {
                const TargetMod tg = sT[jj];
                u64 r = reduce128(lo[jj], hi[jj], tg);
                if (tg.post) r = shoup(r, tg.post, tg.postp, tg.p);
                if (EXACT) r = submod(r, bd.alphaq[(size_t)alpha * bd.sp + j0 + jj], tg.p);
                dst[(((b * dstT) + dst0 + j0 + jj) << logN) + n] = r;
            }
        }
    }
}

// Grid: a few waves of CTAs (SMs x resident CTAs of this kernel), shared between the column blocks; every CTA stages
// the tables once and walks over its coefficient tiles.  Eight waves measured best on the B200 (one wave leaves the SMs
// in lock step, one tile per CTA pays the table staging per tile); FHE_B200_BASIS_WAVES overrides (tuning only).
static int env_int(const char* name, int dflt) {
    const char* e = getenv(name);
    return e && *e ? atoi(e) : dflt;
}
// FHE_B200_TOWER_SPLIT=0: one thread walks all towers of its coefficients even when the input is small (for A/B runs)
static bool tower_split_on() {
    static const int on = env_int("FHE_B200_TOWER_SPLIT", 1);
    return on != 0;
}
template <class K>
static unsigned persistent_grid_x(K kernel, size_t smem, size_t tiles, int nblk) {
    static const int waves = env_int("FHE_B200_BASIS_WAVES", 8);
    int dev = 0, sms = 148, occ = 1;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kernel, BS_THREADS, smem) != cudaSuccess || occ < 1) occ = 1;
    size_t cap = (size_t)sms * occ * (waves > 0 ? waves : 1) / nblk;
    if (cap < 1) cap = 1;
    return (unsigned)(tiles < cap ? tiles : cap);
}
// Column blocking: as few blocks of target towers as the accumulator registers allow, all of (nearly) the same width, so
// that no thread accumulates padding columns (sp = 10 runs as one block of 10, sp = 30 as three of 10, sp = 17 as 9 + 8).
static void column_blocks(uint32_t sp, bool narrow, int& jt, int& nblk) {
    static const int jenv = env_int("FHE_B200_BASIS_JT", 0);
    int jmax = narrow ? BS_JT_MAX : 8;
    if (jenv >= 1 && jenv < jmax) jmax = jenv;
    nblk = (int)(sp + jmax - 1) / jmax;
    jt = (int)(sp + nblk - 1) / nblk;
}

template <int JT, bool EXACT>
static void launch_basis_switch_jt(int jt, bool split, int nblk, size_t smem, cudaStream_t st, u64* d, const u64* s, size_t items,
                                   u32 logN, const BasisDev& bd, u32 srcT, u32 src0, u32 dstT, u32 dst0) {
    if constexpr (JT >= 1) {
        if (jt == JT) {
            const size_t tiles = (items + BS_THREADS - 1) / BS_THREADS;
            if (split) {
                auto k = basis_switch_kernel<JT, EXACT, true>;
                k<<<dim3(persistent_grid_x(k, smem, tiles, nblk), nblk), BS_THREADS, smem, st>>>(d, s, items, logN, bd, srcT, src0, dstT, dst0);
            } else {
                auto k = basis_switch_kernel<(JT > 8 ? 8 : JT), EXACT, false>;
                k<<<dim3(persistent_grid_x(k, smem, tiles, nblk), nblk), BS_THREADS, smem, st>>>(d, s, items, logN, bd, srcT, src0, dstT, dst0);
            }
        } else {
            launch_basis_switch_jt<JT - 1, EXACT>(jt, split, nblk, smem, st, d, s, items, logN, bd, srcT, src0, dstT, dst0);
        }
    }
}

template <bool EXACT>
static cudaError_t launch_basis_switch_t(const BasisDev& bd, u64* d, uint32_t dstT, uint32_t dst0, const u64* s, uint32_t srcT,
                                         uint32_t src0, size_t items, uint32_t logN, cudaStream_t st) {
    int jt, nblk;
    column_blocks(bd.sp, bd.narrow != 0, jt, nblk);
    nblk = (int)(bd.sp + jt - 1) / jt;
    size_t smem = ((size_t)bd.sq * jt * (bd.narrow ? 2 : 1) + 4 * (size_t)bd.sq) * 8 + (size_t)jt * sizeof(TargetMod);
    launch_basis_switch_jt<BS_JT_MAX, EXACT>(jt, bd.narrow != 0, nblk, smem, st, d, s, items, logN, bd, srcT, src0, dstT, dst0);
    kernel_launch_count_add(1);
    return cudaGetLastError();
}

bool basis_switch_tc_applies(const BasisDev& bd, uint32_t logN);
static cudaError_t launch_basis_switch_tc(const BasisDev& bd, u64* d, uint32_t dstT, uint32_t dst0, const u64* s, uint32_t srcT,
                                          uint32_t src0, size_t items, uint32_t logN, bool exact, cudaStream_t st);

cudaError_t launch_basis_switch(const BasisDev& bd, uint64_t* dst, uint32_t dstT, uint32_t dst0, const uint64_t* src,
                                uint32_t srcT, uint32_t src0, size_t batch, uint32_t logN, bool exact, cudaStream_t st) {
    size_t items = batch << logN;
    if (items == 0) return cudaSuccess;
    u64* d = reinterpret_cast<u64*>(dst);
    const u64* s = reinterpret_cast<const u64*>(src);
    if (basis_switch_tc_applies(bd, logN)) return launch_basis_switch_tc(bd, d, dstT, dst0, s, srcT, src0, items, logN, exact, st);
    return exact ? launch_basis_switch_t<true>(bd, d, dstT, dst0, s, srcT, src0, items, logN, st)
                 : launch_basis_switch_t<false>(bd, d, dstT, dst0, s, srcT, src0, items, logN, st);
}

// DCRTPolyImpl::ScaleAndRound(paramsOutput, ...) (dcrtpoly-impl.h:1876-1960, HAVE_INT128 branch).  Input [B][sI+sO][N];
// in0 / out0 = first input / output tower.  Per coefficient: nu = 0.5 + sum_i frac[i] * double(x_i) (IEEE double, in
// order, not fused), alpha = integer part of nu (64-bit, or 128-bit when nu >= 2^64), and for every output modulus o_j:
// (sum_i x_i * modo[j][i] + x_out_j * modo[j][sI]) mod o_j + alpha mod o_j.
template <int JT, bool SPLIT>
__global__ void __launch_bounds__(BS_THREADS, SPLIT ? FHE_MINB : 2 * FHE_MINB)
scale_round_kernel(u64* __restrict__ dst, const u64* __restrict__ src, size_t items, u32 logN, ScaleDev sd) {
    extern __shared__ __align__(16) u64 sm[];
    constexpr int HW = SPLIT ? 2 : 1;                           // words per table entry
    u64* sM = sm;                                               // [sI+1][JT] (split form: uint4 {h0, h1, h0+h1, -})
    double* sF = reinterpret_cast<double*>(sM + (size_t)(sd.sI + 1) * JT * HW);   // [sI]
    TargetMod* sT = reinterpret_cast<TargetMod*>(sF + sd.sI);                // [JT]
    const u32 j0 = blockIdx.y * JT;
    if (threadIdx.x < JT && j0 + threadIdx.x < sd.sO) {
        const BasisTarget g = sd.tgt[j0 + threadIdx.x];
        sT[threadIdx.x] = TargetMod{g.p, g.c64, g.c64p, g.pinv, g.post, g.postp};
    }
    for (u32 k = threadIdx.x; k < (sd.sI + 1) * JT; k += BS_THREADS) {
        u32 i = k / JT, jj = k % JT;
        u64 h = (j0 + jj < sd.sO) ? sd.modo[(size_t)(j0 + jj) * (sd.sI + 1) + i] : 0ull;
        if (SPLIT) reinterpret_cast<uint4*>(sM)[k] = pack30(h);
        else sM[k] = h;
    }
    for (u32 i = threadIdx.x; i < sd.sI; i += BS_THREADS) sF[i] = sd.frac[i];
    __syncthreads();
    for (size_t e = (size_t)blockIdx.x * BS_THREADS + threadIdx.x; e < items; e += (size_t)gridDim.x * BS_THREADS) {
        const size_t b = e >> logN;
        const u32 n = (u32)e & ((1u << logN) - 1);
        const u32 T = sd.sI + sd.sO;
        const u64* ip = src + (((b * T) + sd.in0) << logN) + n;
        u64 lo[JT], hi[JT];
        double nu = 0.5;
        auto load = [&](u32 i) -> u64 { return ip[(size_t)i << logN]; };
        auto term = [&](u32 i, u64 x) -> u64 {
            nu = __dadd_rn(nu, __dmul_rn(sF[i], __ull2double_rn(x)));
            return x;
        };
        Acc30 acc[SPLIT ? JT : 1];
        if constexpr (SPLIT) {
#pragma unroll
            for (int jj = 0; jj < JT; jj++) acc[jj].clear();
            mac_towers<JT>(acc, sd.sI, reinterpret_cast<const uint4*>(sM), load, term);
        } else {
#pragma unroll
            for (int jj = 0; jj < JT; jj++) { lo[jj] = 0; hi[jj] = 0; }
            u64 x = ip[0];
            for (u32 i = 0; i < sd.sI; i++) {
                u64 xn = (i + 1 < sd.sI) ? ip[(size_t)(i + 1) << logN] : 0ull;
                nu = __dadd_rn(nu, __dmul_rn(sF[i], __ull2double_rn(x)));
#pragma unroll
                for (int jj = 0; jj < JT; jj++) mac128(lo[jj], hi[jj], x, sM[i * JT + jj]);
                x = xn;
            }
        }
        // integer part of nu: utilities.h:133-137 decides between a 64-bit and a 128-bit conversion
        u64 a_lo, a_hi = 0;
        if (nu < 18446744073709551616.0) {
            a_lo = __double2ull_rz(nu);
        } else {
            a_hi = __double2ull_rz(nu * 5.421010862427522e-20);                  // nu / 2^64: exact (power of two)
            a_lo = __double2ull_rz(nu - __ull2double_rn(a_hi) * 18446744073709551616.0);
        }
#pragma unroll
        for (int jj = 0; jj < JT; jj++) {
            if (j0 + jj < sd.sO) {
                const TargetMod tg = sT[jj];
                u64 xo = src[(((b * T) + sd.out0 + j0 + jj) << logN) + n];
                if constexpr (SPLIT) {                   // the output tower's own term: limbs are normalised, one more fits
                    const uint4 h = reinterpret_cast<const uint4*>(sM)[sd.sI * JT + jj];
                    u32 x0, x1;
                    split30(xo, x0, x1);
                    const u64 t0 = (u64)x0 * h.x, t2 = (u64)x1 * h.y, tm = (u64)(x0 + x1) * h.z;
                    Acc30& A = acc[jj];
                    A.a0 += t0; A.a1 += tm - t0 - t2; A.a2 += t2;
                    A.normalise();
                    A.value(lo[jj], hi[jj]);
                } else {
                    mac128(lo[jj], hi[jj], xo, sM[sd.sI * JT + jj]);
                }
                u64 r = reduce128(lo[jj], hi[jj], tg);
                u64 a = reduce128(a_lo, a_hi, tg);
                dst[(((b * sd.sO) + j0 + jj) << logN) + n] = addmod(r, a, tg.p);
            }
        }
    }
}

template <int JT>
static void launch_scale_round_jt(int jt, bool split, int nblk, size_t smem, cudaStream_t st, u64* d, const u64* s, size_t items,
                                  u32 logN, const ScaleDev& sd) {
    if constexpr (JT >= 1) {
        if (jt == JT) {
            const size_t tiles = (items + BS_THREADS - 1) / BS_THREADS;
            if (split) {
                auto k = scale_round_kernel<JT, true>;
                k<<<dim3(persistent_grid_x(k, smem, tiles, nblk), nblk), BS_THREADS, smem, st>>>(d, s, items, logN, sd);
            } else {
                auto k = scale_round_kernel<(JT > 8 ? 8 : JT), false>;
                k<<<dim3(persistent_grid_x(k, smem, tiles, nblk), nblk), BS_THREADS, smem, st>>>(d, s, items, logN, sd);
            }
        } else {
            launch_scale_round_jt<JT - 1>(jt, split, nblk, smem, st, d, s, items, logN, sd);
        }
    }
}

cudaError_t launch_scale_round(const ScaleDev& sd, uint64_t* dst, const uint64_t* src, size_t batch, uint32_t logN,
                               cudaStream_t st) {
    size_t items = batch << logN;
    if (items == 0) return cudaSuccess;
    int jt, nblk;
    column_blocks(sd.sO, sd.narrow != 0, jt, nblk);
    nblk = (int)(sd.sO + jt - 1) / jt;
    size_t smem = ((size_t)(sd.sI + 1) * jt * (sd.narrow ? 2 : 1) + sd.sI) * 8 + (size_t)jt * sizeof(TargetMod);
    launch_scale_round_jt<BS_JT_MAX>(jt, sd.narrow != 0, nblk, smem, st, reinterpret_cast<u64*>(dst),
                                     reinterpret_cast<const u64*>(src), items, logN, sd);
    kernel_launch_count_add(1);
    return cudaGetLastError();
}

// ---- one-pass tails -------------------------------------------------------------------------------------------
// out[b][i][n] over [B][l][N];  x read from [B][xT][N] towers 0..l-1;  v from [B][l][N];  per tower k = {c, c', t, t'}.
//   mode 0 (ApproxModDown :1173)          out = (x - v) * c
//   mode 1 (DropLastElementAndScale :761) out = x * c + v
//   mode 2 (ModReduce :809-810)           out = (x + v * t) * c
//   mode 3 (mode 0 accumulated)           out = out + (x - v) * c      (a key switch added to the ciphertext it relinearises)
template <int MODE>
__global__ void __launch_bounds__(256)
tower_affine_kernel(u64* out, const u64* x, const u64* v, size_t npairs, u32 logN,
                    u32 l, u32 xT, const TowerMod* __restrict__ mods, const ulonglong4* __restrict__ k) {
    for (size_t p = (size_t)blockIdx.x * 256 + threadIdx.x; p < npairs; p += (size_t)gridDim.x * 256) {
        size_t e = 2 * p;
        size_t unit = e >> logN;                             // b * l + i
        u32 i = (u32)(unit % l);
        size_t b = unit / l;
        u32 n = (u32)e & ((1u << logN) - 1);
        const u64 q = mods[i].q;
        const ulonglong4 c = k[i];
        ulonglong2 xv = *reinterpret_cast<const ulonglong2*>(x + (((b * xT) + i) << logN) + n);
        ulonglong2 vv = *reinterpret_cast<const ulonglong2*>(v + e);
        auto f = [&](u64 xx, u64 yy) -> u64 {
            if (MODE == 0 || MODE == 3) return shoup(submod(xx, yy, q), c.x, c.y, q);
            if (MODE == 1) return addmod(shoup(xx, c.x, c.y, q), yy, q);
            return shoup(addmod(xx, shoup(yy, c.z, c.w, q), q), c.x, c.y, q);
        };
        ulonglong2 r = make_ulonglong2(f(xv.x, vv.x), f(xv.y, vv.y));
        if (MODE == 3) {
            const ulonglong2 acc = *reinterpret_cast<const ulonglong2*>(out + e);
            r = make_ulonglong2(addmod(acc.x, r.x, q), addmod(acc.y, r.y, q));
        }
        *reinterpret_cast<ulonglong2*>(out + e) = r;
    }
}

// scalar form of the same for N = 1 (no aligned pairs)
template <int MODE>
__global__ void tower_affine_scalar_kernel(u64* out, const u64* x, const u64* v,
                                           size_t total, u32 logN, u32 l, u32 xT, const TowerMod* __restrict__ mods,
                                           const ulonglong4* __restrict__ k) {
    for (size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (size_t)gridDim.x * blockDim.x) {
        size_t unit = e >> logN;
        u32 i = (u32)(unit % l);
        size_t b = unit / l;
        u32 n = (u32)e & ((1u << logN) - 1);
        const u64 q = mods[i].q;
        const ulonglong4 c = k[i];
        u64 xx = x[(((b * xT) + i) << logN) + n], yy = v[e];
        u64 r;
        if (MODE == 0) r = shoup(submod(xx, yy, q), c.x, c.y, q);
        else if (MODE == 1) r = addmod(shoup(xx, c.x, c.y, q), yy, q);
        else r = shoup(addmod(xx, shoup(yy, c.z, c.w, q), q), c.x, c.y, q);
        out[e] = r;
    }
}

cudaError_t launch_tower_affine(int mode, uint64_t* out, const uint64_t* x, const uint64_t* v, size_t batch, uint32_t logN,
                                uint32_t l, uint32_t xT, const TowerMod* mods, const ulonglong4* k, cudaStream_t st) {
    size_t total = (batch * l) << logN;
    if (total == 0) return cudaSuccess;
    if (mode == 3 && logN == 0) return cudaErrorInvalidValue;        // the accumulating form exists for N >= 2 only
    u64* o = reinterpret_cast<u64*>(out);
    const u64* xx = reinterpret_cast<const u64*>(x);
    const u64* vv = reinterpret_cast<const u64*>(v);
    if (logN == 0) {
        unsigned grid = (unsigned)((total + 255) / 256 > 4096 ? 4096 : (total + 255) / 256);
        if (mode == 0) tower_affine_scalar_kernel<0><<<grid, 256, 0, st>>>(o, xx, vv, total, logN, l, xT, mods, k);
        else if (mode == 1) tower_affine_scalar_kernel<1><<<grid, 256, 0, st>>>(o, xx, vv, total, logN, l, xT, mods, k);
        else tower_affine_scalar_kernel<2><<<grid, 256, 0, st>>>(o, xx, vv, total, logN, l, xT, mods, k);
    } else {
        size_t npairs = total / 2;
        size_t g = (npairs + 255) / 256;
        unsigned grid = (unsigned)(g > 148 * 64 ? 148 * 64 : g);
        if (mode == 0) tower_affine_kernel<0><<<grid, 256, 0, st>>>(o, xx, vv, npairs, logN, l, xT, mods, k);
        else if (mode == 3) tower_affine_kernel<3><<<grid, 256, 0, st>>>(o, xx, vv, npairs, logN, l, xT, mods, k);
        else if (mode == 1) tower_affine_kernel<1><<<grid, 256, 0, st>>>(o, xx, vv, npairs, logN, l, xT, mods, k);
        else tower_affine_kernel<2><<<grid, 256, 0, st>>>(o, xx, vv, npairs, logN, l, xT, mods, k);
    }
    kernel_launch_count_add(1);
    return cudaGetLastError();
}

// tmp[b][i][n] (i < l) from last[b][n] (COEFFICIENT form of the dropped tower, modulus ql):
//   bgv = 0 (DropLastElementAndScale :756-758): SwitchModulus(ql -> q_i), then * c_i
//   bgv = 1 (ModReduce :798, :805-806):         * negtInv mod ql first, then SwitchModulus(ql -> q_i)
// sw[i] = {half = ql >> 1, q_i, diff, up};  k[i] = {c_i, c_i'};  pre = {negtInv, companion}
__global__ void __launch_bounds__(256)
rescale_spread_kernel(u64* __restrict__ tmp, const u64* __restrict__ last, size_t items, u32 logN, u32 l, u64 ql,
                      const ulonglong4* __restrict__ sw, const ulonglong4* __restrict__ k, ulonglong2 pre, int bgv) {
    for (size_t e = (size_t)blockIdx.x * 256 + threadIdx.x; e < items; e += (size_t)gridDim.x * 256) {
        size_t b = e >> logN;
        u32 n = (u32)e & ((1u << logN) - 1);
        u64 v = last[e];
        if (bgv) v = shoup(v, pre.x, pre.y, ql);
        for (u32 i = blockIdx.y; i < l; i += gridDim.y) {      // blockIdx.y: towers split between CTAs for small inputs
            const ulonglong4 s = sw[i];
            u64 r = switch_modulus(v, s.x, s.y, s.z, s.w != 0);
            if (!bgv) { const ulonglong4 c = k[i]; r = shoup(r, c.x, c.y, s.y); }
            tmp[(((b * l) + i) << logN) + n] = r;
        }
    }
}

cudaError_t launch_rescale_spread(uint64_t* tmp, const uint64_t* last, size_t batch, uint32_t logN, uint32_t l, uint64_t ql,
                                  const ulonglong4* sw, const ulonglong4* k, ulonglong2 pre, int bgv, cudaStream_t st) {
    size_t items = batch << logN;
    if (items == 0 || l == 0) return cudaSuccess;
    size_t g = (items + 255) / 256;
    dim3 grid((unsigned)(g > 148 * 64 ? 148 * 64 : g), 1);
    if (g < 148 * 4 && tower_split_on()) grid.y = (unsigned)std::min<size_t>(l, (148 * 4 + g - 1) / g);
    rescale_spread_kernel<<<grid, 256, 0, st>>>(reinterpret_cast<u64*>(tmp), reinterpret_cast<const u64*>(last), items, logN, l,
                                               ql, sw, k, pre, bgv);
    kernel_launch_count_add(1);
    return cudaGetLastError();
}


// ---- DCRTPolyImpl::CRTDecompose(0) (dcrtpoly-impl.h:266-288), the digits of KeySwitchBV -------------------------------
// coef [batch][T][N] (COEFFICIENT) -> out [T][batch][T][N]: digit i, tower k = SwitchModulus(coef tower i, q_i -> q_k), and
// the tower itself for k = i; the caller then transforms every tower of out (for k = i that restores the stored
// EVALUATION tower exactly).  sw[i * T + k] = {half = q_i >> 1, q_k, diff, up}.  One thread = one coefficient pair of one
// source tower, T 16-byte stores: HBM-bound on the writes (8 B read, 8 T B written per coefficient).
__global__ void __launch_bounds__(256)
crt_spread_kernel(u64* __restrict__ out, const u64* __restrict__ coef, size_t pairs, u32 logN, u32 T, size_t batch,
                  const ulonglong4* __restrict__ sw) {
    const u32 half = logN - 1;
    for (size_t e = (size_t)blockIdx.x * 256 + threadIdx.x; e < pairs; e += (size_t)gridDim.x * 256) {
        const size_t unit = e >> half;                         // (batch item, source tower)
        const u32 i = (u32)(unit % T);
        const size_t b = unit / T;
        const u32 n = (u32)((e & ((1u << half) - 1)) << 1);
        const ulonglong2 v = *reinterpret_cast<const ulonglong2*>(coef + (unit << logN) + n);
        u64* o = out + ((((size_t)i * batch + b) * T) << logN) + n;
        for (u32 k = 0; k < T; k++) {
            ulonglong2 r = v;
            if (k != i) {
                const ulonglong4 s = sw[(size_t)i * T + k];
                r.x = switch_modulus(v.x, s.x, s.y, s.z, s.w != 0);
                r.y = switch_modulus(v.y, s.x, s.y, s.z, s.w != 0);
            }
            *reinterpret_cast<ulonglong2*>(o + ((size_t)k << logN)) = r;
        }
    }
}

cudaError_t launch_crt_spread(uint64_t* out, const uint64_t* coef, size_t batch, uint32_t logN, uint32_t T, const ulonglong4* sw,
                              cudaStream_t st) {
    if (batch == 0 || T == 0) return cudaSuccess;
    if (logN < 1) return cudaErrorInvalidValue;
    const size_t pairs = (batch * T) << (logN - 1);
    size_t g = (pairs + 255) / 256;
    unsigned grid = (unsigned)(g > 148 * 32 ? 148 * 32 : g);
    crt_spread_kernel<<<grid, 256, 0, st>>>(reinterpret_cast<u64*>(out), reinterpret_cast<const u64*>(coef), pairs, logN, T, batch, sw);
    kernel_launch_count_add(1);
    return cudaGetLastError();
}

// ---- BFV HPSPOVERQ tails ----------------------------------------------------------------------------------------
// One thread per coefficient pair of one batch item, towers in a loop (every row segment is contiguous in n).
// k[i] = {c_i, c_i', M_i = q_i * floor(2^63 / q_i), -}.
//   mode 0 (ExpandCRTBasisQlHat, dcrtpoly-impl.h:1514-1534): x over [B][sl][N] -> out over [B][T][N]:
//          out_i = x_i * c_i (i < sl), 0 (sl <= i < T)
//   mode 1 (ScaleAndRoundPOverQ, dcrtpoly-impl.h:2052-2066): x over [B][T+1][N] -> out over [B][T][N]:
//          out_i = (x_i - (x_T mod q_i)) * c_i; the last tower's residue can exceed q_i (ModSubEq reduces it first).
//          Here the difference is formed as x_i + M_i - x_T with M_i the multiple of q_i next below 2^63: x_T < 2^62 <= M_i,
//          the sum stays below 2^64, and the Shoup product is exact for any 64-bit operand — one product per coefficient
template <int MODE>
__global__ void __launch_bounds__(256)
tower_scale_kernel(u64* out, const u64* x, size_t npairs, u32 logN, u32 sl, u32 T,
                   const TowerMod* __restrict__ mods, const ulonglong4* __restrict__ k) {
    const u32 xT = MODE == 0 ? sl : T + 1;
    for (size_t p = (size_t)blockIdx.x * 256 + threadIdx.x; p < npairs; p += (size_t)gridDim.x * 256) {
        size_t e = 2 * p;
        size_t b = e >> logN;
        u32 n = (u32)e & ((1u << logN) - 1);
        const u64* xb = x + ((b * xT) << logN) + n;
        u64* ob = out + ((b * T) << logN) + n;
        ulonglong2 last = make_ulonglong2(0, 0);
        if (MODE == 1) last = *reinterpret_cast<const ulonglong2*>(xb + ((size_t)T << logN));
        // towers four at a time: the loads of a group are issued together (4 x 16 B in flight per thread — one tower
        // at a time leaves too little in flight to cover HBM latency)
        constexpr u32 U = 4;
        const u32 nread = MODE == 1 ? T : sl;
        // blockIdx.y splits the towers between CTAs when there are few coefficients (one poly of a scheme-level caller):
        // one thread walking all 30-40 towers of its coefficient pair left a B200 with 32 busy CTAs
        for (u32 i0 = blockIdx.y * U; i0 < T; i0 += gridDim.y * U) {
            ulonglong2 xv[U];
#pragma unroll
            for (u32 u = 0; u < U; u++)
                if (i0 + u < nread) xv[u] = *reinterpret_cast<const ulonglong2*>(xb + ((size_t)(i0 + u) << logN));
#pragma unroll
            for (u32 u = 0; u < U; u++) {
                const u32 i = i0 + u;
                if (i >= T) break;
                ulonglong2 r = make_ulonglong2(0, 0);
                if (i < nread) {
                    const u64 q = mods[i].q;
                    const ulonglong4 c = k[i];
                    ulonglong2 v = xv[u];
                    if (MODE == 1) {                    // x_i + M_i - x_p: congruent to x_i - x_p, non-negative, below 2^64
                        v.x = v.x + c.z - last.x;
                        v.y = v.y + c.z - last.y;
                    }
                    r = make_ulonglong2(shoup(v.x, c.x, c.y, q), shoup(v.y, c.x, c.y, q));
                }
                *reinterpret_cast<ulonglong2*>(ob + ((size_t)i << logN)) = r;
            }
        }
    }
}

// scalar form for N = 1
template <int MODE>
__global__ void tower_scale_scalar_kernel(u64* out, const u64* x, size_t batch, u32 sl, u32 T,
                                          const TowerMod* __restrict__ mods, const ulonglong4* __restrict__ k) {
    const u32 xT = MODE == 0 ? sl : T + 1;
    for (size_t b = (size_t)blockIdx.x * blockDim.x + threadIdx.x; b < batch; b += (size_t)gridDim.x * blockDim.x)
        for (u32 i = 0; i < T; i++) {
            u64 r = 0;
            if (MODE == 1 || i < sl) {
                const u64 q = mods[i].q;
                const ulonglong4 c = k[i];
                u64 xv = x[b * xT + i];
                if (MODE == 1) xv = xv + c.z - x[b * xT + T];
                r = shoup(xv, c.x, c.y, q);
            }
            out[b * T + i] = r;
        }
}

cudaError_t launch_tower_scale(int mode, uint64_t* out, const uint64_t* x, size_t batch, uint32_t logN, uint32_t sl,
                               uint32_t T, const TowerMod* mods, const ulonglong4* k, cudaStream_t st) {
    if (batch == 0 || T == 0) return cudaSuccess;
    u64* o = reinterpret_cast<u64*>(out);
    const u64* xx = reinterpret_cast<const u64*>(x);
    if (logN == 0) {
        unsigned grid = (unsigned)((batch + 255) / 256 > 4096 ? 4096 : (batch + 255) / 256);
        if (mode == 0) tower_scale_scalar_kernel<0><<<grid, 256, 0, st>>>(o, xx, batch, sl, T, mods, k);
        else tower_scale_scalar_kernel<1><<<grid, 256, 0, st>>>(o, xx, batch, sl, T, mods, k);
    } else {
        size_t npairs = (batch << logN) / 2;
        size_t g = (npairs + 255) / 256;
        dim3 grid((unsigned)(g > 148 * 64 ? 148 * 64 : g), 1);
        if (g < 148 * 4 && tower_split_on()) grid.y = (unsigned)std::min<size_t>((T + 3) / 4, (148 * 4 + g - 1) / g);
        if (mode == 0) tower_scale_kernel<0><<<grid, 256, 0, st>>>(o, xx, npairs, logN, sl, T, mods, k);
        else tower_scale_kernel<1><<<grid, 256, 0, st>>>(o, xx, npairs, logN, sl, T, mods, k);
    }
    kernel_launch_count_add(1);
    return cudaGetLastError();
}



// ------------------------------------------------------------------------------------------------------------------
// The same conversion on the 5th-generation tensor cores (tcgen05.mma kind::i8, accumulators in TMEM).
//
// sum_i y_i * [Q/q_i]_{p_j} for all coefficients of a tile and all target towers IS a matrix product — the one dense
// contraction on this path.  Both factors are cut into 8-bit limbs, y_i = sum_a y_ia 2^(8a), h_ij = sum_b h_ijb 2^(8b):
//     sum_i y_i h_ij = sum_d 2^(8d) * D[j][d],    D[j][d] = sum_i sum_{a+b=d} y_ia * h_ijb        (d = 0..14)
// and D = A x B^T with
//     A [128 coefficients][K = 8 * sq]      A[m][8i + a] = y_ia                       (built per tile in shared memory)
//     B [16 * targets]  [K]                 B[16j + d][8i + a] = h_ij(d-a) or 0       (a Toeplitz expansion of the table,
//                                                                                      built once per basis on the host)
// u8 x u8 products accumulated in s32: sq <= 32 gives 32 * 8 * 255^2 < 2^25 per entry, far from overflow.  The epilogue
// reads the 15 diagonals of a target from TMEM (tcgen05.ld), joins them into the 128-bit sum and reduces it with the same
// reduce128 as the integer kernel, so every value stored is the canonical residue the CPU path produces.
//
// One CTA = 128 threads = one 128-coefficient tile at a time (thread m owns row m: it forms the y_i of its coefficient
// with Shoup products, writes them as 16-byte rows of the canonical K-major core-matrix layout — no swizzle: chunk kc of
// 16 bytes of K at kc * 2048 + m * 16 — and later reads lane m of TMEM).  One elected thread issues K/32 MMAs
// (M = 128, N = 16 * targets <= 256, K = 32 each) and commits them to an mbarrier.  Two CTAs per SM (256 TMEM columns
// each) overlap one another's phases.  Operands: shared memory descriptors built by hand (version 1, SWIZZLE_NONE,
// leading-dimension byte offset = the distance of the two K chunks, stride-dimension byte offset = 128 = one 8-row core
// matrix), instruction descriptor = {C s32, A/B unsigned 8-bit, K-major both, N >> 3, M >> 4}.
// ------------------------------------------------------------------------------------------------------------------
constexpr int TC_M = 128;

__device__ __forceinline__ u32 tc_smem(const void* p) { return (u32)__cvta_generic_to_shared(p); }
__device__ __forceinline__ u64 tc_desc(u32 saddr, u32 lbo_bytes, u32 sbo_bytes) {
    return (u64)((saddr >> 4) & 0x3fffu) | ((u64)((lbo_bytes >> 4) & 0x3fffu) << 16) | ((u64)((sbo_bytes >> 4) & 0x3fffu) << 32) |
           (1ull << 46);                           // version 1 (Blackwell); base offset 0, SWIZZLE_NONE
}

// A CTA holds SUB independent sub-CTAs of 128 * TPR threads (own A tile, own accumulator columns in TMEM, own mbarrier,
// own named barrier), which share the B matrix and the constants: SUB tiles are in flight per SM, and while one sub-CTA
// waits for its MMAs or reads TMEM the others run their Shoup products.  TPR threads share a row of a tile: thread
// `st` of a sub-CTA owns row m = st % 128 together with st + 128, ...; they split the source towers of the prologue
// (pairs of towers, round robin) and the targets of the epilogue between them (all of them sit in warps with the same
// index modulo 4, i.e. in the TMEM lane quarter of their row).
template <bool EXACT, int TPR, int SUB>
__global__ void __launch_bounds__(TC_M * TPR * SUB, SUB == 1 ? 2 : 1)
basis_switch_tc_kernel(u64* __restrict__ dst, const u64* __restrict__ src, u32 tiles, u32 logN, BasisDev bd, u32 srcT, u32 src0,
                       u32 dstT, u32 dst0) {
    constexpr int STHREADS = TC_M * TPR, THREADS = STHREADS * SUB;
    constexpr int PP = (16 + TPR - 1) / TPR;                         // tower pairs per thread (sq <= 32)
    constexpr u32 TCOLS = SUB == 1 ? 256u : 512u;                    // SUB = 1: two CTAs per SM, half of TMEM each
    extern __shared__ __align__(1024) unsigned char tcsm[];
    const u32 kpad = bd.tc_kpad, jg = bd.tc_jg, ngemm = 16 * jg;
    const u32 cstride = (ngemm + 31) & ~31u;                         // accumulator columns per sub-CTA
    unsigned char* sB = tcsm;                                        // [kpad/16][ngemm][16], shared by the sub-CTAs
    unsigned char* sAall = sB + (size_t)kpad * ngemm;                // SUB x [kpad/16][128][16]
    u64* sQ = reinterpret_cast<u64*>(sAall + (size_t)SUB * kpad * TC_M);   // [32]
    u64* sC = sQ + 32;                                               // [32][2]
    double* sI = reinterpret_cast<double*>(sC + 64);                 // [32]
    TargetMod* sT = reinterpret_cast<TargetMod*>(sI + 32);           // [16]
    u64* bars = reinterpret_cast<u64*>(sT + 16);                     // [SUB]
    u32* tslot = reinterpret_cast<u32*>(bars + 4);
    double* sPall = reinterpret_cast<double*>(bars + 6);             // EXACT: SUB x [32][128] the terms of nu, summed in index order
    const u32 tid = threadIdx.x, sub = tid / STHREADS, st = tid % STHREADS;
    const u32 warp = tid >> 5, row = st & (TC_M - 1), part = st >> 7;
    unsigned char* sA = sAall + (size_t)sub * kpad * TC_M;
    double* sP = sPall + (size_t)sub * 32 * TC_M;
    u64* bar = bars + sub;
    const u32 j0 = blockIdx.y * jg;
    const u32 nj = bd.sp - j0 < jg ? bd.sp - j0 : jg;                // targets of this group

    // ---- setup: TMEM, barriers, the group's B matrix, constants ----
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tc_smem(tslot)), "r"(TCOLS));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    if (tid < SUB) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(tc_smem(bars + tid)));
        asm volatile("fence.mbarrier_init.release.cluster;");
    }
    {
        const uint4* gB = reinterpret_cast<const uint4*>(bd.tc_b + (size_t)blockIdx.y * kpad * ngemm);
        uint4* s4 = reinterpret_cast<uint4*>(sB);
        for (u32 k = tid; k < kpad * ngemm / 16; k += THREADS) s4[k] = gB[k];
        uint4* a4 = reinterpret_cast<uint4*>(sAall);
        for (u32 k = tid; k < SUB * kpad * TC_M / 16; k += THREADS) a4[k] = make_uint4(0, 0, 0, 0);     // K padding stays zero
    }
    for (u32 i = tid; i < 32; i += THREADS) {
        const bool in = i < bd.sq;
        sQ[i] = in ? bd.q[i] : 4ull;
        sC[2 * i] = in ? bd.pre[i].x : 0ull;
        sC[2 * i + 1] = in ? bd.pre[i].y : 0ull;
        sI[i] = (EXACT && in) ? bd.qinv[i] : 0.0;
    }
    if (tid < nj) {
        const BasisTarget g = bd.tgt[j0 + tid];
        sT[tid] = TargetMod{g.p, g.c64, g.c64p, g.pinv, g.post, g.postp};
    }
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;");
    const u32 tmem = *tslot + sub * cstride;
    const u32 idesc = (2u << 4) | ((ngemm >> 3) << 17) | ((TC_M >> 4) << 24);      // C = s32, A / B unsigned 8-bit, K-major
    const u32 aBase = tc_smem(sA), bBase = tc_smem(sB);
    const u32 npairs = (bd.sq + 1) >> 1;
    auto sub_sync = [&]() { asm volatile("bar.sync %0, %1;" ::"r"(sub + 1), "n"(STHREADS) : "memory"); };
    u32 phase = 0;
    const u32 tstep = gridDim.x * SUB, tfirst = blockIdx.x * SUB + sub;

    // the residues of a tile are fetched one tile ahead: the loads of tile t + 1 fly during the MMA and the epilogue of tile t
    u64 xr[2 * PP];
    auto fetch = [&](u32 t) {
        const size_t e = (size_t)t * TC_M + row;
        const u64* sp = src + ((((e >> logN) * srcT) + src0) << logN) + ((u32)e & ((1u << logN) - 1));
#pragma unroll
        for (int k = 0; k < PP; k++) {
            const u32 i = 2 * (part + k * TPR);
            xr[2 * k] = i < bd.sq ? sp[(size_t)i << logN] : 0ull;
            xr[2 * k + 1] = i + 1 < bd.sq ? sp[(size_t)(i + 1) << logN] : 0ull;
        }
    };
    if (tfirst < tiles) fetch(tfirst);
    for (u32 t = tfirst; t < tiles; t += tstep) {
        const size_t e = (size_t)t * TC_M + row;
        const size_t b = e >> logN;
        const u32 n = (u32)e & ((1u << logN) - 1);
        // ---- prologue: y_i of this row, two towers per 16-byte row chunk ----
#pragma unroll
        for (int k = 0; k < PP; k++) {
            const u32 pr = part + k * TPR, i = 2 * pr;
            if (pr < npairs) {
                const u64 y0 = shoup(xr[2 * k], sC[2 * i], sC[2 * i + 1], sQ[i]);
                const u64 y1 = i + 1 < bd.sq ? shoup(xr[2 * k + 1], sC[2 * i + 2], sC[2 * i + 3], sQ[i + 1]) : 0ull;
                if (EXACT) {
                    sP[i * TC_M + row] = __dmul_rn(__ull2double_rn(y0), sI[i]);
                    sP[(i + 1) * TC_M + row] = __dmul_rn(__ull2double_rn(y1), sI[i + 1]);
                }
                *reinterpret_cast<ulonglong2*>(sA + (size_t)pr * (TC_M * 16) + row * 16) = make_ulonglong2(y0, y1);
            }
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");      // generic-proxy writes -> visible to the tensor core
        asm volatile("tcgen05.fence::before_thread_sync;");
        sub_sync();
        // ---- one thread issues the MMAs of the tile and commits them to the sub-CTA's barrier ----
        if (st == 0) {
            asm volatile("tcgen05.fence::after_thread_sync;");
            for (u32 kb = 0; kb < kpad / 32; kb++) {
                const u64 da = tc_desc(aBase + kb * 2 * (TC_M * 16), TC_M * 16, 128);
                const u64 db = tc_desc(bBase + kb * 2 * (ngemm * 16), ngemm * 16, 128);
                const u32 acc = kb ? 1u : 0u;
                asm volatile(
                    "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                    "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}"
                    ::"r"(tmem), "l"(da), "l"(db), "r"(idesc), "r"(acc) : "memory");
            }
            asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(tc_smem(bar)) : "memory");
        }
        if (t + tstep < tiles) fetch(t + tstep);
        u32 alpha = 0;
        if (EXACT) {                        // nu = 0.5 + sum_i double(y_i) * (1 / q_i), IEEE double in index order (:1197-1203)
            double nu = 0.5;
            for (u32 i = 0; i < bd.sq; i++) nu = __dadd_rn(nu, sP[i * TC_M + row]);
            alpha = (u32)nu;
        }
        {   // everyone waits for the accumulators (bounded: a barrier that never completes traps instead of hanging the GPU)
            const u32 baddr = tc_smem(bar);
            u32 done = 0, spins = 0;
            while (!done) {
                asm volatile(
                    "{\n\t.reg .pred P1;\n\t"
                    "mbarrier.try_wait.parity.shared::cta.b64 P1, [%1], %2, %3;\n\t"
                    "selp.u32 %0, 1, 0, P1;\n\t}" : "=r"(done) : "r"(baddr), "r"(phase), "r"(1000000u) : "memory");
                if (!done && ++spins > 4000u) __trap();
            }
            phase ^= 1;
        }
        asm volatile("tcgen05.fence::after_thread_sync;");
        // ---- epilogue: the 15 diagonals of this thread's targets, from the TMEM lane of its row ----
        auto tload = [&](u32 (&d)[16], u32 jj) {
            const u32 taddr = tmem + (((warp & 3u) * 32u) << 16) + jj * 16u;
            asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
                         : "=r"(d[0]), "=r"(d[1]), "=r"(d[2]), "=r"(d[3]), "=r"(d[4]), "=r"(d[5]), "=r"(d[6]), "=r"(d[7]), "=r"(d[8]),
                           "=r"(d[9]), "=r"(d[10]), "=r"(d[11]), "=r"(d[12]), "=r"(d[13]), "=r"(d[14]), "=r"(d[15])
                         : "r"(taddr));
        };
        auto finish = [&](const u32 (&d)[16], u32 jj) {
            // sum_d d[d] * 2^(8d).  Every d[] is below 2^25, so the diagonals 4 apart tile a 128-bit number without overlap:
            // N_r = d[r] + d[4+r] 2^32 + d[8+r] 2^64 + d[12+r] 2^96 costs nothing, and the sum is N_0 + N_1 2^8 + N_2 2^16 + N_3 2^24
            // (three 128-bit shifts and additions; the total is below 2^125, so nothing is shifted out)
            auto pack = [](u32 l, u32 h) { u64 v; asm("mov.b64 %0, {%1, %2};" : "=l"(v) : "r"(l), "r"(h)); return v; };
            u64 lo = pack(d[0], d[4]), hi = pack(d[8], d[12]);
#pragma unroll
            for (int r = 1; r < 4; r++) {
                const u64 l = pack(d[r], d[4 + r]), h = pack(d[8 + r], r == 3 ? 0u : d[12 + r]);
                const u64 ls = l << (8 * r), hs = (h << (8 * r)) | (l >> (64 - 8 * r));
                lo += ls;
                hi += hs + (lo < ls ? 1ull : 0ull);
            }
            const TargetMod tg = sT[jj];
            u64 r = reduce128(lo, hi, tg);
            if (tg.post) r = shoup(r, tg.post, tg.postp, tg.p);
            if (EXACT) r = submod(r, bd.alphaq[(size_t)alpha * bd.sp + j0 + jj], tg.p);
            dst[(((b * dstT) + dst0 + j0 + jj) << logN) + n] = r;
        };
        // tcgen05.ld is warp-collective (.sync.aligned): every warp runs the same number of rounds, guarded per target
        for (u32 base = 0; base < nj; base += 2 * TPR) {
            const u32 ja = base + part, jb = base + TPR + part;
            u32 d0[16], d1[16];
            if (ja < nj) tload(d0, ja);
            if (jb < nj) tload(d1, jb);
            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
            if (ja < nj) finish(d0, ja);
            if (jb < nj) finish(d1, jb);
        }
        asm volatile("tcgen05.fence::before_thread_sync;");
        sub_sync();                          // the sub-CTA's TMEM columns and A tile are free for its next tile
    }
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(*tslot), "r"(TCOLS));
}

bool basis_switch_tc_applies(const BasisDev& bd, uint32_t logN) {
    static const int mode = env_int("FHE_B200_BASIS_TC", 1);
    return mode != 0 && bd.tc_b != nullptr && logN >= 7;
}

template <bool EXACT, int TPR, int SUB>
static cudaError_t launch_basis_switch_tc_t(const BasisDev& bd, u64* d, uint32_t dstT, uint32_t dst0, const u64* s, uint32_t srcT,
                                            uint32_t src0, size_t tiles, uint32_t logN, cudaStream_t st) {
    const u32 ngemm = 16 * bd.tc_jg;
    size_t smem = (size_t)bd.tc_kpad * (SUB * TC_M + ngemm) + 32 * 32 + 16 * sizeof(TargetMod) + 48 + (EXACT ? (size_t)SUB * 32 * TC_M * 8 : 0) + 64;
    // SUB > 1: one CTA per SM (it takes all 512 TMEM columns); SUB = 1: two CTAs per SM with 256 columns each, never three
    const size_t floor_bytes = SUB == 1 ? 80 * 1024 : 116 * 1024;
    if (smem < floor_bytes) smem = floor_bytes;
    if (smem > 227 * 1024) return cudaErrorInvalidValue;
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    auto kern = basis_switch_tc_kernel<EXACT, TPR, SUB>;
    // The attribute is per function and device, and `smem` depends on the plan: two host threads that set it per launch to
    // their own sizes can lower it under each other's launch (seen as "invalid argument" from 16 threads).  It only bounds
    // what a launch may ask for, so it is raised to the device's limit once per device.
    static std::atomic<unsigned long long> attr_done{0};
    static std::atomic<int> sm_cache[64];
    int sms = dev < 64 ? sm_cache[dev].load(std::memory_order_relaxed) : 0;
    if (dev >= 64 || !((attr_done.load(std::memory_order_acquire) >> dev) & 1) || sms <= 0) {
        e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
        if (e != cudaSuccess) return e;
        sms = 148;
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        if (dev < 64) {
            sm_cache[dev].store(sms, std::memory_order_relaxed);
            attr_done.fetch_or(1ull << dev, std::memory_order_release);
        }
    }
    const size_t per_group = std::max<size_t>(1, (size_t)(SUB == 1 ? 2 : 1) * sms / bd.tc_groups);
    dim3 grid((unsigned)std::min((tiles + SUB - 1) / SUB, per_group), bd.tc_groups);
    kern<<<grid, TC_M * TPR * SUB, smem, st>>>(d, s, (u32)tiles, logN, bd, srcT, src0, dstT, dst0);
    kernel_launch_count_add(1);
    return cudaGetLastError();
}

static cudaError_t launch_basis_switch_tc(const BasisDev& bd, u64* d, uint32_t dstT, uint32_t dst0, const u64* s, uint32_t srcT,
                                          uint32_t src0, size_t items, uint32_t logN, bool exact, cudaStream_t st) {
    const size_t tiles = items / TC_M;                     // N >= 128: tiles never straddle a polynomial
    if (tiles > 0xffffffffull) return cudaErrorInvalidValue;
    // accumulators of 16 columns per target: three sub-CTAs fit the 512 TMEM columns up to 10 targets per group, two beyond
    static const int sub_env = env_int("FHE_B200_BASIS_TC_SUB", 0);
    const u32 cstride = (16 * bd.tc_jg + 31) & ~31u;
    (void)cstride;
    int sub = 1;                                           // measured best on the B200: two CTAs per SM, one tile each
    if (sub_env >= 1 && sub_env <= 3 && ((u32)sub_env * cstride <= 512)) sub = sub_env;
    if (exact && sub == 3 && (size_t)bd.tc_kpad * (3 * TC_M + 16 * bd.tc_jg) + 3 * 32 * TC_M * 8 + 4096 > 227 * 1024) sub = 2;
#define TC_GO(E, S) launch_basis_switch_tc_t<E, 2, S>(bd, d, dstT, dst0, s, srcT, src0, tiles, logN, st)
    if (exact) return sub == 3 ? TC_GO(true, 3) : sub == 2 ? TC_GO(true, 2) : TC_GO(true, 1);
    return sub == 3 ? TC_GO(false, 3) : sub == 2 ? TC_GO(false, 2) : TC_GO(false, 1);
#undef TC_GO
}

}  // namespace fhe
