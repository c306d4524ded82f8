// basis.cu — RNS basis conversion and scaling kernels for sm_100a.
//
// Device replacement for (citations relative to the reference tree, lattice/hal/default/dcrtpoly-impl.h):
//   DCRTPolyImpl::ApproxSwitchCRTBasis   :1034-1082  (HAVE_INT128 branch; Mul128 / BarrettUint128ModUint64,
//                                                     utils/utilities-int.h:47-110)
//   the element-wise tails of ApproxModDown :1166-1174, DropLastElementAndScale :753-767, ModReduce :803-811
//   NativeVectorT::SwitchModulus           math/hal/intnat/mubintvecnat.cpp:112-135
//
// basis_switch_kernel: one thread per coefficient index.  For each source tower i it forms
// y_i = x_i * [(Q/q_i)^-1]_{q_i} mod q_i (Shoup, canonical) and accumulates y_i * [Q/q_i]_{p_j} into 128-bit sums for
// JT target towers at a time (IMAD.WIDE chains with carry), then reduces each sum modulo p_j with the reference's
// 128-by-64 Barrett step.  Every value stored is the canonical residue the CPU path produces.  The [Q/q_i]_{p_j}
// slice, the Shoup constants and the moduli sit in shared memory.  Integer-pipe bound: 4 IMAD.WIDE per (i, j) pair
// per coefficient; HBM traffic is 8 B per source tower and 8 B per target tower per coefficient.
// tower_affine_kernel / rescale_spread_kernel: one pass, 128-bit loads and stores, HBM-bound.
#include "modarith.cuh"
#include "kernels.h"

namespace fhe {

// acc += a * b on a 128-bit accumulator.  Written with unsigned __int128 so that ptxas emits the four 32x32 partial
// products once (4 IMAD.WIDE.U32 with carry chains); the mad.lo.cc / madc.hi PTX pair recomputes the low product
// (about 6 wide multiplies per MAC, measured with tools/sass_mix.py).
__device__ __forceinline__ void mac128(u64& lo, u64& hi, u64 a, u64 b) {
    unsigned __int128 acc = ((unsigned __int128)hi << 64) | lo;
    acc += (unsigned __int128)a * b;
    lo = (u64)acc;
    hi = (u64)(acc >> 64);
}

// utils/utilities-int.h:60-110: a mod m for a 128-bit a with mu = floor(2^128 / m)
__device__ __forceinline__ u64 barrett128(u64 a_lo, u64 a_hi, u64 m, u64 mu_lo, u64 mu_hi) {
    u64 left_hi = __umul64hi(a_lo, mu_lo);
    u64 mlo = a_lo * mu_hi, mhi = __umul64hi(a_lo, mu_hi);
    u64 tmp1 = mlo + left_hi;
    u64 tmp2 = mhi + (tmp1 < mlo ? 1ull : 0ull);
    u64 m2lo = a_hi * mu_lo, m2hi = __umul64hi(a_hi, mu_lo);
    u64 left2 = m2hi + ((u64)(m2lo + tmp1) < m2lo ? 1ull : 0ull);
    u64 t = a_hi * mu_hi + tmp2 + left2;
    u64 r = a_lo - t * m;
    while (r >= m) r -= m;
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

constexpr int BS_THREADS = 256;

// EXACT: DCRTPolyImpl::SwitchCRTBasis (dcrtpoly-impl.h:1178-1230): additionally counts the q-overflows
// alpha = (unsigned)(0.5 + sum_i double(y_i) * qinv[i]) in IEEE double (round-to-nearest multiply and add, in index
// order, never fused — what the CPU path computes) and subtracts [alpha*Q]_{p_j}.
template <int JT, bool EXACT>
__global__ void __launch_bounds__(BS_THREADS)
basis_switch_kernel(u64* __restrict__ dst, const u64* __restrict__ src, size_t items, u32 logN, BasisDev bd,
                    u32 srcT, u32 src0, u32 dstT, u32 dst0) {
    extern __shared__ __align__(16) u64 sm[];
    u64* sH = sm;                                          // [sq][JT]
    u64* sQ = sH + (size_t)bd.sq * JT;                     // [sq]
    u64* sC = sQ + bd.sq;                                  // [sq][2] {c, companion}
    double* sI = reinterpret_cast<double*>(sC + 2 * bd.sq); // [sq] 1/q_i (EXACT only)
    const u32 j0 = blockIdx.y * JT;
    for (u32 k = threadIdx.x; k < bd.sq * JT; k += BS_THREADS) {
        u32 i = k / JT, jj = k % JT;
        sH[k] = (j0 + jj < bd.sp) ? bd.hat[(size_t)i * bd.sp + j0 + jj] : 0ull;
    }
    for (u32 i = threadIdx.x; i < bd.sq; i += BS_THREADS) {
        sQ[i] = bd.q[i];
        sC[2 * i] = bd.pre[i].x;
        sC[2 * i + 1] = bd.pre[i].y;
        if (EXACT) sI[i] = bd.qinv[i];
    }
    __syncthreads();
    const size_t e = (size_t)blockIdx.x * BS_THREADS + threadIdx.x;
    if (e >= items) return;
    const size_t b = e >> logN;
    const u32 n = (u32)e & ((1u << logN) - 1);
    u64 lo[JT], hi[JT];
#pragma unroll
    for (int jj = 0; jj < JT; jj++) { lo[jj] = 0; hi[jj] = 0; }
    const u64* sp = src + (((b * srcT) + src0) << logN) + n;
    u64 x = sp[0];
    double nu = 0.5;
    for (u32 i = 0; i < bd.sq; i++) {
        u64 xn = (i + 1 < bd.sq) ? sp[(size_t)(i + 1) << logN] : 0ull;     // prefetch the next tower's coefficient
        u64 y = shoup(x, sC[2 * i], sC[2 * i + 1], sQ[i]);
        if (EXACT) nu = __dadd_rn(nu, __dmul_rn(__ull2double_rn(y), sI[i]));
#pragma unroll
        for (int jj = 0; jj < JT; jj++) mac128(lo[jj], hi[jj], y, sH[i * JT + jj]);
        x = xn;
    }
    const u32 alpha = EXACT ? (u32)nu : 0u;
#pragma unroll
    for (int jj = 0; jj < JT; jj++) {
        if (j0 + jj < bd.sp) {
            const BasisTarget tg = bd.tgt[j0 + jj];
            u64 r = barrett128(lo[jj], hi[jj], tg.p, tg.mu_lo, tg.mu_hi);
            if (tg.post) r = shoup(r, tg.post, tg.postp, tg.p);
            if (EXACT) r = submod(r, bd.alphaq[(size_t)alpha * bd.sp + j0 + jj], tg.p);
            dst[(((b * dstT) + dst0 + j0 + jj) << logN) + n] = r;
        }
    }
}

template <bool EXACT>
static cudaError_t launch_basis_switch_t(const BasisDev& bd, u64* d, uint32_t dstT, uint32_t dst0, const u64* s, uint32_t srcT,
                                         uint32_t src0, size_t items, uint32_t logN, cudaStream_t st) {
    int jt = bd.sp >= 8 ? 8 : bd.sp >= 4 ? 4 : bd.sp >= 2 ? 2 : 1;
    dim3 grid((unsigned)((items + BS_THREADS - 1) / BS_THREADS), (bd.sp + jt - 1) / jt);
    size_t smem = ((size_t)bd.sq * jt + 4 * (size_t)bd.sq) * 8;
    switch (jt) {
        case 8: basis_switch_kernel<8, EXACT><<<grid, BS_THREADS, smem, st>>>(d, s, items, logN, bd, srcT, src0, dstT, dst0); break;
        case 4: basis_switch_kernel<4, EXACT><<<grid, BS_THREADS, smem, st>>>(d, s, items, logN, bd, srcT, src0, dstT, dst0); break;
        case 2: basis_switch_kernel<2, EXACT><<<grid, BS_THREADS, smem, st>>>(d, s, items, logN, bd, srcT, src0, dstT, dst0); break;
        default: basis_switch_kernel<1, EXACT><<<grid, BS_THREADS, smem, st>>>(d, s, items, logN, bd, srcT, src0, dstT, dst0); break;
    }
    kernel_launch_count_add(1);
    return cudaGetLastError();
}

cudaError_t launch_basis_switch(const BasisDev& bd, uint64_t* dst, uint32_t dstT, uint32_t dst0, const uint64_t* src,
                                uint32_t srcT, uint32_t src0, size_t batch, uint32_t logN, bool exact, cudaStream_t st) {
    size_t items = batch << logN;
    if (items == 0) return cudaSuccess;
    u64* d = reinterpret_cast<u64*>(dst);
    const u64* s = reinterpret_cast<const u64*>(src);
    return exact ? launch_basis_switch_t<true>(bd, d, dstT, dst0, s, srcT, src0, items, logN, st)
                 : launch_basis_switch_t<false>(bd, d, dstT, dst0, s, srcT, src0, items, logN, st);
}

// DCRTPolyImpl::ScaleAndRound(paramsOutput, ...) (dcrtpoly-impl.h:1876-1960, HAVE_INT128 branch).  Input [B][sI+sO][N];
// in0 / out0 = first input / output tower.  Per coefficient: nu = 0.5 + sum_i frac[i] * double(x_i) (IEEE double, in
// order, not fused), alpha = integer part of nu (64-bit, or 128-bit when nu >= 2^64), and for every output modulus o_j:
// (sum_i x_i * modo[j][i] + x_out_j * modo[j][sI]) mod o_j + alpha mod o_j.
template <int JT>
__global__ void __launch_bounds__(BS_THREADS)
scale_round_kernel(u64* __restrict__ dst, const u64* __restrict__ src, size_t items, u32 logN, ScaleDev sd) {
    extern __shared__ __align__(16) u64 sm[];
    u64* sM = sm;                                               // [sI+1][JT]
    double* sF = reinterpret_cast<double*>(sM + (size_t)(sd.sI + 1) * JT);   // [sI]
    const u32 j0 = blockIdx.y * JT;
    for (u32 k = threadIdx.x; k < (sd.sI + 1) * JT; k += BS_THREADS) {
        u32 i = k / JT, jj = k % JT;
        sM[k] = (j0 + jj < sd.sO) ? sd.modo[(size_t)(j0 + jj) * (sd.sI + 1) + i] : 0ull;
    }
    for (u32 i = threadIdx.x; i < sd.sI; i += BS_THREADS) sF[i] = sd.frac[i];
    __syncthreads();
    const size_t e = (size_t)blockIdx.x * BS_THREADS + threadIdx.x;
    if (e >= items) return;
    const size_t b = e >> logN;
    const u32 n = (u32)e & ((1u << logN) - 1);
    const u32 T = sd.sI + sd.sO;
    u64 lo[JT], hi[JT];
#pragma unroll
    for (int jj = 0; jj < JT; jj++) { lo[jj] = 0; hi[jj] = 0; }
    const u64* ip = src + (((b * T) + sd.in0) << logN) + n;
    double nu = 0.5;
    u64 x = ip[0];
    for (u32 i = 0; i < sd.sI; i++) {
        u64 xn = (i + 1 < sd.sI) ? ip[(size_t)(i + 1) << logN] : 0ull;
        nu = __dadd_rn(nu, __dmul_rn(sF[i], __ull2double_rn(x)));
#pragma unroll
        for (int jj = 0; jj < JT; jj++) mac128(lo[jj], hi[jj], x, sM[i * JT + jj]);
        x = xn;
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
            const BasisTarget tg = sd.tgt[j0 + jj];
            u64 xo = src[(((b * T) + sd.out0 + j0 + jj) << logN) + n];
            mac128(lo[jj], hi[jj], xo, sM[sd.sI * JT + jj]);
            u64 r = barrett128(lo[jj], hi[jj], tg.p, tg.mu_lo, tg.mu_hi);
            u64 a = barrett128(a_lo, a_hi, tg.p, tg.mu_lo, tg.mu_hi);
            dst[(((b * sd.sO) + j0 + jj) << logN) + n] = addmod(r, a, tg.p);
        }
    }
}

cudaError_t launch_scale_round(const ScaleDev& sd, uint64_t* dst, const uint64_t* src, size_t batch, uint32_t logN,
                               cudaStream_t st) {
    size_t items = batch << logN;
    if (items == 0) return cudaSuccess;
    int jt = sd.sO >= 8 ? 8 : sd.sO >= 4 ? 4 : sd.sO >= 2 ? 2 : 1;
    dim3 grid((unsigned)((items + BS_THREADS - 1) / BS_THREADS), (sd.sO + jt - 1) / jt);
    size_t smem = ((size_t)(sd.sI + 1) * jt + sd.sI) * 8;
    u64* d = reinterpret_cast<u64*>(dst);
    const u64* s = reinterpret_cast<const u64*>(src);
    switch (jt) {
        case 8: scale_round_kernel<8><<<grid, BS_THREADS, smem, st>>>(d, s, items, logN, sd); break;
        case 4: scale_round_kernel<4><<<grid, BS_THREADS, smem, st>>>(d, s, items, logN, sd); break;
        case 2: scale_round_kernel<2><<<grid, BS_THREADS, smem, st>>>(d, s, items, logN, sd); break;
        default: scale_round_kernel<1><<<grid, BS_THREADS, smem, st>>>(d, s, items, logN, sd); break;
    }
    kernel_launch_count_add(1);
    return cudaGetLastError();
}

// ---- one-pass tails -------------------------------------------------------------------------------------------
// out[b][i][n] over [B][l][N];  x read from [B][xT][N] towers 0..l-1;  v from [B][l][N];  per tower k = {c, c', t, t'}.
//   mode 0 (ApproxModDown :1173)          out = (x - v) * c
//   mode 1 (DropLastElementAndScale :761) out = x * c + v
//   mode 2 (ModReduce :809-810)           out = (x + v * t) * c
template <int MODE>
__global__ void __launch_bounds__(256)
tower_affine_kernel(u64* __restrict__ out, const u64* __restrict__ x, const u64* __restrict__ v, size_t npairs, u32 logN,
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
            if (MODE == 0) return shoup(submod(xx, yy, q), c.x, c.y, q);
            if (MODE == 1) return addmod(shoup(xx, c.x, c.y, q), yy, q);
            return shoup(addmod(xx, shoup(yy, c.z, c.w, q), q), c.x, c.y, q);
        };
        ulonglong2 r = make_ulonglong2(f(xv.x, vv.x), f(xv.y, vv.y));
        *reinterpret_cast<ulonglong2*>(out + e) = r;
    }
}

// scalar form of the same for N = 1 (no aligned pairs)
template <int MODE>
__global__ void tower_affine_scalar_kernel(u64* __restrict__ out, const u64* __restrict__ x, const u64* __restrict__ v,
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
        for (u32 i = 0; i < l; i++) {
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
    unsigned grid = (unsigned)(g > 148 * 64 ? 148 * 64 : g);
    rescale_spread_kernel<<<grid, 256, 0, st>>>(reinterpret_cast<u64*>(tmp), reinterpret_cast<const u64*>(last), items, logN, l,
                                               ql, sw, k, pre, bgv);
    kernel_launch_count_add(1);
    return cudaGetLastError();
}

}  // namespace fhe
