// behz.cu — the element-wise halves of the BEHZ variant of BFV multiplication and decryption.
//
// Reference (relative to the reference tree, src/core/include/lattice/hal/default/dcrtpoly-impl.h):
//   FastBaseConvqToBskMontgomery   :2069-2209   Q -> Q u Bsk with the small Montgomery reduction modulo mtilde = 2^16
//   FastRNSFloorq                  :2212-2271   floor(t/Q * x) in the basis Bsk
//   FastBaseConvSK                 :2309-2420   Bsk -> Q with the Shenoy-Kumaresan correction through msk
//   ScaleAndRound(moduliQ, t, tgamma, ...)  :2005-2050   decryption, gamma = 2^26
// Each of the first three is "a fast base conversion, then a per-coefficient correction".  The conversions are
// ApproxSwitchCRTBasis's own statements with other tables (a Shoup twist per source tower, 128-bit sums against a
// [source][target] table, one 128-bit Barrett reduction per target), so they run on basis.cu's kernels, tensor cores
// included; the kernels here are the corrections.  One thread owns two adjacent coefficients of one polynomial and
// loops over the towers (16-byte loads, 512 B per warp per tower); the per-tower constants sit in shared memory.
// Every value stored is a canonical residue computed by the reference's own sequence of word operations
// (ModMulFastConst = the fully reduced Shoup product, ModAddFast / ModSubFast one conditional correction, 64-bit
// wrapping sums modulo 2^16), so results are bit-identical.  All four are HBM-bound streaming kernels.
#include "modarith.cuh"
#include "kernels.h"

namespace fhe {

namespace {
constexpr int BZ_THREADS = 256;

__device__ __forceinline__ ulonglong2 ld2(const u64* p) { return *reinterpret_cast<const ulonglong2*>(p); }
__device__ __forceinline__ void st2(u64* p, ulonglong2 v) { *reinterpret_cast<ulonglong2*>(p) = v; }

// grid-stride over (polynomial, coefficient pair); returns false past the end
struct PairIndex {
    size_t b;
    u32 n;       // first coefficient of the pair
};
__device__ __forceinline__ PairIndex pair_at(size_t e, u32 logN) {
    const u32 half = logN - 1;
    return PairIndex{e >> half, (u32)((e & ((1u << half) - 1)) << 1)};
}
}  // namespace

// Steps 0 (modulo mtilde) and 1 of FastBaseConvqToBskMontgomery (:2148-2190).  coef [batch][nq][N]: the input in
// COEFFICIENT format.  x [batch][nq+nb][N]: towers nq.. hold FastBaseConv(|x*mtilde|_q, q, Bsk) on entry (written by
// the basis kernel) and the Montgomery-reduced value on exit.
__global__ void __launch_bounds__(BZ_THREADS)
behz_montgomery_kernel(u64* __restrict__ x, const u64* __restrict__ coef, size_t pairs, u32 logN, BehzDev d) {
    extern __shared__ __align__(16) unsigned char smraw[];
    ulonglong4* sQ = reinterpret_cast<ulonglong4*>(smraw);        // [nq] {q, c, c', [Q/q_i]_mtilde}
    ulonglong4* sB = sQ + d.nq;                                   // [nb] {[Q]_bsk, ', [mtilde^-1]_bsk, '}
    u64* sM = reinterpret_cast<u64*>(sB + d.nb);                  // [nb] bsk_j
    for (u32 i = threadIdx.x; i < d.nq; i += BZ_THREADS) sQ[i] = d.qk[i];
    for (u32 j = threadIdx.x; j < d.nb; j += BZ_THREADS) { sB[j] = d.bk[j]; sM[j] = d.bsk[j]; }
    __syncthreads();
    const u32 T = d.nq + d.nb;
    for (size_t e = (size_t)blockIdx.x * BZ_THREADS + threadIdx.x; e < pairs; e += (size_t)gridDim.x * BZ_THREADS) {
        const PairIndex pi = pair_at(e, logN);
        const u64* cp = coef + ((pi.b * d.nq) << logN) + pi.n;
        u64 r0 = 0, r1 = 0;
        for (u32 i = 0; i < d.nq; i++) {
            const ulonglong2 v = ld2(cp + ((size_t)i << logN));
            const ulonglong4 k = sQ[i];
            r0 += shoup(v.x, k.y, k.z, k.x) * k.w;
            r1 += shoup(v.y, k.y, k.z, k.x) * k.w;
        }
        r0 = ((r0 & 0xffffull) * d.neg_qinv_mtilde) & 0xffffull;
        r1 = ((r1 & 0xffffull) * d.neg_qinv_mtilde) & 0xffffull;
        u64* xp = x + ((pi.b * T + d.nq) << logN) + pi.n;
        for (u32 j = 0; j < d.nb; j++) {
            const ulonglong4 k = sB[j];
            const u64 m = sM[j];
            u64* p = xp + ((size_t)j << logN);
            ulonglong2 c = ld2(p);
            u64 a0 = r0 >= 0x8000ull ? r0 + m - 0x10000ull : r0;          // centred remainder (:2179-2181)
            u64 a1 = r1 >= 0x8000ull ? r1 + m - 0x10000ull : r1;
            a0 = addmod(shoup(a0, k.x, k.y, m), c.x, m);
            a1 = addmod(shoup(a1, k.x, k.y, m), c.y, m);
            c.x = shoup(a0, k.z, k.w, m);
            c.y = shoup(a1, k.z, k.w, m);
            st2(p, c);
        }
    }
}

// FastRNSFloorq after its conversion (:2233-2243, :2260-2269), in place on x [batch][nq+nb][N]: towers of Q are
// multiplied by [t*(Q/q_i)^-1]_{q_i} (only when update_q: the value the reference leaves behind, dead in the
// scheme's use), towers of Bsk become x*[t*Q^-1]_{bsk} - conv.  conv [batch][nb][N].  Flat over every pair.
__global__ void __launch_bounds__(BZ_THREADS)
behz_floor_kernel(u64* __restrict__ x, const u64* __restrict__ conv, size_t pairs, u32 logN, BehzDev d, int update_q) {
    extern __shared__ __align__(16) unsigned char smraw[];
    ulonglong4* sK = reinterpret_cast<ulonglong4*>(smraw);        // [nq+nb] {c, c', modulus, -}
    const u32 T = d.nq + d.nb;
    for (u32 i = threadIdx.x; i < T; i += BZ_THREADS) sK[i] = d.fk[i];
    __syncthreads();
    const u32 half = logN - 1;
    for (size_t e = (size_t)blockIdx.x * BZ_THREADS + threadIdx.x; e < pairs; e += (size_t)gridDim.x * BZ_THREADS) {
        const size_t unit = e >> half;                            // (polynomial, tower)
        const u32 tw = (u32)(unit % T);
        if (tw < d.nq && !update_q) continue;
        const size_t b = unit / T;
        const u32 n = (u32)((e & ((1u << half) - 1)) << 1);
        const ulonglong4 k = sK[tw];
        u64* p = x + (unit << logN) + n;
        ulonglong2 v = ld2(p);
        v.x = shoup(v.x, k.x, k.y, k.z);
        v.y = shoup(v.y, k.x, k.y, k.z);
        if (tw >= d.nq) {
            const ulonglong2 c = ld2(conv + ((b * d.nb + (tw - d.nq)) << logN) + n);
            v.x = submod(v.x, c.x, k.z);
            v.y = submod(v.y, c.y, k.z);
        }
        st2(p, v);
    }
}

// FastBaseConvSK after its conversions (:2366-2388).  conv [batch][nq+1][N] = FastBaseConv(x, B, q) in towers 0..nq-1
// and FastBaseConv(x, B, msk) in tower nq; xsk = the msk tower of the input, stride xT towers per polynomial;
// dst [batch][nq][N].
__global__ void __launch_bounds__(BZ_THREADS)
behz_sk_kernel(u64* __restrict__ dst, const u64* __restrict__ conv, const u64* __restrict__ xsk, u32 xT, size_t pairs, u32 logN,
               BehzDev d) {
    extern __shared__ __align__(16) unsigned char smraw[];
    ulonglong4* sK = reinterpret_cast<ulonglong4*>(smraw);        // [nq] {[B]_q, ', q, -}
    for (u32 i = threadIdx.x; i < d.nq; i += BZ_THREADS) sK[i] = d.sk[i];
    __syncthreads();
    const u64 msk = d.msk, half_msk = d.msk >> 1;
    for (size_t e = (size_t)blockIdx.x * BZ_THREADS + threadIdx.x; e < pairs; e += (size_t)gridDim.x * BZ_THREADS) {
        const PairIndex pi = pair_at(e, logN);
        const u64* cp = conv + ((pi.b * (d.nq + 1)) << logN) + pi.n;
        const ulonglong2 s = ld2(xsk + ((pi.b * xT) << logN) + pi.n);
        ulonglong2 a = ld2(cp + ((size_t)d.nq << logN));
        a.x = shoup(submod(a.x, s.x, msk), d.binv_msk, d.binv_mskp, msk);           // alpha_sk (:2366-2371)
        a.y = shoup(submod(a.y, s.y, msk), d.binv_msk, d.binv_mskp, msk);
        u64* op = dst + ((pi.b * d.nq) << logN) + pi.n;
        for (u32 i = 0; i < d.nq; i++) {
            const ulonglong4 k = sK[i];
            ulonglong2 v = ld2(cp + ((size_t)i << logN));
            // ModSubFast(msk, q_i) of a value below msk is a + q_i - msk in 64-bit words, whatever the sizes (:2381-2382)
            const u64 t0 = a.x > half_msk ? a.x + k.z - msk : a.x;
            const u64 t1 = a.y > half_msk ? a.y + k.z - msk : a.y;
            v.x = submod(v.x, shoup(t0, k.x, k.y, k.z), k.z);
            v.y = submod(v.y, shoup(t1, k.x, k.y, k.z), k.z);
            st2(op + ((size_t)i << logN), v);
        }
    }
}

// ScaleAndRound(moduliQ, t, tgamma, ...) -> NativePoly (:2005-2050): src [batch][nq][N] COEFFICIENT -> dst [batch][N]
__global__ void __launch_bounds__(BZ_THREADS)
behz_decrypt_kernel(u64* __restrict__ dst, const u64* __restrict__ src, size_t pairs, u32 logN, BehzDev d) {
    extern __shared__ __align__(16) unsigned char smraw[];
    ulonglong4* sK = reinterpret_cast<ulonglong4*>(smraw);        // [nq] {[t*gamma*(Q/q_i)^-1]_q, ', [-q_i^-1]_{t*gamma}, '}
    u64* sM = reinterpret_cast<u64*>(sK + d.nq);                  // [nq] q_i
    for (u32 i = threadIdx.x; i < d.nq; i += BZ_THREADS) { sK[i] = d.dk[i]; sM[i] = d.qk[i].x; }
    __syncthreads();
    const u64 tg = d.tgamma;
    for (size_t e = (size_t)blockIdx.x * BZ_THREADS + threadIdx.x; e < pairs; e += (size_t)gridDim.x * BZ_THREADS) {
        const PairIndex pi = pair_at(e, logN);
        const u64* sp = src + ((pi.b * d.nq) << logN) + pi.n;
        u64 s0 = 0, s1 = 0;
        for (u32 i = 0; i < d.nq; i++) {
            const ulonglong2 v = ld2(sp + ((size_t)i << logN));
            const ulonglong4 k = sK[i];
            const u64 q = sM[i];
            s0 = addmod(s0, shoup(shoup(v.x, k.x, k.y, q), k.z, k.w, tg), tg);
            s1 = addmod(s1, shoup(shoup(v.y, k.x, k.y, q), k.z, k.w, tg), tg);
        }
        s0 += s0 & 0x3ffffffull;                                  // s + (s mod gamma), then the shift by log2(gamma)
        s1 += s1 & 0x3ffffffull;
        st2(dst + (pi.b << logN) + pi.n, make_ulonglong2(s0 >> 26, s1 >> 26));
    }
}

static unsigned bz_grid(size_t pairs) {
    size_t blocks = (pairs + BZ_THREADS - 1) / BZ_THREADS;
    const size_t cap = 148 * 16;                                  // 148 SMs x resident CTAs, grid-stride beyond
    return (unsigned)(blocks < cap ? blocks : cap);
}

cudaError_t launch_behz_montgomery(const BehzDev& d, uint64_t* x, const uint64_t* coef, size_t batch, uint32_t logN, cudaStream_t st) {
    if (batch == 0) return cudaSuccess;
    if (logN < 1) return cudaErrorInvalidValue;
    const size_t pairs = batch << (logN - 1);
    const size_t smem = (size_t)(d.nq + d.nb) * sizeof(ulonglong4) + (size_t)d.nb * 8;
    if (smem > 48 * 1024) return cudaErrorInvalidValue;
    behz_montgomery_kernel<<<bz_grid(pairs), BZ_THREADS, smem, st>>>(reinterpret_cast<u64*>(x), reinterpret_cast<const u64*>(coef),
                                                                    pairs, logN, d);
    kernel_launch_count_add(1);
    return cudaGetLastError();
}

cudaError_t launch_behz_floor(const BehzDev& d, uint64_t* x, const uint64_t* conv, size_t batch, uint32_t logN, bool update_q,
                              cudaStream_t st) {
    if (batch == 0) return cudaSuccess;
    if (logN < 1) return cudaErrorInvalidValue;
    const size_t pairs = (batch * (d.nq + d.nb)) << (logN - 1);
    const size_t smem = (size_t)(d.nq + d.nb) * sizeof(ulonglong4);
    if (smem > 48 * 1024) return cudaErrorInvalidValue;
    behz_floor_kernel<<<bz_grid(pairs), BZ_THREADS, smem, st>>>(reinterpret_cast<u64*>(x), reinterpret_cast<const u64*>(conv), pairs,
                                                               logN, d, update_q ? 1 : 0);
    kernel_launch_count_add(1);
    return cudaGetLastError();
}

cudaError_t launch_behz_sk(const BehzDev& d, uint64_t* dst, const uint64_t* conv, const uint64_t* xsk, uint32_t xT, size_t batch,
                           uint32_t logN, cudaStream_t st) {
    if (batch == 0) return cudaSuccess;
    if (logN < 1) return cudaErrorInvalidValue;
    const size_t pairs = batch << (logN - 1);
    const size_t smem = (size_t)d.nq * sizeof(ulonglong4);
    if (smem > 48 * 1024) return cudaErrorInvalidValue;
    behz_sk_kernel<<<bz_grid(pairs), BZ_THREADS, smem, st>>>(reinterpret_cast<u64*>(dst), reinterpret_cast<const u64*>(conv),
                                                            reinterpret_cast<const u64*>(xsk), xT, pairs, logN, d);
    kernel_launch_count_add(1);
    return cudaGetLastError();
}

cudaError_t launch_behz_decrypt(const BehzDev& d, uint64_t* dst, const uint64_t* src, size_t batch, uint32_t logN, cudaStream_t st) {
    if (batch == 0) return cudaSuccess;
    if (logN < 1) return cudaErrorInvalidValue;
    const size_t pairs = batch << (logN - 1);
    const size_t smem = (size_t)d.nq * (sizeof(ulonglong4) + 8);
    if (smem > 48 * 1024) return cudaErrorInvalidValue;
    behz_decrypt_kernel<<<bz_grid(pairs), BZ_THREADS, smem, st>>>(reinterpret_cast<u64*>(dst), reinterpret_cast<const u64*>(src), pairs,
                                                                 logN, d);
    kernel_launch_count_add(1);
    return cudaGetLastError();
}

}  // namespace fhe
