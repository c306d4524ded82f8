// ntt.cu — batched negacyclic NTT / INTT over RNS towers for sm_100a.
//
// Replaces, on the device, the reference's CPU kernels (citations relative to the reference tree):
//   NumberTheoreticTransformNat::ForwardTransformToBitReverseInPlace   transformnat-impl.h:300-354
//   NumberTheoreticTransformNat::InverseTransformFromBitReverseInPlace transformnat-impl.h:492-552
// with the same index pattern (Cooley-Tukey natural->bit-reversed with twiddle table[m+i];
// Gentleman-Sande bit-reversed->natural with N^-1 folded in), so outputs are bit-identical.
//
// Data layout in HBM: [batch][tower][N] u64, tower-major and coefficient-contiguous like the
// reference's DCRTPoly.  Twiddles: per tower N records {w, floor(w*2^64/q)} (16 B, one LDG.128).
//
// Decomposition for N = 2^L, 9 <= L <= 16: N = N1 x 256.
//   strided pass: N1-point sub-NTTs over elements 256 apart (stages 0..L-9), all columns and all
//                 batch items of a tower share the same N1-1 twiddles;
//   contig pass : 256-point sub-NTTs over contiguous blocks (stages L-8..L-1); a tile takes the
//                 SAME block of up to 16 different batch items so twiddle loads are shared.
// Each thread keeps 16 coefficients in registers and runs 4 butterfly stages per step (radix-16),
// one shared-memory exchange per pass.  Lazy reduction with compile-time range tracking: values stay below 16q
// (q < 2^60) and the butterfly multiplier returns [0,4q) (quotient estimated from three of the four partial
// products, modarith.cuh); the last store of a transform reduces to [0,q), so outputs equal the reference's.
// HBM traffic: 32 B per coefficient per transform (two passes); twiddles amortised over the batch.
// N <= 256 (and any N <= 4096 on request) uses one CTA per tower entirely in shared memory.
#include "ntt_steps.cuh"
#include "kernels.h"
#include <atomic>

#include <algorithm>
#include <cstdlib>

namespace fhe {

constexpr long NTT_L2_CHUNK_MB_DEFAULT = 0;

// ------------------------------------------------------------------------------------------
// Strided pass: R = 2^RL point sub-NTTs over elements 256 apart.  Tile = R rows x C columns.
// Forward: stages 0..RL-1 of the transform (twiddle prefix 1).  Inverse: the same stages in
// reverse order, run AFTER the contig pass, ending with the N^-1 stage.
// ------------------------------------------------------------------------------------------
// TL = log2(coefficients per tile): 12 -> 256 threads, 11 -> 128, 10 -> 64, 9 -> one warp.  Smaller tiles mean more,
// smaller CTAs per SM whose load / compute / store phases interleave.
template <int RL, int TL>
struct StridedCfg {
    static constexpr int A = RL < 4 ? RL : 4;          // stages done on the top bits of the row index
    static constexpr int B = RL - A;                   // stages done on the low bits
    static constexpr int R = 1 << RL;
    static constexpr int C = ((1 << TL) / R) < 256 ? ((1 << TL) / R) : 256;
    static constexpr int THREADS = R * C / 16;
    static constexpr int GA = 16 >> A;                 // register groups per thread in step A
    static constexpr int GB = 16 >> B;
    // exchange buffer [row][C]; C < 16: one extra row of padding per 2^B rows keeps the c-major reads
    // of two neighbouring sub-NTTs in different banks
    static constexpr int PAD = (B > 0 && C < 16) ? C : 0;
    static constexpr int SMEM = (B > 0) ? (R * C + (R >> B) * PAD) * 8 : 0;
    static_assert(THREADS >= 32 || RL < 4, "tile too small");
};

template <int RL, int TL, bool INV>
__global__ void __launch_bounds__(StridedCfg<RL, TL>::THREADS)
ntt_strided_kernel(u64* __restrict__ data, const u64* __restrict__ in, const ulonglong2* __restrict__ twAll,
                   const TowerMod* __restrict__ mods, const ulonglong2* __restrict__ invc,
                   u32 T, u32 logN, u32 outT, u32 inT) {
    using K = StridedCfg<RL, TL>;
    auto sidx = [](u32 row, u32 c) -> u32 { return row * K::C + c + (row >> K::B) * K::PAD; };
    extern __shared__ __align__(16) u64 sm[];
    constexpr int A = K::A, B = K::B, C = K::C;
    constexpr u32 tilesPerUnit = 256 / C;
    const u32 tower = blockIdx.x / tilesPerUnit;      // grid: x = (tower, column group), y/z = batch item
    const u32 cg    = blockIdx.x % tilesPerUnit;
    const u64 q = mods[tower].q, nq = 0ull - q;
    const QMul Q = make_qmul(q);
    const ulonglong2* tw = twAll + ((size_t)tower << logN);
    // towers of one batch item sit outT (inT) towers apart in memory: a view of a wider poly transforms in place,
    // and the first pass of a transform may read another buffer (in) than the one it leaves its result in (data)
    const u32 bitem = blockIdx.z * gridDim.y + blockIdx.y;
    u64* base = data + (((size_t)bitem * outT + tower) << logN) + cg * C;
#ifdef FHE_NTT_INPLACE_ONLY
    const u64* ibase = base;
#else
    const u64* ibase = in + (((size_t)bitem * inT + tower) << logN) + cg * C;
#endif
    const u32 tid = threadIdx.x;

    if (!INV) {
        // step A: top A bits of the row index live in registers
#pragma unroll
        for (int g = 0; g < K::GA; g++) {
            u32 item = tid + g * K::THREADS;
            u32 c = item % C, ulow = item / C;
            u64 x[1 << A];
#pragma unroll
            for (int e = 0; e < (1 << A); e++) x[e] = ibase[((size_t)((e << B) + ulow) << 8) + c];
            fwd_step<A, 1>(x, tw, 1u, Q, nq);
            if (B == 0) {
#pragma unroll
                for (int e = 0; e < (1 << A); e++) base[((size_t)((e << B) + ulow) << 8) + c] = x[e];
            } else {
#pragma unroll
                for (int e = 0; e < (1 << A); e++) sm[sidx((e << B) + ulow, c)] = x[e];
            }
        }
        if (B > 0) {
            __syncthreads();
#pragma unroll
            for (int g = 0; g < K::GB; g++) {
                u32 item = tid + g * K::THREADS;
                u32 c = item % C, uhi = item / C;
                u64 x[1 << B];
#pragma unroll
                for (int e = 0; e < (1 << B); e++) x[e] = sm[sidx((uhi << B) + e, c)];
                fwd_step<B, FwdStep<A, 1>::OUT>(x, tw, (1u << A) + uhi, Q, nq);
#pragma unroll
                for (int e = 0; e < (1 << B); e++) base[((size_t)((uhi << B) + e) << 8) + c] = x[e];
            }
        }
    } else {
        const ulonglong2 sc = invc[2 * tower], w1 = invc[2 * tower + 1];
        if (B > 0) {
#pragma unroll
            for (int g = 0; g < K::GB; g++) {
                u32 item = tid + g * K::THREADS;
                u32 c = item % C, uhi = item / C;
                u64 x[1 << B];
#pragma unroll
                for (int e = 0; e < (1 << B); e++) x[e] = base[((size_t)((uhi << B) + e) << 8) + c];
                inv_step<B, INV_NORM, INV_NORM>(x, tw, (1u << A) + uhi, Q, nq);
#pragma unroll
                for (int e = 0; e < (1 << B); e++) sm[sidx((uhi << B) + e, c)] = x[e];
            }
            __syncthreads();
        }
#pragma unroll
        for (int g = 0; g < K::GA; g++) {
            u32 item = tid + g * K::THREADS;
            u32 c = item % C, ulow = item / C;
            u64 x[1 << A];
            if (B > 0) {
#pragma unroll
                for (int e = 0; e < (1 << A); e++) x[e] = sm[sidx((e << B) + ulow, c)];
            } else {
#pragma unroll
                for (int e = 0; e < (1 << A); e++) x[e] = base[((size_t)((e << B) + ulow) << 8) + c];
            }
            inv_step_last<A, INV_NORM>(x, tw, sc, w1, Q, nq);
#pragma unroll
            for (int e = 0; e < (1 << A); e++) base[((size_t)((e << B) + ulow) << 8) + c] = x[e];
        }
    }
}

// ------------------------------------------------------------------------------------------
// The strided pass as a persistent kernel with register prefetch (round 2).  ncu's stall sampling of the one-tile kernel
// above shows its idle issue cycles are mostly warps waiting for the tile's first loads (long scoreboard on the first
// butterflies of the tile).  Here a CTA walks over tiles and issues the loads of its NEXT tile into the registers that the
// current tile has just left (its values are in shared memory by then), so they fly while the second half of the current
// tile is computed.  The exchange buffer is double-buffered: one barrier per tile.  Same arithmetic, same tile shape.
// Needs A = 4 (RL >= 4), i.e. one register group per thread in the first step.
// ------------------------------------------------------------------------------------------
template <int RL, int TL, bool INV>
__global__ void __launch_bounds__(StridedCfg<RL, TL>::THREADS)
ntt_strided_pf_kernel(u64* __restrict__ data, const u64* __restrict__ in, const ulonglong2* __restrict__ twAll,
                      const TowerMod* __restrict__ mods, const ulonglong2* __restrict__ invc,
                      u32 T, u32 logN, u32 outT, u32 inT, u32 ntiles) {
    using K = StridedCfg<RL, TL>;
    static_assert(K::A == 4 && K::B >= 1 && K::GA == 1, "prefetch form: 16 coefficients per thread in the first step");
    auto sidx = [](u32 row, u32 c) -> u32 { return row * K::C + c + (row >> K::B) * K::PAD; };
    extern __shared__ __align__(16) u64 small[];
    constexpr int A = K::A, B = K::B, C = K::C;
    constexpr u32 tilesPerUnit = 256 / C;
    constexpr u32 SMW = K::SMEM / 8;
    const u32 tid = threadIdx.x;
    // the thread's place in the tile, for the step that touches global memory first (forward: A; inverse: B)
    const u32 cA = tid % C, ulow = tid / C;

    struct Tile { u32 tower; size_t ioff, ooff; };
    auto decode = [&](u32 L) -> Tile {
        const u32 cg = L % tilesPerUnit, rest = L / tilesPerUnit;
        const u32 tower = rest % T, bitem = rest / T;
        Tile t;
        t.tower = tower;
        t.ooff = (((size_t)bitem * outT + tower) << logN) + cg * C;
        t.ioff = (((size_t)bitem * inT + tower) << logN) + cg * C;
        return t;
    };
    // the 16 coefficients this thread reads from global memory for a tile
    auto load = [&](u64 (&x)[16], const Tile& t) {
        const u64* ibase = in + t.ioff;
        if (!INV) {
#pragma unroll
            for (int e = 0; e < 16; e++) x[e] = ibase[((size_t)((e << B) + ulow) << 8) + cA];
        } else {
#pragma unroll
            for (int g = 0; g < K::GB; g++) {
                const u32 item = tid + g * K::THREADS, c = item % C, uhi = item / C;
#pragma unroll
                for (int e = 0; e < (1 << B); e++) x[g * (1 << B) + e] = ibase[((size_t)((uhi << B) + e) << 8) + c];
            }
        }
    };

    u32 L = blockIdx.x;
    if (L >= ntiles) return;
    Tile cur = decode(L);
    u64 x[16];
    load(x, cur);
    for (u32 it = 0;; it++) {
        u64* sm = small + (it & 1u) * SMW;
        const u32 Ln = L + gridDim.x;
        const bool more = Ln < ntiles;
        Tile nxt = cur;
        if (more) nxt = decode(Ln);
        const u64 q = mods[cur.tower].q, nq = 0ull - q;
        const QMul Q = make_qmul(q);
        const ulonglong2* tw = twAll + ((size_t)cur.tower << logN);
        u64* base = data + cur.ooff;
        if (!INV) {
            fwd_step<A, 1>(x, tw, 1u, Q, nq);
#pragma unroll
            for (int e = 0; e < 16; e++) sm[sidx((e << B) + ulow, cA)] = x[e];
            if (more) load(x, nxt);                 // x is free: the next tile's loads fly during the second step
            __syncthreads();
#pragma unroll
            for (int g = 0; g < K::GB; g++) {
                const u32 item = tid + g * K::THREADS, c = item % C, uhi = item / C;
                u64 y[1 << B];
#pragma unroll
                for (int e = 0; e < (1 << B); e++) y[e] = sm[sidx((uhi << B) + e, c)];
                fwd_step<B, FwdStep<A, 1>::OUT>(y, tw, (1u << A) + uhi, Q, nq);
#pragma unroll
                for (int e = 0; e < (1 << B); e++) base[((size_t)((uhi << B) + e) << 8) + c] = y[e];
            }
        } else {
            const ulonglong2 sc = invc[2 * cur.tower], w1 = invc[2 * cur.tower + 1];
#pragma unroll
            for (int g = 0; g < K::GB; g++) {
                const u32 item = tid + g * K::THREADS, c = item % C, uhi = item / C;
                u64 y[1 << B];
#pragma unroll
                for (int e = 0; e < (1 << B); e++) y[e] = x[g * (1 << B) + e];
                inv_step<B, INV_NORM, INV_NORM>(y, tw, (1u << A) + uhi, Q, nq);
#pragma unroll
                for (int e = 0; e < (1 << B); e++) sm[sidx((uhi << B) + e, c)] = y[e];
            }
            if (more) load(x, nxt);
            __syncthreads();
            u64 y[16];
#pragma unroll
            for (int e = 0; e < 16; e++) y[e] = sm[sidx((e << B) + ulow, cA)];
            inv_step_last<A, INV_NORM>(y, tw, sc, w1, Q, nq);
#pragma unroll
            for (int e = 0; e < 16; e++) base[((size_t)((e << B) + ulow) << 8) + cA] = y[e];
        }
        if (!more) break;
        L = Ln;
        cur = nxt;
    }
}

// ------------------------------------------------------------------------------------------
// Contig pass: 256-point sub-NTTs over contiguous blocks (transform stages L-8..L-1).
// Tile = NC = 2^LB sub-NTTs (16, 8 or 4): CB batch items x CH consecutive blocks (CB*CH = NC),
// 16*NC threads.  Shared layout [u][c] with the column XOR-swizzled by bits of u (and, for NC < 16, NC
// words of padding per 16 rows): the u-major, c-major and transposed accesses are all conflict-free.
// ------------------------------------------------------------------------------------------
template <int LB>
struct ContigCfg {
    static constexpr int NC = 1 << LB, THREADS = 16 * NC;
    static constexpr int SMEM = (256 * NC + (LB < 4 ? 16 * NC : 0)) * 8;
    __device__ static __forceinline__ u32 idx(u32 u, u32 c) {
        u32 a = u * NC + (c ^ ((u >> (4 - LB)) & (NC - 1)));
        if (LB < 4) a += (u >> 4) * NC;
        return a;
    }
};

template <int LB, bool INV>
__global__ void __launch_bounds__(ContigCfg<LB>::THREADS)
ntt_contig_kernel(u64* __restrict__ data, const u64* __restrict__ in, const ulonglong2* __restrict__ twAll,
                  const TowerMod* __restrict__ mods, u32 T, u32 logN, u32 batch, u32 logCB, u32 outT, u32 inT) {
    using K = ContigCfg<LB>;
    constexpr u32 NC = K::NC, TH = K::THREADS;
    extern __shared__ __align__(16) u64 sm[];
    const u32 N1 = 1u << (logN - 8);
    const u32 CB = 1u << logCB, CH = NC >> logCB;
    const u32 hgroups = N1 / CH;                      // N1 >= CH is guaranteed by the launcher
    const u32 bgroups = (batch + CB - 1) / CB;
    u32 bid = blockIdx.x;
    const u32 hg = bid % hgroups; bid /= hgroups;
    const u32 bg = bid % bgroups;
    const u32 tower = bid / bgroups;
    const TowerMod tm = mods[tower];
    const u64 q = tm.q, nq = 0ull - q;
    const QMul Q = make_qmul(q);
    const ulonglong2* tw = twAll + ((size_t)tower << logN);
    const u32 tid = threadIdx.x;
    u64* kq = sm + K::SMEM / 8;                        // {0, q, .., 15q} for the closing range reduction
    if (!INV && tid < 16) kq[tid] = tid * q;           // visible after the first barrier

    auto blk_ptr = [&](u32 c, bool& ok, u32& h) -> u64* {
        u32 b = bg * CB + (c & (CB - 1));
        h = hg * CH + (c >> logCB);
        ok = b < batch;
        return data + (((size_t)b * outT + tower) << logN) + ((size_t)h << 8);
    };
    auto in_ptr = [&](u32 c) -> const u64* {         // the same block in the buffer the inverse transform reads
        u32 b = bg * CB + (c & (CB - 1));
#ifdef FHE_NTT_INPLACE_ONLY
        return data + (((size_t)b * outT + tower) << logN) + ((size_t)(hg * CH + (c >> logCB)) << 8);
#endif
        return in + (((size_t)b * inT + tower) << logN) + ((size_t)(hg * CH + (c >> logCB)) << 8);
    };

    if (!INV) {
        {   // step A: lanes run along the low 4 bits of u (coalesced), registers hold the top 4
            u32 ulow = tid & 15, c = tid >> 4, h; bool ok;
            u64* p = blk_ptr(c, ok, h);
            u64 x[16];
#pragma unroll
            for (int e = 0; e < 16; e++) x[e] = ok ? p[e * 16 + ulow] : 0ull;
            fwd_step<4, MAXB>(x, tw, N1 + h, Q, nq);          // entry: whatever the strided pass left, < 16q
#pragma unroll
            for (int e = 0; e < 16; e++) sm[K::idx(e * 16 + ulow, c)] = x[e];
        }
        __syncthreads();
        {   // step B: lanes run along c, registers hold the low 4 bits of u
            u32 c = tid & (NC - 1), uhi = tid >> LB, h; bool ok;
            (void)blk_ptr(c, ok, h);
            u64 x[16];
#pragma unroll
            for (int e = 0; e < 16; e++) x[e] = sm[K::idx(uhi * 16 + e, c)];
            fwd_step<4, FwdStep<4, MAXB>::OUT>(x, tw, ((N1 + h) << 4) + uhi, Q, nq);
#pragma unroll
            for (int e = 0; e < 16; e++)
                sm[K::idx(uhi * 16 + e, c)] = reduce16(x[e], kq, tm.rsh, tm.rinv, q);
        }
        __syncthreads();
#pragma unroll
        for (u32 k = 0; k < NC; k++) {               // coalesced write-back, fully reduced
            u32 h; bool ok;
            u64* p = blk_ptr(k, ok, h);
            if (ok) {
#pragma unroll
                for (u32 u = tid; u < 256; u += TH) p[u] = sm[K::idx(u, k)];
            }
        }
    } else {
#pragma unroll
        for (u32 k = 0; k < NC; k++) {
            u32 h; bool ok;
            (void)blk_ptr(k, ok, h);
            const u64* p = in_ptr(k);
#pragma unroll
            for (u32 u = tid; u < 256; u += TH) sm[K::idx(u, k)] = ok ? p[u] : 0ull;
        }
        __syncthreads();
        {
            u32 c = tid & (NC - 1), uhi = tid >> LB, h; bool ok;
            (void)blk_ptr(c, ok, h);
            u64 x[16];
#pragma unroll
            for (int e = 0; e < 16; e++) x[e] = sm[K::idx(uhi * 16 + e, c)];
            inv_step<4, 1, INV_NORM>(x, tw, ((N1 + h) << 4) + uhi, Q, nq);
#pragma unroll
            for (int e = 0; e < 16; e++) sm[K::idx(uhi * 16 + e, c)] = x[e];
        }
        __syncthreads();
        {
            u32 ulow = tid & 15, c = tid >> 4, h; bool ok;
            u64* p = blk_ptr(c, ok, h);
            u64 x[16];
#pragma unroll
            for (int e = 0; e < 16; e++) x[e] = sm[K::idx(e * 16 + ulow, c)];
            inv_step<4, INV_NORM, INV_NORM>(x, tw, N1 + h, Q, nq);
            if (ok) {
#pragma unroll
                for (int e = 0; e < 16; e++) p[e * 16 + ulow] = x[e];   // lazily reduced, [0, INV_NORM*q)
            }
        }
    }
}

// ------------------------------------------------------------------------------------------
// Both passes in ONE kernel for N = 2^(RL+8) <= 2^14 (round 2): one CTA of N/16 threads per (batch item, tower), the
// whole tower in shared memory (N x 8 B: 128 KB at N = 2^14), so a transform reads and writes HBM once (16 B per
// coefficient instead of 32) and is one launch instead of two.  Same register steps, same twiddle prefixes and same
// bound schedule as the two kernels above; the difference is where the values rest between steps: in place, at
//     L(h, u) = h*256 + (u & 0xF0) + ((u & 15) ^ (u >> 4))          h = 256-block (row of the strided pass), u = offset
// which is conflict-free for the three access patterns (lanes along u for a fixed row: strided steps and the copies;
// lanes along the low 4 bits of u; lanes along the high 4 bits of u), so no second buffer is needed.
// ------------------------------------------------------------------------------------------
template <int RL>
struct FusedCfg {
    static constexpr int A = RL < 4 ? RL : 4, B = RL - A;
    static constexpr int N1 = 1 << RL, THREADS = 16 * N1, N = 256 * N1;
    static constexpr int GA = 16 >> A, GB = 16 >> B;
    static constexpr int SMEM = (N + 16) * 8;
    __device__ static __forceinline__ u32 at(u32 h, u32 u) { return (h << 8) | (u & 0xF0u) | ((u & 15u) ^ (u >> 4)); }
};

template <int RL, bool INV>
__global__ void __launch_bounds__(FusedCfg<RL>::THREADS, 1024 / FusedCfg<RL>::THREADS)      // 1024 threads per SM: 64 registers
ntt_fused_kernel(u64* __restrict__ data, const u64* __restrict__ in, const ulonglong2* __restrict__ twAll,
                 const TowerMod* __restrict__ mods, const ulonglong2* __restrict__ invc, u32 T, u32 outT, u32 inT) {
    using K = FusedCfg<RL>;
    constexpr int A = K::A, B = K::B;
    constexpr u32 N1 = K::N1, TH = K::THREADS, logN = RL + 8;
    extern __shared__ __align__(16) u64 sm[];
    const u32 unit = blockIdx.x, tower = unit % T, bitem = unit / T;
    const TowerMod tm = mods[tower];
    const u64 q = tm.q, nq = 0ull - q;
    const QMul Q = make_qmul(q);
    const ulonglong2* tw = twAll + ((size_t)tower << logN);
    u64* out = data + (((size_t)bitem * outT + tower) << logN);
    const u64* src = in + (((size_t)bitem * inT + tower) << logN);
    const u32 tid = threadIdx.x;
    u64* kq = sm + K::N;                               // {0, q, .., 15q} for the closing range reduction
    if (!INV && tid < 16) kq[tid] = tid * q;           // visible after the first barrier

    if (!INV) {
        // strided step A: the top A bits of the block index live in registers; read from HBM, coalesced along u
#pragma unroll
        for (int g = 0; g < K::GA; g++) {
            const u32 item = tid + g * TH, c = item & 255u, ulow = item >> 8;
            u64 x[1 << A];
#pragma unroll
            for (int e = 0; e < (1 << A); e++) x[e] = src[((size_t)((e << B) + ulow) << 8) + c];
            fwd_step<A, 1>(x, tw, 1u, Q, nq);
#pragma unroll
            for (int e = 0; e < (1 << A); e++) sm[K::at((e << B) + ulow, c)] = x[e];
        }
        __syncthreads();
        if (B > 0) {
#pragma unroll
            for (int g = 0; g < K::GB; g++) {
                const u32 item = tid + g * TH, c = item & 255u, uhi = item >> 8;
                u64 x[1 << B];
#pragma unroll
                for (int e = 0; e < (1 << B); e++) x[e] = sm[K::at((uhi << B) + e, c)];
                fwd_step<B, FwdStep<A, 1>::OUT>(x, tw, (1u << A) + uhi, Q, nq);
#pragma unroll
                for (int e = 0; e < (1 << B); e++) sm[K::at((uhi << B) + e, c)] = x[e];
            }
            __syncthreads();
        }
        const u32 lo4 = tid & 15u, h = tid >> 4;
        {   // contig step A: registers hold the top 4 bits of u
            u64 x[16];
#pragma unroll
            for (int e = 0; e < 16; e++) x[e] = sm[K::at(h, e * 16 + lo4)];
            fwd_step<4, MAXB>(x, tw, N1 + h, Q, nq);
#pragma unroll
            for (int e = 0; e < 16; e++) sm[K::at(h, e * 16 + lo4)] = x[e];
        }
        __syncthreads();
        {   // contig step B: registers hold the low 4 bits of u; fully reduced on the way out
            u64 x[16];
#pragma unroll
            for (int e = 0; e < 16; e++) x[e] = sm[K::at(h, lo4 * 16 + e)];
            fwd_step<4, FwdStep<4, MAXB>::OUT>(x, tw, ((N1 + h) << 4) + lo4, Q, nq);
#pragma unroll
            for (int e = 0; e < 16; e++) sm[K::at(h, lo4 * 16 + e)] = reduce16(x[e], kq, tm.rsh, tm.rinv, q);
        }
        __syncthreads();
#pragma unroll 4
        for (u32 i = tid; i < (u32)K::N; i += TH) out[i] = sm[K::at(i >> 8, i & 255u)];
    } else {
        const ulonglong2 sc = invc[2 * tower], w1 = invc[2 * tower + 1];
#pragma unroll 4
        for (u32 i = tid; i < (u32)K::N; i += TH) sm[K::at(i >> 8, i & 255u)] = src[i];
        __syncthreads();
        const u32 lo4 = tid & 15u, h = tid >> 4;
        {
            u64 x[16];
#pragma unroll
            for (int e = 0; e < 16; e++) x[e] = sm[K::at(h, lo4 * 16 + e)];
            inv_step<4, 1, INV_NORM>(x, tw, ((N1 + h) << 4) + lo4, Q, nq);
#pragma unroll
            for (int e = 0; e < 16; e++) sm[K::at(h, lo4 * 16 + e)] = x[e];
        }
        __syncthreads();
        {
            u64 x[16];
#pragma unroll
            for (int e = 0; e < 16; e++) x[e] = sm[K::at(h, e * 16 + lo4)];
            inv_step<4, INV_NORM, INV_NORM>(x, tw, N1 + h, Q, nq);
#pragma unroll
            for (int e = 0; e < 16; e++) sm[K::at(h, e * 16 + lo4)] = x[e];
        }
        __syncthreads();
        if (B > 0) {
#pragma unroll
            for (int g = 0; g < K::GB; g++) {
                const u32 item = tid + g * TH, c = item & 255u, uhi = item >> 8;
                u64 x[1 << B];
#pragma unroll
                for (int e = 0; e < (1 << B); e++) x[e] = sm[K::at((uhi << B) + e, c)];
                inv_step<B, INV_NORM, INV_NORM>(x, tw, (1u << A) + uhi, Q, nq);
#pragma unroll
                for (int e = 0; e < (1 << B); e++) sm[K::at((uhi << B) + e, c)] = x[e];
            }
            __syncthreads();
        }
#pragma unroll
        for (int g = 0; g < K::GA; g++) {
            const u32 item = tid + g * TH, c = item & 255u, ulow = item >> 8;
            u64 x[1 << A];
#pragma unroll
            for (int e = 0; e < (1 << A); e++) x[e] = sm[K::at((e << B) + ulow, c)];
            inv_step_last<A, INV_NORM>(x, tw, sc, w1, Q, nq);
#pragma unroll
            for (int e = 0; e < (1 << A); e++) out[((size_t)((e << B) + ulow) << 8) + c] = x[e];
        }
    }
}

// ------------------------------------------------------------------------------------------
// Small rings (N <= 4096): one CTA per (batch, tower) unit, whole tower in shared memory,
// radix-2 stages with fully reduced butterflies exactly as the reference orders them.
// ------------------------------------------------------------------------------------------
template <bool INV>
__global__ void __launch_bounds__(256)
ntt_small_kernel(u64* __restrict__ data, const u64* __restrict__ in, const ulonglong2* __restrict__ twAll,
                 const TowerMod* __restrict__ mods, const ulonglong2* __restrict__ invc,
                 u32 T, u32 logN, u32 outT, u32 inT) {
    extern __shared__ __align__(16) u64 sm[];
    const u32 N = 1u << logN;
    const u32 unit = blockIdx.x, tower = unit % T;
    const u64 q = mods[tower].q;
    const ulonglong2* tw = twAll + ((size_t)tower << logN);
    u64* p = data + (((size_t)(unit / T) * outT + tower) << logN);
    const u64* pi = in + (((size_t)(unit / T) * inT + tower) << logN);
    for (u32 i = threadIdx.x; i < N; i += blockDim.x) sm[i] = pi[i];
    __syncthreads();
    if (!INV) {
        for (u32 s = 0; s < logN; s++) {              // m = 2^s, t = N >> (s+1)
            u32 lt = logN - 1 - s;
            for (u32 k = threadIdx.x; k < N / 2; k += blockDim.x) {
                u32 i = k >> lt, j = ((i << (lt + 1)) | (k & ((1u << lt) - 1)));
                ulonglong2 w = ldtw(tw + (1u << s) + i);
                u64 lo = sm[j], of = shoup(sm[j + (1u << lt)], w.x, w.y, q);
                sm[j] = addmod(lo, of, q);
                sm[j + (1u << lt)] = submod(lo, of, q);
            }
            __syncthreads();
        }
    } else {
        const ulonglong2 sc = invc[2 * tower];
        for (int s = (int)logN - 1; s >= 0; s--) {
            u32 lt = logN - 1 - s;
            for (u32 k = threadIdx.x; k < N / 2; k += blockDim.x) {
                u32 i = k >> lt, j = ((i << (lt + 1)) | (k & ((1u << lt) - 1)));
                ulonglong2 w = ldtw(tw + (1u << s) + i);
                u64 lo = sm[j], hi = sm[j + (1u << lt)];
                sm[j] = addmod(lo, hi, q);
                sm[j + (1u << lt)] = shoup(submod(lo, hi, q), w.x, w.y, q);
            }
            __syncthreads();
        }
        for (u32 i = threadIdx.x; i < N; i += blockDim.x) sm[i] = shoup(sm[i], sc.x, sc.y, q);
        __syncthreads();
    }
    for (u32 i = threadIdx.x; i < N; i += blockDim.x) p[i] = sm[i];
}

// ------------------------------------------------------------------------------------------
// Launchers
// ------------------------------------------------------------------------------------------
// log2(coefficients per tile) of the strided (which = 0) and contig (which = 1) pass.  The defaults
// can be overridden for tuning: FHE_B200_NTT_TILE_S / FHE_B200_NTT_TILE_C = 9 | 10 | 11 | 12.
static int tile_log(int which) {
    static int tl[2] = {0, 0};
    if (!tl[which]) {
        const char* e = getenv(which ? "FHE_B200_NTT_TILE_C" : "FHE_B200_NTT_TILE_S");
        int v = e ? atoi(e) : 0;
        tl[which] = (v >= 9 && v <= 12) ? v : NTT_TILE_LOG_DEFAULT[which];
    }
    return tl[which];
}

// FHE_B200_NTT_PF: 0 one-tile kernels, 1 the persistent prefetching strided kernel for large batches, 2 always (tests)
static int ntt_pf_mode() {
    const char* e = getenv("FHE_B200_NTT_PF");        // read every time: the tests switch it
    return e && *e ? atoi(e) : NTT_PF_DEFAULT;
}

template <int RL, int TL, bool INV>
static cudaError_t launch_strided_t(const RingDev& r, u64* data, const u64* in, u32 outT, u32 inT, size_t batch, cudaStream_t st) {
    using K = StridedCfg<RL, TL>;
    if (K::SMEM > 48 * 1024) {                         // a function attribute belongs to ONE device: remember it per device
        static std::atomic<unsigned long long> attr_done{0};
        int dev = 0;
        cudaError_t e = cudaGetDevice(&dev);
        if (e != cudaSuccess) return e;
        if (dev >= 64 || !((attr_done.load(std::memory_order_relaxed) >> dev) & 1)) {
            e = cudaFuncSetAttribute(ntt_strided_kernel<RL, TL, INV>, cudaFuncAttributeMaxDynamicSharedMemorySize, K::SMEM);
            if (e != cudaSuccess) return e;
            if (dev < 64) attr_done.fetch_or(1ull << dev, std::memory_order_relaxed);
        }
    }
    // batch items along y (and z beyond 32768 of them: the largest divisor of the batch that fits)
    unsigned by = (unsigned)batch, bz = 1;
    while (by > 32768u) {
        unsigned d = 2;
        while (by % d) d++;
        by /= d; bz *= d;
    }
    if (bz > 65535u) return cudaErrorInvalidValue;
    if constexpr (K::A == 4 && K::B >= 1) {
        const size_t ntiles = (size_t)r.T * (256 / K::C) * batch;
        if (ntt_pf_mode() != 0 && ntiles < 0x7fffffffull && (ntt_pf_mode() == 2 || ntiles >= 148 * 64)) {
            static std::atomic<int> occ_cache{0};
            int occ = occ_cache.load(std::memory_order_relaxed);
            if (occ == 0) {
                if (2 * K::SMEM > 48 * 1024) {
                    cudaError_t e = cudaFuncSetAttribute(ntt_strided_pf_kernel<RL, TL, INV>, cudaFuncAttributeMaxDynamicSharedMemorySize, 2 * K::SMEM);
                    if (e != cudaSuccess) return e;
                }
                if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, ntt_strided_pf_kernel<RL, TL, INV>, K::THREADS, 2 * K::SMEM) != cudaSuccess || occ < 1) occ = 1;
                occ_cache.store(occ, std::memory_order_relaxed);
            }
            int dev = 0, sms = 148;
            cudaGetDevice(&dev);
            cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
            size_t g = (size_t)sms * occ;
            if (g > ntiles) g = ntiles;
            ntt_strided_pf_kernel<RL, TL, INV><<<(unsigned)g, K::THREADS, 2 * K::SMEM, st>>>(
                data, in, INV ? r.twI : r.twF, r.mods, r.invc, r.T, r.logN, outT, inT, (u32)ntiles);
            kernel_launch_count_add(1);
            return cudaGetLastError();
        }
    }
    dim3 grid(r.T * (256 / K::C), by, bz);
    ntt_strided_kernel<RL, TL, INV><<<grid, K::THREADS, K::SMEM, st>>>(
        data, in, INV ? r.twI : r.twF, r.mods, r.invc, r.T, r.logN, outT, inT);
    kernel_launch_count_add(1);
    return cudaGetLastError();
}

template <int RL, bool INV>
static cudaError_t launch_strided_rl(const RingDev& r, u64* data, const u64* in, u32 outT, u32 inT, size_t batch, cudaStream_t st) {
    if constexpr (RL >= 4) {                       // smaller rings: the tile is already below 4096 coefficients
        switch (tile_log(0)) {
            case 9:  if constexpr (RL >= 7) return launch_strided_t<RL, 9, INV>(r, data, in, outT, inT, batch, st);
            case 10: if constexpr (RL >= 6) return launch_strided_t<RL, 10, INV>(r, data, in, outT, inT, batch, st);
            case 11: if constexpr (RL >= 5) return launch_strided_t<RL, 11, INV>(r, data, in, outT, inT, batch, st);
        }
    }
    return launch_strided_t<RL, 12, INV>(r, data, in, outT, inT, batch, st);
}

template <bool INV>
static cudaError_t launch_strided(const RingDev& r, u64* data, const u64* in, u32 outT, u32 inT, size_t batch, cudaStream_t st) {
    switch (r.logN - 8) {
        case 1: return launch_strided_rl<1, INV>(r, data, in, outT, inT, batch, st);
        case 2: return launch_strided_rl<2, INV>(r, data, in, outT, inT, batch, st);
        case 3: return launch_strided_rl<3, INV>(r, data, in, outT, inT, batch, st);
        case 4: return launch_strided_rl<4, INV>(r, data, in, outT, inT, batch, st);
        case 5: return launch_strided_rl<5, INV>(r, data, in, outT, inT, batch, st);
        case 6: return launch_strided_rl<6, INV>(r, data, in, outT, inT, batch, st);
        case 7: return launch_strided_rl<7, INV>(r, data, in, outT, inT, batch, st);
        case 8: return launch_strided_rl<8, INV>(r, data, in, outT, inT, batch, st);
    }
    return cudaErrorInvalidValue;
}

template <int LB, bool INV>
static cudaError_t launch_contig_t(const RingDev& r, u64* data, const u64* in, u32 outT, u32 inT, size_t batch, cudaStream_t st) {
    using K = ContigCfg<LB>;
    const u32 N1 = r.N >> 8, NC = K::NC;
    u32 logCB = 0;                                   // batch items per tile: largest 2^k <= min(batch, NC) ...
    while (logCB < (u32)LB && (2u << logCB) <= batch) logCB++;
    while ((NC >> logCB) > N1) logCB++;              // ... but never more blocks per tile than the ring has
    const u32 CB = 1u << logCB, CH = NC >> logCB;
    size_t grid = (size_t)r.T * ((batch + CB - 1) / CB) * (N1 / CH);
    ntt_contig_kernel<LB, INV><<<(unsigned)grid, K::THREADS, K::SMEM + 128, st>>>(data, in, INV ? r.twI : r.twF, r.mods,
                                                                          r.T, r.logN, (u32)batch, logCB, outT, inT);
    kernel_launch_count_add(1);
    return cudaGetLastError();
}

template <bool INV>
static cudaError_t launch_contig(const RingDev& r, u64* data, const u64* in, u32 outT, u32 inT, size_t batch, cudaStream_t st) {
    const u32 N1 = r.N >> 8;
    switch (tile_log(1)) {
        case 9:  return launch_contig_t<1, INV>(r, data, in, outT, inT, batch, st);
        case 10: if (N1 >= 4 || batch >= 4) return launch_contig_t<2, INV>(r, data, in, outT, inT, batch, st);
        case 11: if (N1 >= 8 || batch >= 8) return launch_contig_t<3, INV>(r, data, in, outT, inT, batch, st);
    }
    return launch_contig_t<4, INV>(r, data, in, outT, inT, batch, st);
}

template <bool INV>
static cudaError_t launch_small(const RingDev& r, u64* data, const u64* in, u32 outT, u32 inT, size_t batch, cudaStream_t st) {
    size_t smem = (size_t)r.N * 8;
    ntt_small_kernel<INV><<<(unsigned)(batch * r.T), 256, smem, st>>>(data, in, INV ? r.twI : r.twF,
                                                                     r.mods, r.invc, r.T, r.logN, outT, inT);
    kernel_launch_count_add(1);
    return cudaGetLastError();
}

// Where the single-pass kernel beats the two passes (measured on the B200): up to N = 2^11 in both
// directions (+8..21 % at large batches: 148-159 against 132-138 G coeff/s), at N = 2^12 forward (+1.5 %) and, for grids of
// at most one wave, inverse too (13 against 15 us: one launch).  At N = 2^13 and 2^14 it loses (109 / 89 against 127 / 125
// G coeff/s, and 17 / 29 against 15 us for one wave): the transform is bound by the integer pipe, not by HBM, and one big CTA
// per tower runs its five phases behind block-wide barriers instead of overlapping them with other CTAs' phases.
static bool ntt_fused_wins(const RingDev& r, size_t batch, bool inverse) {
    if (r.logN < 9 || r.logN > 12) return false;
    if (r.logN <= 11) return true;
    return !inverse || batch * r.T <= 148 * 4;
}

// FHE_B200_NTT_FUSED: 0 two passes, 1 the single-pass kernel where it was measured to win, 2 always for 2^9 <= N <= 2^14 (tests)
static int ntt_fused_mode() {
    const char* e = getenv("FHE_B200_NTT_FUSED");       // read every time: the tests switch it
    return e && *e ? atoi(e) : NTT_FUSED_DEFAULT;
}

template <int RL, bool INV>
static cudaError_t launch_fused_t(const RingDev& r, u64* data, const u64* in, u32 outT, u32 inT, size_t batch, cudaStream_t st) {
    using K = FusedCfg<RL>;
    if (K::SMEM > 48 * 1024) {
        static std::atomic<unsigned long long> attr_done{0};
        int dev = 0;
        cudaError_t e = cudaGetDevice(&dev);
        if (e != cudaSuccess) return e;
        if (dev >= 64 || !((attr_done.load(std::memory_order_relaxed) >> dev) & 1)) {
            e = cudaFuncSetAttribute(ntt_fused_kernel<RL, INV>, cudaFuncAttributeMaxDynamicSharedMemorySize, K::SMEM);
            if (e != cudaSuccess) return e;
            if (dev < 64) attr_done.fetch_or(1ull << dev, std::memory_order_relaxed);
        }
    }
    const size_t units = batch * r.T;
    if (units > 0x7fffffffull) return cudaErrorInvalidValue;
    ntt_fused_kernel<RL, INV><<<(unsigned)units, K::THREADS, K::SMEM, st>>>(data, in, INV ? r.twI : r.twF, r.mods, r.invc, r.T, outT, inT);
    kernel_launch_count_add(1);
    return cudaGetLastError();
}

template <bool INV>
static cudaError_t launch_fused(const RingDev& r, u64* data, const u64* in, u32 outT, u32 inT, size_t batch, cudaStream_t st) {
    switch (r.logN) {
        case 9: return launch_fused_t<1, INV>(r, data, in, outT, inT, batch, st);
        case 10: return launch_fused_t<2, INV>(r, data, in, outT, inT, batch, st);
        case 11: return launch_fused_t<3, INV>(r, data, in, outT, inT, batch, st);
        case 12: return launch_fused_t<4, INV>(r, data, in, outT, inT, batch, st);
        case 13: return launch_fused_t<5, INV>(r, data, in, outT, inT, batch, st);
        case 14: return launch_fused_t<6, INV>(r, data, in, outT, inT, batch, st);
        default: return cudaErrorInvalidValue;
    }
}

cudaError_t launch_ntt(const RingDev& r, uint64_t* data, size_t batch, bool inverse,
                       cudaStream_t st, int algo, const uint64_t* src, uint32_t outT, uint32_t inT) {
    if (batch == 0) return cudaSuccess;
    if (batch * r.T * 16 > 0x7fffffffull) return cudaErrorInvalidValue;   // grid.x limit
    u64* d = reinterpret_cast<u64*>(data);
    const u64* in = src ? reinterpret_cast<const u64*>(src) : d;         // the first pass reads src, everything else is in place
    if (!outT) outT = r.T;
    if (!inT) inT = src ? r.T : outT;
    if (r.logN <= 8 || (algo == 1 && r.logN <= 12))
        return inverse ? launch_small<true>(r, d, in, outT, inT, batch, st) : launch_small<false>(r, d, in, outT, inT, batch, st);
    if (r.logN > 16) return cudaErrorInvalidValue;
    if (algo == 0 && r.logN <= 14) {
        const int fm = ntt_fused_mode();
        if (fm == 2 || (fm == 1 && ntt_fused_wins(r, batch, inverse)))
            return inverse ? launch_fused<true>(r, d, in, outT, inT, batch, st) : launch_fused<false>(r, d, in, outT, inT, batch, st);
    }
    if (algo != 2 && r.logN >= 14 && (algo == 3 || ntt_ws_applies(r, batch)))
        return launch_ntt_ws(r, data, batch, inverse, st, reinterpret_cast<const uint64_t*>(in), outT, inT);
    // L2-resident hand-over between the two passes: the batch is walked in chunks small enough that what the first
    // pass of a chunk writes is still in the 126 MB L2 when the second pass reads it, so a transform costs one HBM
    // round trip (16 B per coefficient) instead of two.  FHE_B200_NTT_L2_MB = chunk size (0: whole batch per launch).
    static long l2_mb = -1;
    if (l2_mb < 0) {
        const char* e = getenv("FHE_B200_NTT_L2_MB");
        l2_mb = e && *e ? atol(e) : NTT_L2_CHUNK_MB_DEFAULT;
    }
    size_t chunk = batch;
    if (l2_mb > 0) {
        const size_t item_bytes = (size_t)r.T * r.N * 8;
        chunk = std::max<size_t>(1, ((size_t)l2_mb << 20) / item_bytes);
    }
    for (size_t b0 = 0; b0 < batch; b0 += chunk) {
        const size_t nb = std::min(chunk, batch - b0);
        u64* dc = d + b0 * outT * (size_t)r.N;
        const u64* ic = in + b0 * inT * (size_t)r.N;
        cudaError_t e;
        if (!inverse) {
            e = launch_strided<false>(r, dc, ic, outT, inT, nb, st);
            if (e == cudaSuccess) e = launch_contig<false>(r, dc, dc, outT, outT, nb, st);
        } else {
            e = launch_contig<true>(r, dc, ic, outT, inT, nb, st);
            if (e == cudaSuccess) e = launch_strided<true>(r, dc, dc, outT, outT, nb, st);
        }
        if (e != cudaSuccess) return e;
    }
    return cudaSuccess;
}

}  // namespace fhe
