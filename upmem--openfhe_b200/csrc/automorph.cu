// automorph.cu — ring automorphism X -> X^k (k odd) over batched RNS polynomials for sm_100a.
//
// Device replacement for (citations relative to the reference tree):
//   DCRTPolyImpl::AutomorphismTransform(k) / (k, precomp)   lattice/hal/default/dcrtpoly-impl.h:348-369
//   PolyImpl::AutomorphismTransform(k)                       poly-impl.h:312-365
//   PolyImpl::AutomorphismTransform(k, precomp)              poly-impl.h:368-380
//   PrecomputeAutoMap                                        src/core/lib/math/nbtheory2.cpp:274-286
//
// EVALUATION format (values in bit-reversed order): out[rev(j)] = in[rev(((2j+1)k mod 2N) >> 1)].  The index map
// is computed on the fly (brev + one 32-bit multiply), no table is read.  Because idx = j*k + (k-1)/2 mod N has
// low bits that depend only on the low bits of j, the permutation maps every aligned block of 2^s outputs onto
// an aligned block of 2^s inputs: a warp's 32 outputs come from one aligned 256-byte block, and an aligned output
// pair (o, o+1) comes from an aligned input pair in either order — so one LDG.128 and one STG.128 per pair, every
// sector fully used.  HBM-bound: 16 B per coefficient.
// COEFFICIENT format: out[jk mod N] = +-in[j]; done as a gather with k^-1 mod 2N so that the stores coalesce
// (the loads are a stride-k^-1 walk over one tower, which the L2 absorbs: a tower is at most 512 KB).  A thread
// produces four aligned output pairs and issues its eight gathers before the first store; with one gather in
// flight per thread the kernel ran at 4.8 TB/s, with eight at the copy rate (6.9 TB/s).
#include "modarith.cuh"
#include "kernels.h"

namespace fhe {

__device__ __forceinline__ u32 rev_bits(u32 x, u32 logn) { return logn ? (__brev(x) >> (32 - logn)) : 0u; }

constexpr int AM_THREADS = 256;
constexpr int AM_UNROLL  = 4;

// one thread = AM_UNROLL aligned output pairs; npairs = batch*T*N/2, N >= 2
__global__ void __launch_bounds__(AM_THREADS)
automorph_eval_kernel(u64* __restrict__ dst, const u64* __restrict__ src, size_t npairs, u32 logN, u32 k) {
    const u32 mask = (1u << logN) - 1;
    const u32 half = (k - 1) >> 1;
    const size_t p0 = (size_t)blockIdx.x * (AM_THREADS * AM_UNROLL) + threadIdx.x;
    ulonglong2 v[AM_UNROLL];
    u32 swap[AM_UNROLL];
#pragma unroll
    for (int u = 0; u < AM_UNROLL; u++) {
        size_t p = p0 + (size_t)u * AM_THREADS;
        if (p < npairs) {
            size_t e = 2 * p;                             // even output index (global)
            u32 o = (u32)e & mask;
            size_t unit = e >> logN;
            u32 j = rev_bits(o, logN);
            u32 idx = (j * k + half) & mask;              // ((2j+1)k mod 2N) >> 1
            u32 s = rev_bits(idx, logN);                  // source of output o; output o+1 reads s ^ 1
            swap[u] = s & 1;
            const ulonglong2* sp = reinterpret_cast<const ulonglong2*>(src + (unit << logN) + (s & ~1u));
            asm volatile("ld.global.L1::no_allocate.v2.u64 {%0, %1}, [%2];" : "=l"(v[u].x), "=l"(v[u].y) : "l"(sp));
        }
    }
#pragma unroll
    for (int u = 0; u < AM_UNROLL; u++) {
        size_t p = p0 + (size_t)u * AM_THREADS;
        if (p < npairs) {
            ulonglong2 r = swap[u] ? make_ulonglong2(v[u].y, v[u].x) : v[u];
            ulonglong2* dp = reinterpret_cast<ulonglong2*>(dst) + p;
            asm volatile("st.global.L1::no_allocate.v2.u64 [%0], {%1, %2};" ::"l"(dp), "l"(r.x), "l"(r.y) : "memory");
        }
    }
}

// COEFFICIENT format, one thread per output coefficient.  kinv = k^-1 mod 2N.
__global__ void __launch_bounds__(AM_THREADS)
automorph_coef_kernel(u64* __restrict__ dst, const u64* __restrict__ src, size_t total, u32 logN, u32 T,
                      const TowerMod* __restrict__ mods, u32 kinv) {
    const u32 mask = (1u << logN) - 1;
    for (size_t e = (size_t)blockIdx.x * AM_THREADS + threadIdx.x; e < total; e += (size_t)gridDim.x * AM_THREADS) {
        u32 d = (u32)e & mask;
        size_t unit = e >> logN;
        u32 j2 = (d * kinv) & (2 * mask + 1);             // j2 * k = d (mod 2N)
        u64 v = src[(unit << logN) + (j2 & mask)];
        if (j2 >> logN) v = mods[(u32)(unit % T)].q - v;  // X^(jk) with jk in [N, 2N): sign flips (q - 0 stays q, like the reference)
        dst[e] = v;
    }
}

// The same for N >= 2: one thread = AM_UNROLL aligned output pairs, all 2 * AM_UNROLL gathers issued before the first store
// (one gather at a time leaves too little in flight), 16-byte stores.
__global__ void __launch_bounds__(AM_THREADS)
automorph_coef_pairs_kernel(u64* __restrict__ dst, const u64* __restrict__ src, size_t npairs, u32 logN, u32 T,
                            const TowerMod* __restrict__ mods, u32 kinv) {
    const u32 mask = (1u << logN) - 1, mask2 = 2 * mask + 1;
    const size_t p0 = (size_t)blockIdx.x * (AM_THREADS * AM_UNROLL) + threadIdx.x;
    ulonglong2 v[AM_UNROLL];
    u32 neg[AM_UNROLL];
    u64 q[AM_UNROLL];
#pragma unroll
    for (int u = 0; u < AM_UNROLL; u++) {
        size_t p = p0 + (size_t)u * AM_THREADS;
        if (p < npairs) {
            size_t e = 2 * p;
            u32 d = (u32)e & mask;
            size_t unit = e >> logN;
            u32 ja = (d * kinv) & mask2, jb = (ja + kinv) & mask2;     // j * k = d, d + 1 (mod 2N)
            const u64* base = src + (unit << logN);
            v[u].x = __ldg(base + (ja & mask));
            v[u].y = __ldg(base + (jb & mask));
            neg[u] = (ja >> logN) | ((jb >> logN) << 1);
            q[u] = mods[(u32)(unit % T)].q;
        }
    }
#pragma unroll
    for (int u = 0; u < AM_UNROLL; u++) {
        size_t p = p0 + (size_t)u * AM_THREADS;
        if (p < npairs) {
            ulonglong2 r = v[u];
            if (neg[u] & 1) r.x = q[u] - r.x;                          // q - 0 stays q, like the reference
            if (neg[u] & 2) r.y = q[u] - r.y;
            ulonglong2* dp = reinterpret_cast<ulonglong2*>(dst) + p;
            asm volatile("st.global.L1::no_allocate.v2.u64 [%0], {%1, %2};" ::"l"(dp), "l"(r.x), "l"(r.y) : "memory");
        }
    }
}

// two-argument form: out[j] = in[map[j]] per tower, map resident in HBM (N uint32)
__global__ void __launch_bounds__(AM_THREADS)
automorph_map_kernel(u64* __restrict__ dst, const u64* __restrict__ src, size_t total, u32 logN,
                     const u32* __restrict__ map) {
    const u32 mask = (1u << logN) - 1;
    for (size_t e = (size_t)blockIdx.x * AM_THREADS + threadIdx.x; e < total; e += (size_t)gridDim.x * AM_THREADS) {
        size_t unit = e >> logN;
        dst[e] = src[(unit << logN) + (__ldg(map + ((u32)e & mask)) & mask)];
    }
}

// the table form for N >= 2: aligned output pairs, AM_UNROLL pairs per thread, gathers issued before the stores
__global__ void __launch_bounds__(AM_THREADS)
automorph_map_pairs_kernel(u64* __restrict__ dst, const u64* __restrict__ src, size_t npairs, u32 logN,
                           const u32* __restrict__ map) {
    const u32 mask = (1u << logN) - 1;
    const size_t p0 = (size_t)blockIdx.x * (AM_THREADS * AM_UNROLL) + threadIdx.x;
    ulonglong2 v[AM_UNROLL];
#pragma unroll
    for (int u = 0; u < AM_UNROLL; u++) {
        size_t p = p0 + (size_t)u * AM_THREADS;
        if (p < npairs) {
            size_t e = 2 * p;
            const u64* base = src + ((e >> logN) << logN);
            uint2 m = __ldg(reinterpret_cast<const uint2*>(map + ((u32)e & mask)));
            v[u].x = __ldg(base + (m.x & mask));
            v[u].y = __ldg(base + (m.y & mask));
        }
    }
#pragma unroll
    for (int u = 0; u < AM_UNROLL; u++) {
        size_t p = p0 + (size_t)u * AM_THREADS;
        if (p < npairs) {
            ulonglong2* dp = reinterpret_cast<ulonglong2*>(dst) + p;
            asm volatile("st.global.L1::no_allocate.v2.u64 [%0], {%1, %2};" ::"l"(dp), "l"(v[u].x), "l"(v[u].y) : "memory");
        }
    }
}

static u32 inv_mod_pow2(u32 k) {          // k odd: inverse modulo 2^32 by Newton iteration
    u32 x = k;                             // correct to 3 bits
    for (int i = 0; i < 5; i++) x *= 2u - k * x;
    return x;
}

cudaError_t launch_automorphism(const RingDev& r, uint64_t* dst, const uint64_t* src, size_t batch, uint32_t k,
                                int format, cudaStream_t st) {
    size_t total = batch * r.T * (size_t)r.N;
    if (total == 0) return cudaSuccess;
    if (format == 0 && r.logN >= 1) {
        size_t npairs = total / 2;
        size_t grid = (npairs + AM_THREADS * AM_UNROLL - 1) / (AM_THREADS * AM_UNROLL);
        if (grid > 0x7fffffffull) return cudaErrorInvalidValue;
        automorph_eval_kernel<<<(unsigned)grid, AM_THREADS, 0, st>>>(reinterpret_cast<u64*>(dst),
                                                                   reinterpret_cast<const u64*>(src), npairs, r.logN, k);
    } else if (r.logN >= 1) {
        u32 kinv = inv_mod_pow2(k) & (2 * r.N - 1);
        size_t npairs = total / 2;
        size_t grid = (npairs + AM_THREADS * AM_UNROLL - 1) / (AM_THREADS * AM_UNROLL);
        if (grid > 0x7fffffffull) return cudaErrorInvalidValue;
        automorph_coef_pairs_kernel<<<(unsigned)grid, AM_THREADS, 0, st>>>(reinterpret_cast<u64*>(dst),
                                                                         reinterpret_cast<const u64*>(src), npairs, r.logN,
                                                                         r.T, r.mods, kinv);
    } else {
        // N = 1 in EVALUATION format is the identity, which the gather with kinv = 1 also gives
        u32 kinv = format == 0 ? 1u : (inv_mod_pow2(k) & (2 * r.N - 1));
        size_t grid = (total + AM_THREADS - 1) / AM_THREADS;
        if (grid > 148 * 32) grid = 148 * 32;
        automorph_coef_kernel<<<(unsigned)grid, AM_THREADS, 0, st>>>(reinterpret_cast<u64*>(dst),
                                                                   reinterpret_cast<const u64*>(src), total, r.logN, r.T,
                                                                   r.mods, kinv);
    }
    kernel_launch_count_add(1);
    return cudaGetLastError();
}

cudaError_t launch_automorphism_map(const RingDev& r, uint64_t* dst, const uint64_t* src, size_t batch,
                                    const uint32_t* map, cudaStream_t st) {
    size_t total = batch * r.T * (size_t)r.N;
    if (total == 0) return cudaSuccess;
    if (r.logN >= 1 && (reinterpret_cast<uintptr_t>(map) & 7) == 0) {
        size_t npairs = total / 2;
        size_t grid = (npairs + AM_THREADS * AM_UNROLL - 1) / (AM_THREADS * AM_UNROLL);
        if (grid > 0x7fffffffull) return cudaErrorInvalidValue;
        automorph_map_pairs_kernel<<<(unsigned)grid, AM_THREADS, 0, st>>>(reinterpret_cast<u64*>(dst),
                                                                        reinterpret_cast<const u64*>(src), npairs, r.logN, map);
        kernel_launch_count_add(1);
        return cudaGetLastError();
    }
    size_t grid = (total + AM_THREADS - 1) / AM_THREADS;
    if (grid > 148 * 32) grid = 148 * 32;
    automorph_map_kernel<<<(unsigned)grid, AM_THREADS, 0, st>>>(reinterpret_cast<u64*>(dst),
                                                              reinterpret_cast<const u64*>(src), total, r.logN, map);
    kernel_launch_count_add(1);
    return cudaGetLastError();
}

}  // namespace fhe
