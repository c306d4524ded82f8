// elementwise.cu — tower-wise modular add / sub / mul over batched RNS polynomials for sm_100a.
//
// Device replacement for (citations relative to the reference tree):
//   DCRTPolyImpl::Plus / operator+= / Minus / operator-=   dcrtpoly.h:164-179, dcrtpoly-impl.h:398-416
//   DCRTPolyImpl::Times / operator*=                        dcrtpoly.h:142-148, 185-200
//   NativeVectorT::ModAddNoCheckEq / ModSubEq / ModMulNoCheckEq   mubintvecnat.h:426-432, 501-513
// and for what the UPMEM kernels add-mod.c / sub-mod.c / mult-mod.c (vector, vector_eq, scalar,
// scalar_eq variants) were meant to compute.
//
// HBM-bound: 24 B of traffic per coefficient (read a, read b, write dst).  One launch covers the
// whole [batch][tower][N] block; every thread moves 4 x 16 B per operand with all loads issued
// before the first use, streaming (no L1 allocation).  The modulus record of the tower is looked
// up once per CTA when a CTA's 2048-coefficient tile lies inside one tower (N >= 2048).
#include "modarith.cuh"
#include "kernels.h"

#include <atomic>

namespace fhe {

static std::atomic<int> g_launches{0};
int kernel_launch_count() { return g_launches.load(); }
void kernel_launch_count_add(int n) { g_launches.fetch_add(n); }

__device__ __forceinline__ ulonglong2 ld_stream(const ulonglong2* p) {
    ulonglong2 v;
    asm volatile("ld.global.L1::no_allocate.v2.u64 {%0, %1}, [%2];" : "=l"(v.x), "=l"(v.y) : "l"(p));
    return v;
}
__device__ __forceinline__ void st_stream(ulonglong2* p, ulonglong2 v) {
    asm volatile("st.global.L1::no_allocate.v2.u64 [%0], {%1, %2};" ::"l"(p), "l"(v.x), "l"(v.y) : "memory");
}

template <int OP>
__device__ __forceinline__ u64 ew_apply(u64 a, u64 b, const TowerMod& m) {
    if (OP == EW_ADD) return addmod(a, b, m.q);
    if (OP == EW_SUB) return submod(a, b, m.q);
    return mulmod_barrett(a, b, m);
}

constexpr int EW_THREADS = 256;
constexpr int EW_UNROLL  = 4;
constexpr int EW_TILE    = EW_THREADS * EW_UNROLL * 2;     // coefficients per CTA (2048)

// total = batch*T*N coefficients, nvec = total/2.  UNIFORM: N >= EW_TILE, so a CTA sees one tower.
template <int OP, bool UNIFORM>
__global__ void __launch_bounds__(EW_THREADS)
ew_ring_kernel(u64* dst, const u64* a, const u64* b,
               size_t nvec, u32 logN, u32 T, const TowerMod* __restrict__ mods) {
    const size_t v0 = (size_t)blockIdx.x * (EW_THREADS * EW_UNROLL) + threadIdx.x;
    const ulonglong2* av = reinterpret_cast<const ulonglong2*>(a);
    const ulonglong2* bv = reinterpret_cast<const ulonglong2*>(b);
    ulonglong2* dv = reinterpret_cast<ulonglong2*>(dst);
    ulonglong2 x[EW_UNROLL], y[EW_UNROLL];
#pragma unroll
    for (int k = 0; k < EW_UNROLL; k++) {
        size_t v = v0 + (size_t)k * EW_THREADS;
        if (v < nvec) { x[k] = ld_stream(av + v); y[k] = ld_stream(bv + v); }
    }
    TowerMod m;
    if (UNIFORM) m = mods[(u32)((((size_t)blockIdx.x * EW_TILE) >> logN) % T)];
#pragma unroll
    for (int k = 0; k < EW_UNROLL; k++) {
        size_t v = v0 + (size_t)k * EW_THREADS;
        if (v < nvec) {
            ulonglong2 r;
            if (UNIFORM || logN >= 1) {
                if (!UNIFORM) m = mods[(u32)(((2 * v) >> logN) % T)];
                r.x = ew_apply<OP>(x[k].x, y[k].x, m);
                r.y = ew_apply<OP>(x[k].y, y[k].y, m);
            } else {                                       // N = 1: neighbours belong to different towers
                r.x = ew_apply<OP>(x[k].x, y[k].x, mods[(u32)((2 * v) % T)]);
                r.y = ew_apply<OP>(x[k].y, y[k].y, mods[(u32)((2 * v + 1) % T)]);
            }
            st_stream(dv + v, r);
        }
    }
}

// dst = a (+|-) s_tower: one constant per tower added to (subtracted from) every coefficient — DCRTPolyImpl::Plus /
// Minus(const std::vector<Integer>&) (dcrtpoly-impl.h:556-563, 577-584; CKKS EvalAdd / EvalSub with a real constant).  The
// constants (already reduced) travel in the launch parameters.
struct TowerScalars { u64 s[EW_SCALARS_MAX]; };
template <int OP>
__global__ void __launch_bounds__(EW_THREADS)
ew_tower_scalar_kernel(u64* dst, const u64* a, size_t nvec, u32 logN, u32 T, const TowerMod* __restrict__ mods, TowerScalars sc) {
    const ulonglong2* av = reinterpret_cast<const ulonglong2*>(a);
    ulonglong2* dv = reinterpret_cast<ulonglong2*>(dst);
    for (size_t v = (size_t)blockIdx.x * EW_THREADS + threadIdx.x; v < nvec; v += (size_t)gridDim.x * EW_THREADS) {
        const u32 tw = (u32)(((2 * v) >> logN) % T);
        const u64 q = mods[tw].q, s = sc.s[tw];
        ulonglong2 x = av[v];
        x.x = OP == EW_ADD ? addmod(x.x, s, q) : submod(x.x, s, q);
        x.y = OP == EW_ADD ? addmod(x.y, s, q) : submod(x.y, s, q);
        dv[v] = x;
    }
}

cudaError_t launch_tower_scalars(const RingDev& r, int op, uint64_t* dst, const uint64_t* a, const uint64_t* scalars, size_t batch,
                                 cudaStream_t st) {
    const size_t total = batch * r.T * (size_t)r.N;
    if (total == 0) return cudaSuccess;
    if (r.T > EW_SCALARS_MAX || r.N < 2 || (op != EW_ADD && op != EW_SUB)) return cudaErrorInvalidValue;
    if ((((uintptr_t)dst | (uintptr_t)a)) & 15) return cudaErrorMisalignedAddress;
    TowerScalars sc;
    for (uint32_t i = 0; i < r.T; i++) sc.s[i] = scalars[i];
    const size_t nvec = total / 2;
    size_t g = (nvec + EW_THREADS - 1) / EW_THREADS;
    const unsigned grid = (unsigned)(g > 148 * 32 ? 148 * 32 : g);
    u64* d = reinterpret_cast<u64*>(dst);
    const u64* pa = reinterpret_cast<const u64*>(a);
    if (op == EW_ADD) ew_tower_scalar_kernel<EW_ADD><<<grid, EW_THREADS, 0, st>>>(d, pa, nvec, r.logN, r.T, r.mods, sc);
    else ew_tower_scalar_kernel<EW_SUB><<<grid, EW_THREADS, 0, st>>>(d, pa, nvec, r.logN, r.T, r.mods, sc);
    kernel_launch_count_add(1);
    return cudaGetLastError();
}

// Ciphertext tensor product (LeveledSHEBase::EvalMultCore, 2 x 2 case, base-leveledshe.cpp:667-672):
//   c0 = a0*b0, c1 = a1*b0 + a0*b1, c2 = a1*b1   per tower, per coefficient.
// One pass instead of four multiplies and an add: 56 B of HBM traffic per coefficient slot instead of
// 120 B.  c1 is formed as (a0+a1)(b0+b1) - c0 - c2 (three Barrett products, not four), which keeps the
// kernel on the HBM side of the ridge; every intermediate is a canonical residue, so the outputs are
// the same words the reference's four Times and one += produce.
constexpr int TP_THREADS = 256;
constexpr int TP_UNROLL  = 2;

template <bool UNIFORM>
__global__ void __launch_bounds__(TP_THREADS)
ew_tensor_kernel(u64* c0, u64* c1, u64* c2, const u64* a0, const u64* a1, const u64* b0, const u64* b1,
                 size_t nvec, u32 logN, u32 T, const TowerMod* __restrict__ mods) {
    const size_t v0 = (size_t)blockIdx.x * (TP_THREADS * TP_UNROLL) + threadIdx.x;
    ulonglong2 xa0[TP_UNROLL], xa1[TP_UNROLL], xb0[TP_UNROLL], xb1[TP_UNROLL];
#pragma unroll
    for (int k = 0; k < TP_UNROLL; k++) {
        size_t v = v0 + (size_t)k * TP_THREADS;
        if (v < nvec) {
            xa0[k] = ld_stream(reinterpret_cast<const ulonglong2*>(a0) + v);
            xa1[k] = ld_stream(reinterpret_cast<const ulonglong2*>(a1) + v);
            xb0[k] = ld_stream(reinterpret_cast<const ulonglong2*>(b0) + v);
            xb1[k] = ld_stream(reinterpret_cast<const ulonglong2*>(b1) + v);
        }
    }
    TowerMod m;
    if (UNIFORM) m = mods[(u32)((((size_t)blockIdx.x * (TP_THREADS * TP_UNROLL * 2)) >> logN) % T)];
    auto one = [&](u64 p0, u64 p1, u64 r0, u64 r1, const TowerMod& mm, u64& o0, u64& o1, u64& o2) {
        o0 = mulmod_barrett(p0, r0, mm);
        o2 = mulmod_barrett(p1, r1, mm);
        u64 cross = mulmod_barrett(addmod(p0, p1, mm.q), addmod(r0, r1, mm.q), mm);
        o1 = submod(submod(cross, o0, mm.q), o2, mm.q);
    };
#pragma unroll
    for (int k = 0; k < TP_UNROLL; k++) {
        size_t v = v0 + (size_t)k * TP_THREADS;
        if (v < nvec) {
            ulonglong2 r0, r1, r2;
            if (UNIFORM || logN >= 1) {
                if (!UNIFORM) m = mods[(u32)(((2 * v) >> logN) % T)];
                one(xa0[k].x, xa1[k].x, xb0[k].x, xb1[k].x, m, r0.x, r1.x, r2.x);
                one(xa0[k].y, xa1[k].y, xb0[k].y, xb1[k].y, m, r0.y, r1.y, r2.y);
            } else {
                one(xa0[k].x, xa1[k].x, xb0[k].x, xb1[k].x, mods[(u32)((2 * v) % T)], r0.x, r1.x, r2.x);
                one(xa0[k].y, xa1[k].y, xb0[k].y, xb1[k].y, mods[(u32)((2 * v + 1) % T)], r0.y, r1.y, r2.y);
            }
            st_stream(reinterpret_cast<ulonglong2*>(c0) + v, r0);
            st_stream(reinterpret_cast<ulonglong2*>(c1) + v, r1);
            st_stream(reinterpret_cast<ulonglong2*>(c2) + v, r2);
        }
    }
}

cudaError_t launch_tensor_product(const RingDev& r, uint64_t* c0, uint64_t* c1, uint64_t* c2,
                                  const uint64_t* a0, const uint64_t* a1, const uint64_t* b0, const uint64_t* b1,
                                  size_t batch, cudaStream_t st) {
    size_t total = batch * r.T * (size_t)r.N;
    if (total == 0) return cudaSuccess;
    uintptr_t all = (uintptr_t)c0 | (uintptr_t)c1 | (uintptr_t)c2 | (uintptr_t)a0 | (uintptr_t)a1 | (uintptr_t)b0 | (uintptr_t)b1;
    if ((total & 1) || (all & 15)) return cudaErrorMisalignedAddress;
    size_t nvec = total / 2;
    size_t blocks = (nvec + TP_THREADS * TP_UNROLL - 1) / (TP_THREADS * TP_UNROLL);
    if (blocks >= 0x7fffffffull) return cudaErrorInvalidValue;
    u64 *d0 = reinterpret_cast<u64*>(c0), *d1 = reinterpret_cast<u64*>(c1), *d2 = reinterpret_cast<u64*>(c2);
    const u64 *pa0 = reinterpret_cast<const u64*>(a0), *pa1 = reinterpret_cast<const u64*>(a1);
    const u64 *pb0 = reinterpret_cast<const u64*>(b0), *pb1 = reinterpret_cast<const u64*>(b1);
    if (r.N >= (uint32_t)(TP_THREADS * TP_UNROLL * 2))
        ew_tensor_kernel<true><<<(unsigned)blocks, TP_THREADS, 0, st>>>(d0, d1, d2, pa0, pa1, pb0, pb1, nvec, r.logN, r.T, r.mods);
    else
        ew_tensor_kernel<false><<<(unsigned)blocks, TP_THREADS, 0, st>>>(d0, d1, d2, pa0, pa1, pb0, pb1, nvec, r.logN, r.T, r.mods);
    kernel_launch_count_add(1);
    return cudaGetLastError();
}

// Flat vector with one modulus; scalar element loop (tails, unaligned buffers, tiny vectors).
// SCALAR: second operand is the broadcast value sb.
template <int OP, bool SCALAR>
__global__ void __launch_bounds__(256)
ew_flat_kernel(u64* dst, const u64* a, const u64* b, u64 sb,
               size_t first, size_t n, TowerMod m) {
    size_t i = first + (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) dst[i] = ew_apply<OP>(a[i], SCALAR ? sb : b[i], m);
}

// Flat vector, 128-bit path (16 B aligned pointers), nvec = n/2.
template <int OP, bool SCALAR>
__global__ void __launch_bounds__(EW_THREADS)
ew_flat_vec_kernel(u64* dst, const u64* a, const u64* b, u64 sb,
                   size_t nvec, TowerMod m) {
    const size_t v0 = (size_t)blockIdx.x * (EW_THREADS * EW_UNROLL) + threadIdx.x;
    const ulonglong2* av = reinterpret_cast<const ulonglong2*>(a);
    const ulonglong2* bv = reinterpret_cast<const ulonglong2*>(b);
    ulonglong2* dv = reinterpret_cast<ulonglong2*>(dst);
    ulonglong2 x[EW_UNROLL], y[EW_UNROLL];
#pragma unroll
    for (int k = 0; k < EW_UNROLL; k++) {
        size_t v = v0 + (size_t)k * EW_THREADS;
        if (v < nvec) {
            x[k] = ld_stream(av + v);
            if (!SCALAR) y[k] = ld_stream(bv + v); else y[k] = make_ulonglong2(sb, sb);
        }
    }
#pragma unroll
    for (int k = 0; k < EW_UNROLL; k++) {
        size_t v = v0 + (size_t)k * EW_THREADS;
        if (v < nvec) {
            ulonglong2 r;
            r.x = ew_apply<OP>(x[k].x, y[k].x, m);
            r.y = ew_apply<OP>(x[k].y, y[k].y, m);
            st_stream(dv + v, r);
        }
    }
}

static TowerMod make_mod_host(uint64_t q) {
    TowerMod m;
    unsigned msb = q ? 64u - (unsigned)__builtin_clzll(q) : 0u;
    m.q = q;
    m.mu = (uint64_t)(((unsigned __int128)1 << (2 * msb + 3)) / q);
    m.n = msb - 2;
    m.s = msb + 5;
    towermod_set_reduce(m);
    return m;
}

template <int OP>
static cudaError_t launch_ring_op(const RingDev& r, u64* dst, const u64* a, const u64* b, size_t total,
                                  cudaStream_t st) {
    size_t nvec = total / 2;
    unsigned grid = (unsigned)((nvec + EW_THREADS * EW_UNROLL - 1) / (EW_THREADS * EW_UNROLL));
    if (r.N >= (uint32_t)EW_TILE)
        ew_ring_kernel<OP, true><<<grid, EW_THREADS, 0, st>>>(dst, a, b, nvec, r.logN, r.T, r.mods);
    else
        ew_ring_kernel<OP, false><<<grid, EW_THREADS, 0, st>>>(dst, a, b, nvec, r.logN, r.T, r.mods);
    kernel_launch_count_add(1);
    return cudaGetLastError();
}

cudaError_t launch_elementwise(const RingDev& r, int op, uint64_t* dst, const uint64_t* a,
                               const uint64_t* b, size_t batch, cudaStream_t st) {
    size_t total = batch * r.T * (size_t)r.N;
    if (total == 0) return cudaSuccess;
    if ((total & 1) || (((uintptr_t)dst | (uintptr_t)a | (uintptr_t)b) & 15)) return cudaErrorMisalignedAddress;
    if (total / (2 * EW_THREADS * EW_UNROLL) >= 0x7fffffffull) return cudaErrorInvalidValue;
    u64* d = reinterpret_cast<u64*>(dst);
    const u64* pa = reinterpret_cast<const u64*>(a);
    const u64* pb = reinterpret_cast<const u64*>(b);
    switch (op) {
        case EW_ADD: return launch_ring_op<EW_ADD>(r, d, pa, pb, total, st);
        case EW_SUB: return launch_ring_op<EW_SUB>(r, d, pa, pb, total, st);
        case EW_MUL: return launch_ring_op<EW_MUL>(r, d, pa, pb, total, st);
    }
    return cudaErrorInvalidValue;
}

template <int OP, bool SCALAR>
static cudaError_t launch_flat(u64* dst, const u64* a, const u64* b, u64 sb, size_t n, uint64_t q,
                               cudaStream_t st) {
    if (n == 0) return cudaSuccess;
    TowerMod m = make_mod_host(q);
    bool aligned = !(((uintptr_t)dst | (uintptr_t)a | (SCALAR ? 0 : (uintptr_t)b)) & 15);
    size_t nvec = aligned ? n / 2 : 0;
    if (nvec) {
        unsigned grid = (unsigned)((nvec + EW_THREADS * EW_UNROLL - 1) / (EW_THREADS * EW_UNROLL));
        ew_flat_vec_kernel<OP, SCALAR><<<grid, EW_THREADS, 0, st>>>(dst, a, b, sb, nvec, m);
        kernel_launch_count_add(1);
    }
    size_t first = nvec * 2;
    if (first < n) {
        unsigned grid = (unsigned)((n - first + 255) / 256);
        ew_flat_kernel<OP, SCALAR><<<grid, 256, 0, st>>>(dst, a, b, sb, first, n, m);
        kernel_launch_count_add(1);
    }
    return cudaGetLastError();
}

cudaError_t launch_vec(int op, uint64_t* dst, const uint64_t* a, const uint64_t* b, size_t n,
                       uint64_t q, cudaStream_t st) {
    u64* d = reinterpret_cast<u64*>(dst);
    const u64* pa = reinterpret_cast<const u64*>(a);
    const u64* pb = reinterpret_cast<const u64*>(b);
    switch (op) {
        case EW_ADD: return launch_flat<EW_ADD, false>(d, pa, pb, 0, n, q, st);
        case EW_SUB: return launch_flat<EW_SUB, false>(d, pa, pb, 0, n, q, st);
        case EW_MUL: return launch_flat<EW_MUL, false>(d, pa, pb, 0, n, q, st);
    }
    return cudaErrorInvalidValue;
}

// Scalar operand is reduced mod q first, as NativeVectorT::ModAddEq(scalar) does (mubintvecnat.cpp:198-218).
cudaError_t launch_vec_scalar(int op, uint64_t* dst, const uint64_t* a, uint64_t s, size_t n,
                              uint64_t q, cudaStream_t st) {
    u64* d = reinterpret_cast<u64*>(dst);
    const u64* pa = reinterpret_cast<const u64*>(a);
    if (s >= q) s %= q;
    switch (op) {
        case EW_ADD: return launch_flat<EW_ADD, true>(d, pa, nullptr, s, n, q, st);
        case EW_SUB: return launch_flat<EW_SUB, true>(d, pa, nullptr, s, n, q, st);
        case EW_MUL: return launch_flat<EW_MUL, true>(d, pa, nullptr, s, n, q, st);
    }
    return cudaErrorInvalidValue;
}

// ---- order-independent digest of a buffer: (sum mod 2^64, xor) -----------------------------------
// Lets sharded results be compared without gathering them (config 4's results stay
// sharded, "only a checksum/sample is gathered"); matches sharding.checksum on the host.
__global__ void __launch_bounds__(256)
checksum_kernel(const u64* __restrict__ x, size_t n, u64* __restrict__ out) {
    u64 s = 0, xr = 0;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        u64 v = x[i];
        s += v; xr ^= v;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        s += __shfl_xor_sync(0xffffffffu, s, o);
        xr ^= __shfl_xor_sync(0xffffffffu, xr, o);
    }
    if ((threadIdx.x & 31) == 0) {
        atomicAdd(out, s);
        atomicXor(out + 1, xr);
    }
}

cudaError_t launch_checksum(const uint64_t* x, size_t n, uint64_t* out2, cudaStream_t st) {
    cudaError_t e = cudaMemsetAsync(out2, 0, 16, st);
    if (e != cudaSuccess || n == 0) return e;
    size_t blocks = (n + 255) / 256;
    if (blocks > 148 * 16) blocks = 148 * 16;
    checksum_kernel<<<(unsigned)blocks, 256, 0, st>>>(reinterpret_cast<const u64*>(x), n, reinterpret_cast<u64*>(out2));
    kernel_launch_count_add(1);
    return cudaGetLastError();
}

// ---- seeded uniform fill: value i of tower t depends only on (seed, i, q_t) ----
__device__ __forceinline__ u64 splitmix64(u64 z) {
    z += 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

__global__ void __launch_bounds__(256)
fill_uniform_kernel(u64* __restrict__ dst, size_t total, u32 logN, u32 T, const TowerMod* __restrict__ mods,
                    u64 seed) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= total) return;
    u64 q = mods[(u32)((i >> logN) % T)].q;
    u32 bits = 64 - __clzll(q - 1);
    u64 mask = bits >= 64 ? ~0ull : ((1ull << bits) - 1);
    u64 s = splitmix64(seed ^ (i * 0xD1342543DE82EF95ull));
    u64 v;
    do { s = splitmix64(s); v = s & mask; } while (v >= q);
    dst[i] = v;
}

cudaError_t launch_fill_uniform(const RingDev& r, uint64_t* dst, size_t batch, uint64_t seed,
                                cudaStream_t st) {
    size_t total = batch * r.T * (size_t)r.N;
    if (total == 0) return cudaSuccess;
    fill_uniform_kernel<<<(unsigned)((total + 255) / 256), 256, 0, st>>>(
        reinterpret_cast<u64*>(dst), total, r.logN, r.T, r.mods, seed);
    kernel_launch_count_add(1);
    return cudaGetLastError();
}

// ---- integer-pipe probe: 8 independent chains of lazy Shoup products (mulhi64 + 2 mullo64), the
// multiply work of one NTT butterfly.  Its rate is the integer roofline the NTT is reported against.
__global__ void imad_probe_kernel(u64* out, int iters) {
    u64 a[8], w = 0x9E3779B97F4A7C15ull + threadIdx.x, wp = w ^ 0x5555555555555555ull;
    const u64 nq = 0ull - (0x0FFFFFFFFFFC0001ull);
#pragma unroll
    for (int k = 0; k < 8; k++) a[k] = w * (k + 3) + blockIdx.x;
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int k = 0; k < 8; k++) a[k] = shoup_lazy(a[k], w, wp, nq);
    }
    u64 s = 0;
#pragma unroll
    for (int k = 0; k < 8; k++) s ^= a[k];
    if (s == 0x1234567ull) out[0] = s;
}

cudaError_t launch_imad_probe(uint64_t* out, int blocks, int threads, int iters, cudaStream_t st) {
    imad_probe_kernel<<<blocks, threads, 0, st>>>(reinterpret_cast<u64*>(out), iters);
    kernel_launch_count_add(1);
    return cudaGetLastError();
}

}  // namespace fhe
