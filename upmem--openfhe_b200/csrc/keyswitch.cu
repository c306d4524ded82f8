// keyswitch.cu — the inner product of hybrid key switching for sm_100a.
//
// KeySwitchHYBRID::EvalFastKeySwitchCoreExt (src/pke/lib/keyswitch/keyswitch-hybrid.cpp:442-480):
//   cTilda0 = sum_j digit_j * b_j,  cTilda1 = sum_j digit_j * a_j      tower by tower over Ql and P,
// with digit_j the extended digits EvalKeySwitchPrecomputeCore produced (:330-420) and (b_j, a_j) the evaluation key.
// The reference does it with one PolyImpl product and one sum per (digit, tower) and output — 6 passes over memory per
// digit; here one pass reads every digit once and writes both outputs: (8 * digits + 16) B per coefficient slot.  The
// key is the same for every ciphertext of the batch: a thread keeps its 2 * digits key words in registers and walks the
// batch, so the key costs no HBM traffic beyond its first read.  Every term is a Barrett product (the reference's
// ModMulFastEq, ubintnat.h:1398-1413) added with ModAddFastEq, so each output is the canonical residue the reference
// stores.  The digit decomposition itself (INTT, ApproxSwitchCRTBasis, NTT) is composed from the existing kernels in
// runtime.cu (fhe_keyswitch_precompute).
#include "modarith.cuh"
#include "kernels.h"

#include <algorithm>

namespace fhe {

constexpr int KS_THREADS = 256;

// digits: [np][batch][T][N]; keys: [np][T][N]; outputs [batch][T][N].  One thread = one aligned coefficient pair of one
// tower; grid.y strides over the batch.
template <int NP>
__global__ void __launch_bounds__(KS_THREADS)
ks_inner_kernel(u64* __restrict__ o0, u64* __restrict__ o1, const u64* __restrict__ digits, const u64* __restrict__ kb,
                const u64* __restrict__ ka, u32 np, size_t batch, u32 logN, u32 T, const TowerMod* __restrict__ mods) {
    const size_t pairsPerItem = ((size_t)T << logN) >> 1;
    const size_t p = (size_t)blockIdx.x * KS_THREADS + threadIdx.x;
    if (p >= pairsPerItem) return;
    const u32 tower = (u32)((2 * p) >> logN);
    const TowerMod m = mods[tower];
    const size_t item = (size_t)T << logN, dstride = batch * item;
    constexpr int R = NP > 0 ? NP : 1;
    ulonglong2 b[R], a[R];
    if (NP > 0) {
#pragma unroll
        for (int j = 0; j < R; j++) {
            b[j] = *reinterpret_cast<const ulonglong2*>(kb + (size_t)j * item + 2 * p);
            a[j] = *reinterpret_cast<const ulonglong2*>(ka + (size_t)j * item + 2 * p);
        }
    }
    for (size_t bi = blockIdx.y; bi < batch; bi += gridDim.y) {
        const u64* d = digits + bi * item + 2 * p;
        ulonglong2 s0 = make_ulonglong2(0, 0), s1 = make_ulonglong2(0, 0);
        if (NP > 0) {
            ulonglong2 x[R];
#pragma unroll
            for (int j = 0; j < R; j++) x[j] = *reinterpret_cast<const ulonglong2*>(d + (size_t)j * dstride);
#pragma unroll
            for (int j = 0; j < R; j++) {
                s0.x = addmod(s0.x, mulmod_barrett(x[j].x, b[j].x, m), m.q);
                s0.y = addmod(s0.y, mulmod_barrett(x[j].y, b[j].y, m), m.q);
                s1.x = addmod(s1.x, mulmod_barrett(x[j].x, a[j].x, m), m.q);
                s1.y = addmod(s1.y, mulmod_barrett(x[j].y, a[j].y, m), m.q);
            }
        } else {                                   // any number of digits: keys re-read (they stay in L2)
            for (u32 j = 0; j < np; j++) {
                const ulonglong2 x = *reinterpret_cast<const ulonglong2*>(d + (size_t)j * dstride);
                const ulonglong2 kbj = *reinterpret_cast<const ulonglong2*>(kb + (size_t)j * item + 2 * p);
                const ulonglong2 kaj = *reinterpret_cast<const ulonglong2*>(ka + (size_t)j * item + 2 * p);
                s0.x = addmod(s0.x, mulmod_barrett(x.x, kbj.x, m), m.q);
                s0.y = addmod(s0.y, mulmod_barrett(x.y, kbj.y, m), m.q);
                s1.x = addmod(s1.x, mulmod_barrett(x.x, kaj.x, m), m.q);
                s1.y = addmod(s1.y, mulmod_barrett(x.y, kaj.y, m), m.q);
            }
        }
        *reinterpret_cast<ulonglong2*>(o0 + bi * item + 2 * p) = s0;
        *reinterpret_cast<ulonglong2*>(o1 + bi * item + 2 * p) = s1;
    }
}

cudaError_t launch_ks_inner(const RingDev& r, uint64_t* o0, uint64_t* o1, const uint64_t* digits, const uint64_t* kb,
                            const uint64_t* ka, uint32_t np, size_t batch, cudaStream_t st) {
    if (batch == 0 || np == 0) return cudaSuccess;
    if (r.N < 2) return cudaErrorInvalidValue;
    if ((((uintptr_t)o0 | (uintptr_t)o1 | (uintptr_t)digits | (uintptr_t)kb | (uintptr_t)ka) & 15)) return cudaErrorMisalignedAddress;
    const size_t pairs = ((size_t)r.T << r.logN) >> 1;
    // enough CTAs along the batch to fill the machine, few enough that a thread amortises its key loads
    unsigned gx = (unsigned)((pairs + KS_THREADS - 1) / KS_THREADS);
    unsigned gy = (unsigned)std::min<size_t>(batch, std::max<size_t>(1, (148 * 16 + gx - 1) / gx));
    dim3 grid(gx, gy);
    u64 *p0 = reinterpret_cast<u64*>(o0), *p1 = reinterpret_cast<u64*>(o1);
    const u64 *pd = reinterpret_cast<const u64*>(digits), *pb = reinterpret_cast<const u64*>(kb), *pa = reinterpret_cast<const u64*>(ka);
#define KS_CASE(K) case K: ks_inner_kernel<K><<<grid, KS_THREADS, 0, st>>>(p0, p1, pd, pb, pa, np, batch, r.logN, r.T, r.mods); break;
    switch (np) {
        KS_CASE(1) KS_CASE(2) KS_CASE(3) KS_CASE(4) KS_CASE(5) KS_CASE(6)
        default: ks_inner_kernel<0><<<grid, KS_THREADS, 0, st>>>(p0, p1, pd, pb, pa, np, batch, r.logN, r.T, r.mods); break;
    }
#undef KS_CASE
    kernel_launch_count_add(1);
    return cudaGetLastError();
}

// EvalKeySwitchPrecomputeCore's "own towers" (keyswitch-hybrid.cpp:410-412): tower t of c[b] belongs to digit t / alpha and is
// stored unchanged at tower t of that digit's poly over Ql u P.  One launch for all digits (16 B per thread and step).
__global__ void __launch_bounds__(256) ks_own_towers_kernel(u64* __restrict__ digits, const u64* __restrict__ c, u32 size_ql, u32 T,
                                                            u32 alpha, u32 logN, size_t batch) {
    const size_t half = (size_t)1 << (logN - 1), total = batch * size_ql * half;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        const size_t n2 = i & (half - 1), bt = i >> (logN - 1);
        const size_t t = bt % size_ql, b = bt / size_ql, j = t / alpha;
        const ulonglong2 v = *reinterpret_cast<const ulonglong2*>(c + ((b * size_ql + t) << logN) + 2 * n2);
        *reinterpret_cast<ulonglong2*>(digits + ((((j * batch + b) * T) + t) << logN) + 2 * n2) = v;
    }
}

cudaError_t launch_ks_own_towers(uint64_t* digits, const uint64_t* c, uint32_t size_ql, uint32_t T, uint32_t alpha, uint32_t logN,
                                 size_t batch, cudaStream_t st) {
    if (batch == 0 || size_ql == 0) return cudaSuccess;
    if (logN < 1 || alpha == 0) return cudaErrorInvalidValue;
    if (((uintptr_t)digits | (uintptr_t)c) & 15) return cudaErrorMisalignedAddress;
    const size_t total = (batch * size_ql) << (logN - 1);
    const unsigned grid = (unsigned)std::min<size_t>((total + 255) / 256, 148 * 16);
    ks_own_towers_kernel<<<grid, 256, 0, st>>>(reinterpret_cast<u64*>(digits), reinterpret_cast<const u64*>(c), size_ql, T, alpha, logN, batch);
    kernel_launch_count_add(1);
    return cudaGetLastError();
}

}  // namespace fhe
