// modops.cu — the "mod-ops" kernels of the PimData surface for sm_100a.
//
// Device replacement for what the reference's DPU binary mod-ops.c (13 mod_kernel entry points,
// src/core/pim/dpu/element-wise/mod-ops.c:25-39; "not yet connected" on its host side,
// pim/PimData.h:171) was meant to compute.  Semantics follow OpenFHE's CPU definitions, not the DPU
// C code (which has bugs):
//   SwitchModulus / Mod / ModEq     NativeVectorT  mubintvecnat.cpp:111-195
//   ModByTwo(Eq)                    mubintvecnat.cpp:370-385
//   ModExp(Eq)                      mubintvecnat.cpp:387-410, NativeIntegerT::ModExp ubintnat.h:1507-1530
//   ModInverse(Eq)                  mubintvecnat.h:546-566,  NativeIntegerT::ModInverse ubintnat.h:1565-1592
//   MultiplyAndRound(Eq)            mubintvecnat.cpp:423-455, ubintnat.h:536-541 (IEEE double, no FMA)
//   DivideAndRound(Eq)              mubintvecnat.cpp:457-487, ubintnat.h:596-606
// All are one-pass element-wise kernels: 16 B of HBM traffic per coefficient (read src, write dst).
#include "modarith.cuh"
#include "kernels.h"

namespace fhe {

typedef unsigned __int128 u128;

__device__ __forceinline__ u64 mulmod_any(u64 a, u64 b, u64 m, bool barrett, const TowerMod& tm) {
    if (barrett) return mulmod_barrett(a, b, tm);
    return (u64)(((u128)a * b) % m);
}

__device__ __forceinline__ u64 mulround(u64 v, double pd, double qd) {
    double ratio = __ddiv_rn(__ull2double_rn(v), qd);
    return __double2ull_rz(__dadd_rn(__dmul_rn(pd, ratio), 0.5));
}

__device__ __forceinline__ u64 divround(u64 v, u64 q) {
    u64 ans = v / q, rem = v - ans * q;
    return rem > (q >> 1) ? ans + 1 : ans;
}

// kind: enum mod_kernel of pim/common.h.  err: set to 1 when an element has no inverse.
__global__ void __launch_bounds__(256)
modop_kernel(int kind, u64* dst, const u64* src, size_t n, u64 mod, u64 p, u64 q,
             TowerMod tm, int barrett, int* __restrict__ err) {
    const u64 half = mod >> 1;
    const double pd = __ull2double_rn(p), qd = __ull2double_rn(q);
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        u64 v = src[i], r = 0;
        switch (kind) {
            case 0: case 1: case 2:                       // SWITCHMODULUS, MOD_NORMAL, MODEQ (p = new modulus)
                if ((kind != 0) && p == 2) { r = 1ull & (v ^ (u64)(v > half)); break; }
                if (p > mod) { r = v > half ? v + (p - mod) : v; }
                else {
                    if (v > half) v += p - (mod % p);
                    r = v >= p ? v % p : v;
                }
                break;
            case 3: case 4:                               // MODBYTWO(EQ)
                r = 1ull & (v ^ (u64)(v > half));
                break;
            case 5: case 6: {                             // MODEXP(EQ), exponent already reduced by the host
                u64 t = v, e = p;
                if (t >= mod) t %= mod;
                r = 1;
                if (e & 1) { r = t; if (r >= mod) r %= mod; }
                while (e >>= 1) {
                    t = mulmod_any(t, t, mod, barrett, tm);
                    if (e & 1) r = mulmod_any(r, t, mod, barrett, tm);
                }
                break;
            }
            case 7: case 8: {                             // MODINVERSE(EQ)
                long long modulus = (long long)mod, a = (long long)(v % mod);
                if (a == 0) { atomicExch(err, 1); r = 0; break; }
                if (modulus == 1) { r = 0; break; }
                long long y = 0, x = 1;
                while (a > 1) {
                    long long t = modulus, qq = a / t;
                    modulus = a - qq * t; a = t; t = y;
                    y = x - qq * y; x = t;
                }
                if (x < 0) x += (long long)mod;
                r = (u64)x;
                break;
            }
            case 9: case 10:                              // MULTIPLYANDROUND(EQ)
                if (v > half) r = mod - mulround(mod - v, pd, qd);
                else r = mulround(v, pd, qd) % mod;
                break;
            case 11: case 12:                             // DIVIDEANDROUND(EQ): divisor in p
                if (v > half) r = mod - divround(mod - v, p);
                else r = divround(v, p);
                break;
        }
        dst[i] = r;
    }
}

cudaError_t launch_modop(int kind, uint64_t* dst, const uint64_t* src, size_t n, uint64_t mod, uint64_t p,
                         uint64_t q, int* err, cudaStream_t st) {
    if (n == 0) return cudaSuccess;
    TowerMod tm{};
    int barrett = (mod >= 4 && mod < (1ull << 61)) ? 1 : 0;
    if (barrett) {
        unsigned msb = 64u - (unsigned)__builtin_clzll(mod);
        tm.q = mod;
        tm.mu = (uint64_t)(((unsigned __int128)1 << (2 * msb + 3)) / mod);
        tm.n = msb - 2;
        tm.s = msb + 5;
        towermod_set_reduce(tm);
    }
    size_t blocks = (n + 255) / 256;
    if (blocks > 148 * 32) blocks = 148 * 32;
    modop_kernel<<<(unsigned)blocks, 256, 0, st>>>(kind, reinterpret_cast<u64*>(dst), reinterpret_cast<const u64*>(src),
                                                   n, mod, p, q, tm, barrett, err);
    kernel_launch_count_add(1);
    return cudaGetLastError();
}

}  // namespace fhe
