// decrypt.cu — BFV decryption's scaling step on the device: DCRTPoly over Q -> NativePoly modulo the plaintext modulus.
//
// DCRTPolyImpl::ScaleAndRound(t, tQHatInvModqDivqModt, ...Precon, tQHatInvModqBDivqModt, ...Precon,
//                             tQHatInvModqDivqFrac, tQHatInvModqBDivqFrac)        dcrtpoly-impl.h:1537-1816
// computes round(t/Q * x) mod t per coefficient from the residues x_i:
//     sum_i x_i * [t*QHatInv_i/q_i]_t   (integers, modulo t)   +   round(sum_i x_i * frac(t*QHatInv_i/q_i))   (IEEE double)
// with EIGHT branches chosen from the bit lengths of q_0, t and the tower count:
//   t a power of two or not            (mask, or a quotient estimated in double: :1672-1674)
//   qMSB + sizeQMSB < 52 or not        (x_i used whole, or split in two halves of qMSB/2 bits with a second table pair)
//   the integer sum fits 63/62/52 bits (plain wrapping MulEqFast / AddEqFast) or not (ModMulFastConstEq modulo t)
// Here one kernel with the three choices as template flags follows each branch statement for statement; the double
// sums use __dmul_rn / __dadd_rn / __dsub_rn in the reference's order (the CPU path is compiled without FMA contraction),
// so every coefficient equals the reference's bit for bit.  One thread per coefficient, towers in a loop (coalesced
// 8-byte loads per tower), 8 B per tower in and 8 B out: HBM-bound.
#include "modarith.cuh"
#include "kernels.h"

namespace fhe {

template <bool POW2, bool SPLIT, bool REDUCE>
__global__ void __launch_bounds__(256)
decrypt_scale_kernel(u64* __restrict__ dst, const u64* __restrict__ src, size_t items, u32 logN, DecryptDev dd) {
    extern __shared__ __align__(16) unsigned char smraw[];
    ulonglong4* sK = reinterpret_cast<ulonglong4*>(smraw);                   // [T] {modt, modt', bmodt, bmodt'}
    double2* sF = reinterpret_cast<double2*>(sK + dd.T);                     // [T] {frac, bfrac}
    for (u32 i = threadIdx.x; i < dd.T; i += 256) { sK[i] = dd.k[i]; sF[i] = dd.f[i]; }
    __syncthreads();
    const u64 t = dd.t;
    for (size_t e = (size_t)blockIdx.x * 256 + threadIdx.x; e < items; e += (size_t)gridDim.x * 256) {
        const size_t b = e >> logN;
        const u32 n = (u32)e & ((1u << logN) - 1);
        const u64* ip = src + ((b * dd.T) << logN) + n;
        double fs = POW2 ? 0.5 : 0.0;
        u64 is = 0;
        u64 x = ip[0];
        for (u32 i = 0; i < dd.T; i++) {
            const u64 xn = (i + 1 < dd.T) ? ip[(size_t)(i + 1) << logN] : 0ull;
            const ulonglong4 k = sK[i];
            const double2 f = sF[i];
            if (!SPLIT) {
                fs = __dadd_rn(fs, __dmul_rn(__ull2double_rn(x), f.x));
                is += REDUCE ? shoup(x, k.x, k.y, t) : x * k.x;
            } else {
                const u64 hi = x >> dd.hbits, lo = x - (hi << dd.hbits);
                fs = __dadd_rn(fs, __dmul_rn(__ull2double_rn(lo), f.x));
                fs = __dadd_rn(fs, __dmul_rn(__ull2double_rn(hi), f.y));
                is += REDUCE ? shoup(lo, k.x, k.y, t) : lo * k.x;
                is += REDUCE ? shoup(hi, k.z, k.w, t) : hi * k.z;
            }
            x = xn;
        }
        u64 r;
        if (POW2) {
            is += __double2ull_rz(fs);
            r = is & (t - 1);
        } else {                                   // :1672-1676: quotient in double, then rounding
            fs = __dadd_rn(fs, __ull2double_rn(is));
            const u64 quot = __double2ull_rz(__dmul_rn(fs, dd.tinv));
            fs = __dsub_rn(fs, __dmul_rn(dd.td, __ull2double_rn(quot)));
            r = __double2ull_rz(__dadd_rn(fs, 0.5));
        }
        dst[e] = r;
    }
}

cudaError_t launch_decrypt_scale(const DecryptDev& dd, uint64_t* dst, const uint64_t* src, size_t batch, uint32_t logN,
                                 cudaStream_t st) {
    const size_t items = batch << logN;
    if (items == 0) return cudaSuccess;
    const size_t smem = (size_t)dd.T * (sizeof(ulonglong4) + sizeof(double2));
    if (smem > 48 * 1024) return cudaErrorInvalidValue;
    size_t blocks = (items + 255) / 256;
    if (blocks > 148 * 32) blocks = 148 * 32;
    u64* d = reinterpret_cast<u64*>(dst);
    const u64* s = reinterpret_cast<const u64*>(src);
    const int sel = (dd.pow2 ? 4 : 0) | (dd.split ? 2 : 0) | (dd.reduce ? 1 : 0);
#define DS_CASE(V, P, S, R) case V: decrypt_scale_kernel<P, S, R><<<(unsigned)blocks, 256, smem, st>>>(d, s, items, logN, dd); break;
    switch (sel) {
        DS_CASE(0, false, false, false) DS_CASE(1, false, false, true) DS_CASE(2, false, true, false) DS_CASE(3, false, true, true)
        DS_CASE(4, true, false, false) DS_CASE(5, true, false, true) DS_CASE(6, true, true, false) DS_CASE(7, true, true, true)
    }
#undef DS_CASE
    kernel_launch_count_add(1);
    return cudaGetLastError();
}

}  // namespace fhe
