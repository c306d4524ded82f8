// kernels.h — host-side launch interface of the sm_100a kernels (internal to libfhe_b200.so).
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>
#include <vector>

namespace fhe {

struct TowerMod;

// Device-resident description of one ring: (N, T, q_i, psi_i) plus the tables PreCompute builds
// in the reference (transformnat-impl.h:708-763), already uploaded to HBM.
struct RingDev {
    uint32_t N, logN, T;
    const TowerMod* mods;        // [T]
    const ulonglong2* twF;       // [T][N]  {psi^bitrev(i), companion}
    const ulonglong2* twI;       // [T][N]  {psi^-bitrev(i), companion}
    const ulonglong2* invc;      // [T][2]  {N^-1, companion}, {WI[1]*N^-1, companion}
};

enum EwOp { EW_ADD = 0, EW_SUB = 1, EW_MUL = 2 };

// dst = a (op) b over [batch][T][N]; dst may alias a or b.
cudaError_t launch_elementwise(const RingDev& r, int op, uint64_t* dst, const uint64_t* a,
                               const uint64_t* b, size_t batch, cudaStream_t st);
// dst = a (op) scalars[tower] over [batch][T][N] (op EW_ADD / EW_SUB; scalars on the host, reduced; T <= EW_SCALARS_MAX)
#define EW_SCALARS_MAX 64
cudaError_t launch_tower_scalars(const RingDev& r, int op, uint64_t* dst, const uint64_t* a, const uint64_t* scalars, size_t batch,
                                 cudaStream_t st);
// c0 = a0*b0, c1 = a1*b0 + a0*b1, c2 = a1*b1 over [batch][T][N]; outputs may alias inputs element-for-element.
cudaError_t launch_tensor_product(const RingDev& r, uint64_t* c0, uint64_t* c1, uint64_t* c2,
                                  const uint64_t* a0, const uint64_t* a1, const uint64_t* b0, const uint64_t* b1,
                                  size_t batch, cudaStream_t st);
// dst[i] = a[i] (op) b[i] mod q over a flat vector of n words (one modulus), any n.
cudaError_t launch_vec(int op, uint64_t* dst, const uint64_t* a, const uint64_t* b, size_t n,
                       uint64_t q, cudaStream_t st);
// dst[i] = a[i] (op) s mod q
cudaError_t launch_vec_scalar(int op, uint64_t* dst, const uint64_t* a, uint64_t s, size_t n,
                              uint64_t q, cudaStream_t st);
// Forward / inverse negacyclic NTT of every tower of every batch item, result in data.
// algo: 0 = auto, 1 = force the single-CTA shared-memory kernel (N <= 4096), 2 = one-tile-per-CTA kernels (ntt.cu) even
// where the persistent ones (ntt_ws.cu) would be chosen, 3 = the persistent ones wherever they cover the ring.
// src: buffer the transform reads (nullptr = data, in place).  outT / inT: towers per batch item in the memory of data /
// src (0 = r.T): the r.T towers transformed are the first ones of each item, so a pointer offset by t0 * N transforms
// towers t0 .. t0 + r.T - 1 of a wider poly in place.
cudaError_t launch_ntt(const RingDev& r, uint64_t* data, size_t batch, bool inverse,
                       cudaStream_t st, int algo, const uint64_t* src = nullptr, uint32_t outT = 0, uint32_t inT = 0);
// The warp-specialised persistent form of the same transform (ntt_ws.cu; N = 2^14..2^16, large batches); launch_ntt
// dispatches to it when ntt_ws_applies().
bool ntt_ws_applies(const RingDev& r, size_t batch);
cudaError_t launch_ntt_ws(const RingDev& r, uint64_t* data, size_t batch, bool inverse, cudaStream_t st,
                          const uint64_t* in, uint32_t outT, uint32_t inT);
// dst = automorphism X -> X^k of src over [batch][T][N] (k odd, dst != src).  format 0 EVALUATION, 1 COEFFICIENT.
cudaError_t launch_automorphism(const RingDev& r, uint64_t* dst, const uint64_t* src, size_t batch, uint32_t k,
                                int format, cudaStream_t st);
// dst[j] = src[map[j]] per tower (map: N uint32 in HBM), the precomputed-table form.
cudaError_t launch_automorphism_map(const RingDev& r, uint64_t* dst, const uint64_t* src, size_t batch,
                                    const uint32_t* map, cudaStream_t st);
// ---- RNS basis conversion (basis.cu) ----
struct BasisTarget {             // one target modulus p_j of ApproxSwitchCRTBasis
    uint64_t p;
    uint64_t c64, c64p;          // 2^64 mod p and its Shoup companion floor(c64 * 2^64 / p)   (128-bit sum -> residue, basis.cu)
    uint64_t pinv;               // floor(2^64 / p)
    uint64_t post, postp;        // optional multiplier applied to the reduced value (0 = none) and its Shoup companion
    uint64_t pad[2];
};
inline void basis_target_set(BasisTarget& g, uint64_t p) {
    g.p = p;
    g.c64 = (uint64_t)((((unsigned __int128)1) << 64) % p);
    g.c64p = (uint64_t)((((unsigned __int128)g.c64) << 64) / p);
    g.pinv = (uint64_t)((((unsigned __int128)1) << 64) / p);
}
struct BasisDev {
    uint32_t sq, sp;
    const uint64_t* q;           // [sq] source moduli
    const ulonglong2* pre;       // [sq] {[(Q/q_i)^-1]_{q_i} (times an optional scalar), Shoup companion}
    const uint64_t* hat;         // [sq][sp] [Q/q_i]_{p_j}
    const BasisTarget* tgt;      // [sp]
    const double* qinv;          // [sq] 1./double(q_i)                      (exact form only)
    const uint64_t* alphaq;      // [sq+1][sp] [alpha*Q]_{p_j}               (exact form only)
    uint32_t narrow;             // 1: every q_i and every hat entry is below 2^60 -> carry-free 30-bit split accumulation
    // tensor-core form (basis.cu, basis_switch_tc_kernel): the table expanded into 8-bit limbs, Toeplitz per target,
    // in the shared-memory core-matrix layout, one block per group of tc_jg targets; null: integer kernels only
    const uint8_t* tc_b;         // [tc_groups][tc_kpad / 16][16 * tc_jg][16]
    uint32_t tc_groups, tc_jg, tc_kpad;
};
// host side: fills `out` ([groups][kpad/16][16*jg][16] bytes) from hat [sq][sp]; returns false when the shape is not covered
bool basis_tc_build(const uint64_t* hat, uint32_t sq, uint32_t sp, uint32_t& groups, uint32_t& jg, uint32_t& kpad,
                    std::vector<uint8_t>& out);
struct ScaleDev {                // DCRTPolyImpl::ScaleAndRound(paramsOutput, ...)
    uint32_t sI, sO, in0, out0;  // input / output tower counts and their first tower inside [sI+sO]
    const double* frac;          // [sI]
    const uint64_t* modo;        // [sO][sI+1]
    const BasisTarget* tgt;      // [sO] output moduli with their reduction constants
    uint32_t narrow;             // 1: input moduli and modo entries below 2^60 (see BasisDev::narrow)
};
// dst towers dst0.. of [batch][dstT][N] = ApproxSwitchCRTBasis of src towers src0..src0+sq-1 of [batch][srcT][N]
cudaError_t launch_basis_switch(const BasisDev& bd, uint64_t* dst, uint32_t dstT, uint32_t dst0, const uint64_t* src,
                                uint32_t srcT, uint32_t src0, size_t batch, uint32_t logN, bool exact, cudaStream_t st);
// dst [batch][sO][N] = ScaleAndRound of src [batch][sI+sO][N]
cudaError_t launch_scale_round(const ScaleDev& sd, uint64_t* dst, const uint64_t* src, size_t batch, uint32_t logN,
                               cudaStream_t st);
// mode 0: out = (x - v) * c; 1: out = x * c + v; 2: out = (x + v * t) * c, per tower k[i] = {c, c', t, t'};
// x from [batch][xT][N] (towers 0..l-1), v and out over [batch][l][N]
cudaError_t launch_tower_affine(int mode, uint64_t* out, const uint64_t* x, const uint64_t* v, size_t batch, uint32_t logN,
                                uint32_t l, uint32_t xT, const TowerMod* mods, const ulonglong4* k, cudaStream_t st);
// tmp[b][i][n] = SwitchModulus(last[b][n], ql -> q_i) (* c_i), see basis.cu
cudaError_t launch_tower_scale(int mode, uint64_t* out, const uint64_t* x, size_t batch, uint32_t logN, uint32_t sl,
                               uint32_t T, const TowerMod* mods, const ulonglong4* k, cudaStream_t st);
cudaError_t launch_rescale_spread(uint64_t* tmp, const uint64_t* last, size_t batch, uint32_t logN, uint32_t l, uint64_t ql,
                                  const ulonglong4* sw, const ulonglong4* k, ulonglong2 pre, int bgv, cudaStream_t st);
// DCRTPolyImpl::CRTDecompose(0): coef [batch][T][N] -> out [T][batch][T][N] in COEFFICIENT form (basis.cu)
cudaError_t launch_crt_spread(uint64_t* out, const uint64_t* coef, size_t batch, uint32_t logN, uint32_t T, const ulonglong4* sw,
                              cudaStream_t st);
// EvalFastKeySwitchCoreExt (keyswitch-hybrid.cpp:442-480): o0 = sum_j digits[j] * kb[j], o1 = sum_j digits[j] * ka[j];
// digits [np][batch][T][N], keys [np][T][N], outputs [batch][T][N] over the ring r
cudaError_t launch_ks_inner(const RingDev& r, uint64_t* o0, uint64_t* o1, const uint64_t* digits, const uint64_t* kb,
                            const uint64_t* ka, uint32_t np, size_t batch, cudaStream_t st);
// each digit's own stored towers copied where they belong in its poly over Ql u P (keyswitch.cu)
cudaError_t launch_ks_own_towers(uint64_t* digits, const uint64_t* c, uint32_t size_ql, uint32_t T, uint32_t alpha, uint32_t logN,
                                 size_t batch, cudaStream_t st);
// DCRTPolyImpl::ScaleAndRound(t, ...) -> NativePoly (dcrtpoly-impl.h:1537-1816), BFV decryption
struct DecryptDev {
    uint32_t T;                  // towers of Q
    uint64_t t;                  // plaintext modulus
    uint32_t pow2, split, reduce, hbits;     // the reference's branch: t power of two; x split at hbits = qMSB/2; products reduced mod t
    const ulonglong4* k;         // [T] {tQHatInvModqDivqModt, its Shoup companion, tQHatInvModqBDivqModt, its companion}
    const double2* f;            // [T] {tQHatInvModqDivqFrac, tQHatInvModqBDivqFrac}
    double td, tinv;             // double(t), 1. / td
};
// dst [batch][N] = ScaleAndRound of src [batch][T][N] (COEFFICIENT)
cudaError_t launch_decrypt_scale(const DecryptDev& dd, uint64_t* dst, const uint64_t* src, size_t batch, uint32_t logN,
                                 cudaStream_t st);
// BEHZ (dcrtpoly-impl.h:2005-2050, 2069-2420): per-tower constants of the corrections that follow the fast base
// conversions (behz.cu).  nq towers of Q, nb towers of Bsk = B u {msk} (msk last).
struct BehzDev {
    uint32_t nq, nb;
    const ulonglong4* qk;        // [nq] {q_i, [mtilde*(Q/q_i)^-1]_{q_i}, its Shoup companion, [Q/q_i]_{mtilde}}
    const ulonglong4* bk;        // [nb] {[Q]_{bsk_j}, companion, [mtilde^-1]_{bsk_j}, companion}
    const uint64_t* bsk;         // [nb] bsk_j
    const ulonglong4* fk;        // [nq+nb] {c, companion, modulus, 0}: c = [t*(Q/q_i)^-1]_{q_i} over Q, [t*Q^-1]_{bsk_j} over Bsk
    const ulonglong4* sk;        // [nq] {[B]_{q_i}, companion, q_i, 0}
    const ulonglong4* dk;        // [nq] {[t*gamma*(Q/q_i)^-1]_{q_i}, companion, [-q_i^-1]_{t*gamma}, companion}
    uint64_t neg_qinv_mtilde;    // [-Q^-1]_{mtilde}
    uint64_t msk, binv_msk, binv_mskp;      // msk, [B^-1]_{msk} and its companion
    uint64_t tgamma;             // t * 2^26
};
// x [batch][nq+nb][N]: towers nq.. = FastBaseConv(|x*mtilde|_q, q, Bsk) on entry, Montgomery-reduced on exit; coef
// [batch][nq][N] = the input in COEFFICIENT format (must not overlap x)
cudaError_t launch_behz_montgomery(const BehzDev& d, uint64_t* x, const uint64_t* coef, size_t batch, uint32_t logN, cudaStream_t st);
// in place on x [batch][nq+nb][N] with conv [batch][nb][N] = FastBaseConv(|t*x|_q, q, Bsk); update_q: also leave the towers of Q
// multiplied by [t*(Q/q_i)^-1]_{q_i} as the reference does
cudaError_t launch_behz_floor(const BehzDev& d, uint64_t* x, const uint64_t* conv, size_t batch, uint32_t logN, bool update_q,
                              cudaStream_t st);
// dst [batch][nq][N] from conv [batch][nq+1][N] (FastBaseConv(x, B, q) and FastBaseConv(x, B, msk)) and the msk tower of the
// input (xsk, xT towers per polynomial)
cudaError_t launch_behz_sk(const BehzDev& d, uint64_t* dst, const uint64_t* conv, const uint64_t* xsk, uint32_t xT, size_t batch,
                           uint32_t logN, cudaStream_t st);
// dst [batch][N] = ScaleAndRound(moduliQ, t, tgamma, ...) of src [batch][nq][N] (COEFFICIENT)
cudaError_t launch_behz_decrypt(const BehzDev& d, uint64_t* dst, const uint64_t* src, size_t batch, uint32_t logN, cudaStream_t st);
// Counter-based uniform fill in [0, q_t) per tower (same definition as the test oracle's generator).
cudaError_t launch_fill_uniform(const RingDev& r, uint64_t* dst, size_t batch, uint64_t seed,
                                cudaStream_t st);
// mod-ops of the PimData surface; kind = enum mod_kernel (pim/common.h).  *err (device int) is set to 1
// when ModInverse meets an element without an inverse.
cudaError_t launch_modop(int kind, uint64_t* dst, const uint64_t* src, size_t n, uint64_t mod, uint64_t p,
                         uint64_t q, int* err, cudaStream_t st);
// out2[0] = sum of the n words mod 2^64, out2[1] = their xor (out2 is device memory, 16 bytes).
cudaError_t launch_checksum(const uint64_t* x, size_t n, uint64_t* out2, cudaStream_t st);
// Integer-pipe probe: every thread runs `iters` dependent-chain-free 64x64->128 multiply-adds.
cudaError_t launch_imad_probe(uint64_t* out, int blocks, int threads, int iters, cudaStream_t st);

int kernel_launch_count();       // number of kernels launched by this library since load
void kernel_launch_count_add(int n);

}  // namespace fhe
