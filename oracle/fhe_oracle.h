/*
 * fhe_oracle.h — CPU restatement of the reference's hot path (TEST INFRASTRUCTURE ONLY).
 *
 * This is the parity oracle: plain C (gcc, unsigned __int128) following OpenFHE's
 * native-integer CPU path function by function.  It is NOT part of the product:
 * only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs may load it.  The product (libfhe_b200.so) never links or calls it.
 *
 * Parity status: PINNED.  tests/test_oracle_golden.py replays the reference's own
 * known-answer vectors (UnitTestTransform.cpp:59-98, UnitTestDCRTElements.cpp:285-406,
 * UnitTestNTT.cpp:53-133, SURVEY §8 N2 parameters) and, in this container,
 * oracle/_ref/libopenfhe_ref.so (the reference's own sources compiled by
 * oracle/Makefile) is compared against it on random 60-bit inputs
 * (tests/test_oracle_vs_ref.py; vectors committed under tests/golden/).
 *
 * All citations are relative to /root/reference.
 */
#ifndef FHE_ORACLE_H
#define FHE_ORACLE_H
#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- scalar primitives (src/core/include/math/hal/intnat/ubintnat.h) ---- */
uint32_t orc_msb(uint64_t x);                                   /* lbcrypto::GetMSB: bit length, 0 for 0 */
uint64_t orc_compute_mu(uint64_t q);                            /* ubintnat.h:650-656 */
uint64_t orc_modadd_fast(uint64_t a, uint64_t b, uint64_t q);   /* ubintnat.h:760-767 */
uint64_t orc_modsub_fast(uint64_t a, uint64_t b, uint64_t q);   /* ubintnat.h:934-938 */
uint64_t orc_modmul_barrett(uint64_t a, uint64_t b, uint64_t q, uint64_t mu); /* ubintnat.h:1398-1413 */
uint64_t orc_prep_mulconst(uint64_t w, uint64_t q);             /* ubintnat.h:1447-1455 */
uint64_t orc_modmul_const(uint64_t a, uint64_t w, uint64_t q, uint64_t wprecon); /* ubintnat.h:1491-1497 */
uint64_t orc_modexp(uint64_t b, uint64_t e, uint64_t q);
uint64_t orc_modinv(uint64_t a, uint64_t q);

/* ---- parameter generation (src/core/include/math/nbtheory-impl.h) ---- */
int      orc_is_prime(uint64_t p);                              /* stands in for MillerRabinPrimalityTest */
uint64_t orc_first_prime(uint32_t nbits, uint64_t m);           /* nbtheory-impl.h:334-359 */
uint64_t orc_next_prime(uint64_t q, uint64_t m);                /* nbtheory-impl.h:361-369 */
uint64_t orc_previous_prime(uint64_t q, uint64_t m);            /* nbtheory-impl.h:372-379 */
uint64_t orc_root_of_unity(uint64_t m, uint64_t q);             /* nbtheory-impl.h:183-231 (minimum primitive m-th root) */
/* benchmark/src/poly-benchmark-16k.cpp:82-101: descending PreviousPrime chain + roots */
void     orc_gen_dcrt_params(uint32_t nbits, uint32_t N, uint32_t T, uint64_t* q, uint64_t* psi);

/* ---- vector ops, one tower (mubintvecnat.h:426-432, 501-513; mubintvecnat.cpp:245-367) ---- */
void orc_vec_modadd(uint64_t* dst, const uint64_t* a, const uint64_t* b, size_t n, uint64_t q);
void orc_vec_modsub(uint64_t* dst, const uint64_t* a, const uint64_t* b, size_t n, uint64_t q);
void orc_vec_modmul(uint64_t* dst, const uint64_t* a, const uint64_t* b, size_t n, uint64_t q);
void orc_vec_modadd_scalar(uint64_t* dst, const uint64_t* a, uint64_t s, size_t n, uint64_t q);
void orc_vec_modsub_scalar(uint64_t* dst, const uint64_t* a, uint64_t s, size_t n, uint64_t q);
void orc_vec_modmul_scalar(uint64_t* dst, const uint64_t* a, uint64_t s, size_t n, uint64_t q);

/* mod-ops of the PimData surface; kind = enum mod_kernel (pim/common.h:32-46); 0 ok, 1 = the
 * reference would throw (no inverse / zero divisor) */
int orc_vec_modop(int kind, uint64_t* dst, const uint64_t* src, size_t n, uint64_t mod, uint64_t p, uint64_t q);

/* ---- twiddle tables (transformnat-impl.h:708-763) ----
 * W[bitrev(i)] = psi^i, WI[bitrev(i)] = psi^-i (i < N), Wp/WIp their Shoup companions,
 * *ninv = N^-1 mod q and *ninvp its companion (TableCOI[log2 N]). */
void orc_precompute(uint64_t psi, uint32_t N, uint64_t q,
                    uint64_t* W, uint64_t* Wp, uint64_t* WI, uint64_t* WIp,
                    uint64_t* ninv, uint64_t* ninvp);

/* ---- NTT kernels, one tower, in place (transformnat-impl.h:300-354, 492-552) ---- */
void orc_ntt_fwd(uint64_t* x, uint32_t N, uint64_t q, const uint64_t* W, const uint64_t* Wp);
void orc_ntt_inv(uint64_t* x, uint32_t N, uint64_t q, const uint64_t* WI, const uint64_t* WIp,
                 uint64_t ninv, uint64_t ninvp);

/* ---- DCRTPoly level: batch x towers, layout [B][T][N] (dcrtpoly.h:142-200,
 * dcrtpoly-impl.h:410-416, 2518-2524).  op: 0 add, 1 sub, 2 mul.
 * nthreads: OpenMP threads over the (batch, tower) units; 0 = omp default. ---- */
typedef struct orc_ctx orc_ctx;
orc_ctx* orc_ctx_create(uint32_t N, uint32_t T, const uint64_t* q, const uint64_t* psi);
void     orc_ctx_destroy(orc_ctx* c);
void     orc_dcrt_elementwise(const orc_ctx* c, int op, uint64_t* dst, const uint64_t* a,
                              const uint64_t* b, size_t B, int nthreads);
void     orc_dcrt_ntt(const orc_ctx* c, int inverse, uint64_t* x, size_t B, int nthreads);
/* BFV decryption: ScaleAndRound(t, ...) -> NativePoly (dcrtpoly-impl.h:1537-1816); tables of bfvrns-cryptoparameters.cpp:447-494 */
int      orc_decrypt_branch(const uint64_t* q, uint32_t T, uint64_t t);
void     orc_decrypt_tables(const uint64_t* q, uint32_t T, uint64_t t, uint64_t* modt, uint64_t* bmodt, double* frac, double* bfrac);
void     orc_scale_and_round_native(uint64_t* out, const uint64_t* in, size_t B, uint32_t N, const uint64_t* q, uint32_t T, uint64_t t,
                                    const uint64_t* modt, const uint64_t* bmodt, const double* frac, const double* bfrac);
/* hybrid key switching (keyswitch-hybrid.cpp:324-480); QlP = sizeQl towers of Q then P; digits of alpha towers */
uint32_t orc_keyswitch_digits(uint32_t sizeQl, uint32_t alpha);
void     orc_keyswitch_precompute(const orc_ctx* QlP, uint32_t sizeQl, uint32_t alpha, uint64_t* digits, const uint64_t* c, size_t B);
void     orc_keyswitch_inner(const orc_ctx* QlP, uint32_t np, uint64_t* c0, uint64_t* c1, const uint64_t* digits,
                             const uint64_t* evk_b, const uint64_t* evk_a, size_t B);
void     orc_keyswitch_hybrid(const orc_ctx* QlP, uint32_t sizeQl, uint32_t alpha, uint64_t* ct0, uint64_t* ct1, const uint64_t* c,
                              const uint64_t* evk_b, const uint64_t* evk_a, size_t B, uint64_t t);
/* BEHZ (dcrtpoly-impl.h:2005-2050, 2069-2420; tables of bfvrns-cryptoparameters.cpp:662-885).  QBsk = nq towers of Q
 * followed by Bsk = B u {msk}.  Tables by number: 0 mtildeQHatInvModq[nq], 1 QHatModbsk[nq][nb], 2 QHatModmtilde[nq],
 * 3 QModbsk[nb], 4 negQInvModmtilde, 5 mtildeInvModbsk[nb], 6 tQHatInvModq[nq], 7 qInvModbsk[nq][nb], 8 tQInvModbsk[nb],
 * 9 BHatInvModb[nb-1], 10 BHatModq[nb-1][nq], 11 BHatModmsk[nb-1], 12 BInvModmsk, 13 BModq[nq], 14 tgamma,
 * 15 negInvqModtgamma[nq], 16 tgammaQHatInvModq[nq] */
typedef struct orc_behz orc_behz;
orc_behz* orc_behz_create(const orc_ctx* QBsk, uint32_t nq, uint64_t t);
void     orc_behz_destroy(orc_behz* z);
size_t   orc_behz_table(const orc_behz* z, int which, uint64_t* out);
void     orc_behz_q_to_bsk_montgomery(const orc_behz* z, uint64_t* out_qbsk, const uint64_t* in_q, size_t B, int format);
void     orc_behz_fast_rns_floor_q(const orc_behz* z, uint64_t* x_qbsk, size_t B);
void     orc_behz_fast_base_conv_sk(const orc_behz* z, uint64_t* out_q, const uint64_t* in_qbsk, size_t B);
void     orc_behz_scale_and_round(const orc_behz* z, uint64_t* out, const uint64_t* in_q, size_t B);
void     orc_crt_decompose(const orc_ctx* c, uint64_t* out, const uint64_t* in, size_t B, int format);   /* dcrtpoly-impl.h:266-288 */
int      orc_max_threads(void);

/* ---- automorphism X -> X^k, k odd ----
 * orc_automorphism: one tower, PolyImpl::AutomorphismTransform(k) (poly-impl.h:312-365); format 0 =
 * EVALUATION (bit-reversed order), 1 = COEFFICIENT (a negated zero is stored as q, as the reference's
 * `q - value` does).  orc_precompute_automap: PrecomputeAutoMap (nbtheory2.cpp:274-286);
 * orc_automorphism_map: the two-argument form (poly-impl.h:368-380).  The dcrt form applies it to
 * [B][T][N].  Return 1 where the reference throws (k even). */
int  orc_automorphism(uint64_t* out, const uint64_t* in, uint32_t N, uint64_t q, uint32_t k, int format);
void orc_precompute_automap(uint32_t N, uint32_t k, uint32_t* map);
void orc_automorphism_map(uint64_t* out, const uint64_t* in, uint32_t N, const uint32_t* map);
int  orc_dcrt_automorphism(const orc_ctx* c, uint64_t* out, const uint64_t* in, size_t B, uint32_t k, int format);

/* ---- RNS basis conversion and scaling, layouts [B][towers][N] ----
 * Tables (plain modular arithmetic; equal to the BigInteger values of
 * src/pke/lib/schemerns/rns-cryptoparameters.cpp:190-213 and ckksrns-cryptoparameters.cpp:66-88):
 *   hat_inv[i] = (Q/q_i)^-1 mod q_i, hat_modp[i*sp+j] = (Q/q_i) mod p_j, Q = prod q_i
 *   qlql[i] = floor(Ql * (Ql^-1 mod ql) / ql) mod q_i, qlinv[i] = ql^-1 mod q_i (l = T-1, Ql = prod_{i<l} q_i) */
void orc_switch_tables(const uint64_t* q, uint32_t sq, const uint64_t* p, uint32_t sp, uint64_t* hat_inv, uint64_t* hat_modp);
void orc_rescale_tables(const uint64_t* q, uint32_t T, uint64_t* qlql, uint64_t* qlinv);
/* DCRTPolyImpl::ApproxSwitchCRTBasis (dcrtpoly-impl.h:1034-1082, HAVE_INT128 branch; Mul128 and
 * BarrettUint128ModUint64 of utils/utilities-int.h:47-110) */
void orc_approx_switch(uint64_t* out, const uint64_t* in, size_t B, uint32_t N, const uint64_t* q, uint32_t sq,
                       const uint64_t* p, uint32_t sp, const uint64_t* hat_inv, const uint64_t* hat_modp);
/* DCRTPolyImpl::ApproxModUp (dcrtpoly-impl.h:1085-1133): in over Q (format 0 EVALUATION / 1 COEFFICIENT) ->
 * out over QP in EVALUATION */
void orc_approx_mod_up(const orc_ctx* Q, const orc_ctx* P, uint64_t* out, const uint64_t* in, size_t B, int format);
/* DCRTPolyImpl::ApproxModDown (dcrtpoly-impl.h:1136-1175): in over QP, EVALUATION -> out over Q, EVALUATION;
 * t = 0 (CKKS) or the plaintext modulus (BGV) */
void orc_approx_mod_down(const orc_ctx* Q, const orc_ctx* P, uint64_t* out, const uint64_t* in, size_t B, uint64_t t);
/* t = 0: DCRTPolyImpl::DropLastElementAndScale (dcrtpoly-impl.h:746-769); t > 0: DCRTPolyImpl::ModReduce
 * (dcrtpoly-impl.h:792-812).  full = all T towers, rest = first T-1, last = tower T-1.  With format 1 the
 * CKKS form returns forward-transformed towers, as the reference does (dcrtpoly-impl.h:765-766). */
void orc_rescale(const orc_ctx* full, const orc_ctx* rest, const orc_ctx* last, uint64_t* out, const uint64_t* in,
                 size_t B, int format, uint64_t t);

/* ---- BFV (HPS) set: exact basis switch and scaling (second half) ----
 * DCRTPolyImpl::SwitchCRTBasis (dcrtpoly-impl.h:1178-1230): like ApproxSwitchCRTBasis, plus the number of q-overflows
 * alpha = (unsigned)(0.5 + sum_i double(y_i) * qinv[i]) (IEEE double, products and sums in index order, no fused
 * multiply-add) and out_j = sum_j mod p_j - [alpha*Q]_{p_j}.  qinv[i] = 1./double(q_i) and alpha_q_modp[a*sp+j] as
 * bfvrns-cryptoparameters.cpp:270-273, 343-350. */
void orc_exact_tables(const uint64_t* q, uint32_t sq, const uint64_t* p, uint32_t sp, double* qinv, uint64_t* alpha_q_modp);
void orc_switch_crt_basis(uint64_t* out, const uint64_t* in, size_t B, uint32_t N, const uint64_t* q, uint32_t sq,
                          const uint64_t* p, uint32_t sp, const uint64_t* hat_inv, const uint64_t* hat_modp,
                          const double* qinv, const uint64_t* alpha_q_modp);
/* DCRTPolyImpl::ExpandCRTBasis (dcrtpoly-impl.h:1311-1358): in over Q (format 0/1) -> out over QP in result_format */
void orc_expand_crt_basis(const orc_ctx* Q, const orc_ctx* P, uint64_t* out, const uint64_t* in, size_t B, int format,
                          int result_format);
/* DCRTPolyImpl::ExpandCRTBasisReverseOrder (dcrtpoly-impl.h:1361-1410): out over PQ (new towers first) */
void orc_expand_crt_basis_reverse(const orc_ctx* Q, const orc_ctx* P, uint64_t* out, const uint64_t* in, size_t B, int format,
                                  int result_format);
/* ---- BFV HPSPOVERQ / HPSPOVERQLEVELED set ----
 * DCRTPolyImpl::FastExpandCRTBasisPloverQ (dcrtpoly-impl.h:1413-1468): in over q -> out over [ql | pl], tables
 * neg_pl_qhat_inv[sq], qinv_modp[sq][spl] from the scheme (bfvrns-cryptoparameters.cpp:501-523) */
void orc_fast_expand_pl_over_q(uint64_t* out, const uint64_t* in, size_t B, uint32_t N, const uint64_t* q, uint32_t sq,
                               const uint64_t* ql, uint32_t sql, const uint64_t* pl, uint32_t spl,
                               const uint64_t* neg_pl_qhat_inv, const uint64_t* qinv_modp);
void orc_pl_over_q_tables(const uint64_t* q, uint32_t sq, const uint64_t* pl, uint32_t spl, uint64_t* neg, uint64_t* qinv_modp);
/* DCRTPolyImpl::ExpandCRTBasisQlHat (dcrtpoly-impl.h:1514-1534) */
void orc_expand_ql_hat(uint64_t* out, const uint64_t* in, size_t B, uint32_t N, const uint64_t* q, uint32_t sq, uint32_t sl,
                       const uint64_t* c);
/* DCRTPolyImpl::ScaleAndRoundPOverQ (dcrtpoly-impl.h:2052-2066) */
void orc_scale_round_p_over_q(uint64_t* out, const uint64_t* in, size_t B, uint32_t N, const uint64_t* q, uint32_t sq,
                              const uint64_t* pinv_modq);
/* DCRTPolyImpl::ScaleAndRound(paramsOutput, ...) (dcrtpoly-impl.h:1876-1960, HAVE_INT128 branch): in over
 * [B][sI+sO][N]; the output moduli o[sO] are the LAST sO towers when out_last != 0 (input index 0), else the first
 * (input index sO).  frac[sI], modo[sO][sI+1] row-major as the scheme precomputes them. */
void orc_scale_and_round(uint64_t* out, const uint64_t* in, size_t B, uint32_t N, uint32_t sI, const uint64_t* o, uint32_t sO,
                         int out_last, const double* frac, const uint64_t* modo);

/* ---- seeded input generator shared with the product's device-side fill
 * (counter based: value i depends only on (seed, i, q)) ---- */
uint64_t orc_uniform_at(uint64_t seed, uint64_t i, uint64_t q);
void     orc_fill_uniform(uint64_t* x, size_t n, uint64_t q, uint64_t seed, uint64_t first_index);

#ifdef __cplusplus
}
#endif
#endif
