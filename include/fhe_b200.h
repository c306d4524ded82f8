/*
 * fhe_b200.h — C ABI of the B200 backend for the UPMEM--OpenFHE intercept surface.
 *
 * This is the drop-in boundary: plain pointers and sizes, no C++/torch types.  It replaces what
 * sat behind the reference's PIM host runtime (all citations relative to the reference tree):
 *
 *   reference interface                                         replaced by
 *   ----------------------------------------------------------  -----------------------------------
 *   PimManager::getPim / ctor (dpu_alloc, load BOOT)             fhe_init / fhe_shutdown
 *     src/core/include/pim/PimManager.h:23-29,101-111
 *   PimManager::allocate / deallocate + DpuMemory                fhe_alloc / fhe_free (device arena)
 *     src/core/pim/host/PimManager.cpp:56-93, pim/DpuMemory.h:11-82
 *   PimManager::copy_to_pim / copy_from_pim                      fhe_h2d / fhe_d2h
 *     src/core/pim/host/PimManager.cpp:5-54
 *   PimManager::load_kernel + copy_to_pim("meta") + start_kernel fhe_launch (legacy descriptor path)
 *     pim/PimData.h:275-305,307-329; pim/common.h:9-54; pim/kernel.h:4-8
 *   PimData operator+ - * / ModAdd ModSub ModMul (+Eq, scalar)   fhe_vec_op / fhe_vec_op_scalar
 *     pim/PimData.h:46-168
 *   intended hook NativeVectorT::ModAddEq -> Run_On_Pim          fhe_vec_op
 *     src/core/lib/math/hal/intnat/mubintvecnat.cpp:250-259
 *   DCRTPolyImpl::Plus/Minus/Times, += -= *= (OMP over towers)   fhe_dcrt_op (all towers, whole batch,
 *     lattice/hal/default/dcrtpoly.h:142-200,                       one launch)
 *     dcrtpoly-impl.h:398-416
 *   DCRTPolyImpl::SwitchFormat -> ChineseRemainderTransformFTT   fhe_ntt_forward / fhe_ntt_inverse
 *     dcrtpoly-impl.h:2518-2524, poly-impl.h:412-432,
 *     math/hal/intnat/transformnat-impl.h:575-602,637-666
 *   ChineseRemainderTransformFTTNat::PreCompute (twiddle cache)  fhe_ring_create
 *     transformnat-impl.h:708-763
 *   DCRTPolyImpl::AutomorphismTransform (k) / (k, precomp)       fhe_automorphism / fhe_automorphism_map
 *     dcrtpoly-impl.h:348-369, poly-impl.h:312-380
 *   DCRTPolyImpl::ApproxSwitchCRTBasis / ApproxModUp /          fhe_approx_switch_crt_basis / fhe_approx_mod_up /
 *     ApproxModDown   dcrtpoly-impl.h:1034-1175                   fhe_approx_mod_down
 *   DCRTPolyImpl::SwitchCRTBasis / ExpandCRTBasis /              fhe_switch_crt_basis / fhe_expand_crt_basis /
 *     ScaleAndRound   dcrtpoly-impl.h:1178-1358, 1876-1960        fhe_scale_and_round
 *   DCRTPolyImpl::DropLastElementAndScale / ModReduce            fhe_drop_last_element_and_scale / fhe_mod_reduce
 *     dcrtpoly-impl.h:746-812
 *   DCRTPolyImpl::FastBaseConvqToBskMontgomery / FastRNSFloorq / fhe_fast_base_conv_q_to_bsk_montgomery / fhe_fast_rns_floor_q /
 *     FastBaseConvSK / ScaleAndRound(moduliQ, t, tgamma, ...)     fhe_fast_base_conv_sk / fhe_scale_and_round_behz
 *     dcrtpoly-impl.h:2005-2050, 2069-2420 (BFV, BEHZ)
 *   Serial::Serialize / Deserialize (SerType::BINARY) of a       fhe_serialize / fhe_deserialize
 *     DCRTPoly   utils/serial.h:99-115, dcrtpoly.h:349-365,
 *     poly.h:322-338, mubintvecnat.h:666-699
 *   FirstPrime / PreviousPrime / RootOfUnity parameter recipe    fhe_gen_params (host side, setup only)
 *     math/nbtheory-impl.h:183-231,334-379;
 *     benchmark/src/poly-benchmark-16k.cpp:82-101
 *
 * Conventions
 *   - Every function returns 0 on success, a negative FHE_E* code on failure; the message is in
 *     fhe_last_error() (thread local).  There is NO CPU fallback: without a CUDA device fhe_init
 *     fails and nothing else can be called.
 *   - Device buffers are plain device addresses carried as uint64_t (fhe_dptr), so a caller may
 *     also pass memory it allocated itself (cudaMalloc, torch).
 *   - Polynomial data layout: [batch][tower][N] uint64, coefficients contiguous per tower (the
 *     reference's DCRTPoly layout, towers of one polynomial adjacent).  Coefficients must be
 *     < q_tower (the reference's *Fast* precondition, ubintnat.h:740-745).
 *   - `stream` is an index into the context's stream pool (0 <= stream < fhe_num_streams), or
 *     FHE_STREAM_LEGACY (-1) for the CUDA legacy default stream.  Launches are asynchronous;
 *     fhe_sync waits.  Calls on different buffers may be issued concurrently from several
 *     threads (the reference's callers are OpenMP loops over towers, dcrtpoly-impl.h:412).
 */
#ifndef FHE_B200_H
#define FHE_B200_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct fhe_ctx fhe_ctx;     /* one CUDA device: streams, arena, pinned staging */
typedef struct fhe_ring fhe_ring;   /* (N, towers, q_i, psi_i) + device-resident constant tables */
typedef uint64_t fhe_dptr;          /* device address */

enum {
    FHE_OK = 0,
    FHE_EINVAL = -1,      /* bad argument */
    FHE_ECUDA = -2,       /* CUDA runtime error (message has the CUDA string) */
    FHE_ENOMEM = -3,      /* device arena exhausted */
    FHE_ENODEV = -4,      /* no usable CUDA device: there is no CPU fallback */
    FHE_EUNSUPPORTED = -5 /* valid request this build does not implement */
};

enum { FHE_OP_ADD = 0, FHE_OP_SUB = 1, FHE_OP_MUL = 2,
       /* accepted by fhe_dcrt_chain_host only: DCRTPolyImpl::SwitchFormat on the running value,
        * COEFFICIENT -> EVALUATION (forward) and EVALUATION -> COEFFICIENT (inverse) */
       FHE_OP_NTT = 3, FHE_OP_INTT = 4 };
enum { FHE_STREAM_LEGACY = -1 };

/* Kernel ids of the legacy descriptor path: the five binaries named in pim/kernel.h:4-8. */
enum { FHE_KERNEL_BOOT = 0, FHE_KERNEL_MOD_ADD = 1, FHE_KERNEL_MOD_MULT = 2, FHE_KERNEL_MOD_SUB = 3,
       FHE_KERNEL_MOD_OPS = 4 };

/* ---- runtime ------------------------------------------------------------------------------ */
const char* fhe_last_error(void);
const char* fhe_version(void);
/* opts: "key=value,..."; keys: streams (default 4), arena_mb (arena slab size, default 256), arena_grow_mb (how much the
 * arena reserves before an allocation that finds no free block waits for blocks freed behind unfinished work instead of
 * taking another slab; default a quarter of the device's memory), chunk_mb (bytes per operand per chunk of the
 * host-operand pipelines, default 32). */
int fhe_init(int device, const char* opts, fhe_ctx** out);
int fhe_shutdown(fhe_ctx* ctx);
int fhe_device(const fhe_ctx* ctx);
int fhe_num_streams(const fhe_ctx* ctx);
int fhe_sync(fhe_ctx* ctx, int stream);               /* stream = -2: whole device */
/* number of kernels this library has launched since it was loaded */
long fhe_launch_count(void);

/* ---- memory ------------------------------------------------------------------------------- */
/* Alignment: fhe_alloc returns 256-byte aligned blocks.  The polynomial kernels move 128-bit words, so every
 * polynomial pointer handed to an entry point below (a block, or a view into one at a whole number of towers) must be
 * 16-byte aligned when N >= 2; a pointer that is not is refused with FHE_EINVAL, never passed to a kernel.  The flat
 * fhe_vec_* entry points take any 8-byte aligned pointer (they fall back to a scalar element loop). */
int fhe_alloc(fhe_ctx* ctx, size_t bytes, fhe_dptr* out);        /* 256-byte aligned */
/* Stream-ordered: a block may be freed as soon as the calls that use it have RETURNED (their launches may still be
 * pending).  It is handed out again only after everything enqueued, before the fhe_free call, on the streams of this
 * context (the pool and FHE_STREAM_LEGACY) has completed; an allocation that finds the arena full waits for such
 * blocks instead of growing.  Work a caller enqueued on streams of its own must be synchronised by the caller. */
int fhe_free(fhe_ctx* ctx, fhe_dptr p);
int fhe_arena_stats(fhe_ctx* ctx, size_t* reserved_bytes, size_t* in_use_bytes);
int fhe_host_alloc(fhe_ctx* ctx, size_t bytes, void** out);      /* pinned host memory */
int fhe_host_free(fhe_ctx* ctx, void* p);
int fhe_h2d(fhe_ctx* ctx, fhe_dptr dst, const void* src, size_t bytes, int stream);
int fhe_d2h(fhe_ctx* ctx, void* dst, fhe_dptr src, size_t bytes, int stream);
int fhe_d2d(fhe_ctx* ctx, fhe_dptr dst, fhe_dptr src, size_t bytes, int stream);

/* ---- timing on the launching stream (CUDA events) ------------------------------------------ */
int fhe_timer_start(fhe_ctx* ctx, int stream);
int fhe_timer_stop(fhe_ctx* ctx, int stream, float* elapsed_ms);  /* waits for the stop event */

/* ---- parameters (host, setup) --------------------------------------------------------------- */
/* q_0 = PreviousPrime(FirstPrime(nbits, 2N), 2N), q_i = PreviousPrime(q_{i-1}, 2N),
 * psi_i = minimum primitive 2N-th root of unity mod q_i. */
int fhe_gen_params(uint32_t nbits, uint32_t N, uint32_t towers, uint64_t* q, uint64_t* psi);
uint64_t fhe_root_of_unity(uint64_t m, uint64_t q);              /* 0 if m does not divide q-1 */

/* ---- ring --------------------------------------------------------------------------------- */
/* N power of two, 1 <= N <= 65536; 4 <= q_i < 2^61 (< 2^60 when psi is given: the reference's
 * MAX_MODULUS_SIZE, basicint.h:44-45); psi_i a primitive 2N-th root mod q_i
 * (psi may be NULL when only fhe_dcrt_op is needed).  Builds the Barrett constants and, per
 * tower, the four tables PreCompute builds, and keeps them in HBM. */
int fhe_ring_create(fhe_ctx* ctx, uint32_t N, uint32_t towers, const uint64_t* q,
                    const uint64_t* psi, fhe_ring** out);
int fhe_ring_destroy(fhe_ring* ring);
uint32_t fhe_ring_n(const fhe_ring* ring);
uint32_t fhe_ring_towers(const fhe_ring* ring);
/* copies table `which` (0 W, 1 W companion, 2 W^-1, 3 W^-1 companion; N words) of one tower to host */
int fhe_ring_table(const fhe_ring* ring, uint32_t tower, int which, uint64_t* out);

/* ---- the hot path: device-resident operands -------------------------------------------------- */
/* dst = a (op) b for every tower of every polynomial of the batch; dst may alias a or b. */
int fhe_dcrt_op(fhe_ring* ring, int op, fhe_dptr dst, fhe_dptr a, fhe_dptr b, size_t batch, int stream);
/* dst = a (+|-) scalars[tower] for every coefficient of every tower (op FHE_OP_ADD / FHE_OP_SUB; one host constant per tower,
 * reduced modulo q_tower first as NativeVector::ModAdd / ModSub do): DCRTPolyImpl::Plus / Minus(const std::vector<Integer>&)
 * (dcrtpoly-impl.h:556-563, 577-584; Plus in EVALUATION format - in COEFFICIENT format the reference adds to coefficient 0
 * only, which this call does not do).  dst may be a.  At most 64 towers. */
int fhe_dcrt_op_scalars(fhe_ring* ring, int op, fhe_dptr dst, fhe_dptr a, const uint64_t* scalars, size_t batch, int stream);
/* in-place COEFFICIENT -> EVALUATION (bit-reversed order), and back */
int fhe_ntt_forward(fhe_ring* ring, fhe_dptr x, size_t batch, int stream);
int fhe_ntt_inverse(fhe_ring* ring, fhe_dptr x, size_t batch, int stream);
/* same, selecting the kernel family: 0 auto, 1 single-CTA shared-memory kernel (N <= 4096), 2 one-tile-per-CTA two-pass
 * kernels, 3 warp-specialised persistent two-pass kernels (N = 2^14..2^16; measured slower, never picked by auto) */
int fhe_ntt_algo(fhe_ring* ring, fhe_dptr x, size_t batch, int inverse, int algo, int stream);
/* c = INTT(NTT(a) .* NTT(b)): negacyclic product of COEFFICIENT-format polynomials; a and b are
 * left in EVALUATION format (they are transformed in place), c may alias neither. */
int fhe_poly_mul(fhe_ring* ring, fhe_dptr c, fhe_dptr a, fhe_dptr b, size_t batch, int stream);
/* Ciphertext tensor product, 2 x 2 case of LeveledSHEBase::EvalMultCore
 * (src/pke/lib/schemebase/base-leveledshe.cpp:667-672): with a = (a[0], a[1]), b = (b[0], b[1]) batches of
 * DCRTPolys in EVALUATION format, c[0] = a0*b0, c[1] = a1*b0 + a0*b1, c[2] = a1*b1, one fused launch
 * (56 B of HBM traffic per coefficient slot instead of 120 B for four Times and one +=).  The three
 * outputs must be distinct buffers; an output may be the same buffer as an input. */
int fhe_ct_tensor_product(fhe_ring* ring, const fhe_dptr* c, const fhe_dptr* a, const fhe_dptr* b, size_t batch,
                          int stream);
/* The multiplication pipelines BASELINE.json's configs 4 and 5 are worded as, composed from the calls
 * above.  transform 0: tensor product only.  1: forward NTT of the four inputs (in place, they stay in
 * EVALUATION format), tensor product, inverse NTT of the three outputs (COEFFICIENT in, COEFFICIENT
 * out: the primitive part of BFVrns EvalMult, bfvrns-leveledshe.cpp:193-343).  2: inverse NTT of the
 * inputs, tensor product, forward NTT of the outputs (config 4's wording). */
int fhe_ct_multiply(fhe_ring* ring, const fhe_dptr* c, const fhe_dptr* a, const fhe_dptr* b, size_t batch,
                    int transform, int stream);
/* Ring automorphism X -> X^k, k odd (DCRTPolyImpl::AutomorphismTransform, dcrtpoly-impl.h:348-369 ->
 * PolyImpl::AutomorphismTransform, poly-impl.h:312-365), every tower of every polynomial of the batch, one
 * launch.  format 0 = EVALUATION (bit-reversed order; index map computed on the fly), 1 = COEFFICIENT
 * (a negated zero is stored as q, as the reference's `q - value` does).  k even -> FHE_EINVAL
 * ("Automorphism index not odd", the reference's math_error).  dst must not be src. */
int fhe_automorphism(fhe_ring* ring, fhe_dptr dst, fhe_dptr src, size_t batch, uint32_t k, int format, int stream);
/* PrecomputeAutoMap (src/core/lib/math/nbtheory2.cpp:274-286): map[rev(j)] = rev(((2j+1)k mod 2N) >> 1); host. */
int fhe_precompute_automap(uint32_t N, uint32_t k, uint32_t* map);
/* The two-argument form (poly-impl.h:368-380): dst[j] = src[map[j]] per tower, map = N uint32 in HBM. */
int fhe_automorphism_map(fhe_ring* ring, fhe_dptr dst, fhe_dptr src, size_t batch, fhe_dptr map, int stream);
/* ---- RNS basis conversion and scaling ---------------------------------- */
/* A ring over towers [first, first+count) of `parent`, sharing its constant tables (no copy); the parent must
 * outlive the view.  ILDCRTParams of a sub-basis (e.g. Q and P inside QP). */
int fhe_ring_view(fhe_ring* parent, uint32_t first, uint32_t count, fhe_ring** out);

typedef struct fhe_basis fhe_basis;  /* tables of ApproxSwitchCRTBasis from one ring's moduli to another's */
/* qhat_inv_modq [sq] = [(Q/q_i)^-1]_{q_i} and qhat_modp [sq][sp] = [Q/q_i]_{p_j} as the schemes precompute them
 * (src/pke/lib/schemerns/rns-cryptoparameters.cpp:199-229); either may be NULL, then it is derived from the
 * moduli.  Shoup companions and the 128-bit Barrett constants (rns-cryptoparameters.cpp:273-283) are built here.
 * Both rings must have the same N and live on the same context; they must outlive the basis.  A basis keeps
 * scratch buffers, so use one per stream. */
int fhe_basis_create(fhe_ring* from, fhe_ring* to, const uint64_t* qhat_inv_modq, const uint64_t* qhat_modp,
                     fhe_basis** out);
int fhe_basis_destroy(fhe_basis* basis);
/* host copies of the two tables in use */
int fhe_basis_tables(const fhe_basis* basis, uint64_t* qhat_inv_modq, uint64_t* qhat_modp);
/* DCRTPolyImpl::ApproxSwitchCRTBasis (dcrtpoly-impl.h:1034-1082): src over `from` [batch][sq][N] ->
 * dst over `to` [batch][sp][N]; the format is whatever src is in (the schemes call it in COEFFICIENT). */
int fhe_approx_switch_crt_basis(fhe_basis* basis, fhe_dptr dst, fhe_dptr src, size_t batch, int stream);
/* DCRTPolyImpl::ApproxModUp (dcrtpoly-impl.h:1085-1133) with basis = Q -> P: src over Q in `format`
 * (0 EVALUATION, 1 COEFFICIENT) -> dst over QP [batch][sq+sp][N] in EVALUATION.  src is not modified. */
int fhe_approx_mod_up(fhe_basis* q_to_p, fhe_dptr dst_qp, fhe_dptr src_q, size_t batch, int format, int stream);
/* DCRTPolyImpl::ApproxModDown (dcrtpoly-impl.h:1136-1175) with basis = P -> Q: src over QP [batch][sq+sp][N]
 * in EVALUATION -> dst over Q in EVALUATION.  t = 0: CKKS form; t > 0: BGV form (multiplies the P part by
 * [t^-1]_{p_j} and the switched part by t, bgvrns-cryptoparameters.cpp:83-88).  [P^-1]_{q_i}
 * (rns-cryptoparameters.cpp:190-197) is derived from the moduli. */
int fhe_approx_mod_down(fhe_basis* p_to_q, fhe_dptr dst_q, fhe_dptr src_qp, size_t batch, uint64_t t, int stream);

/* DCRTPolyImpl::SwitchCRTBasis (dcrtpoly-impl.h:1178-1230): the exact form — counts the q-overflows in IEEE double
 * (`nu = 0.5 + sum_i double(y_i) * (1./double(q_i))`, in index order, not fused) and subtracts [alpha*Q]_{p_j}; the
 * tables of bfvrns-cryptoparameters.cpp:270-273, 343-350 are derived from the moduli. */
int fhe_switch_crt_basis(fhe_basis* basis, fhe_dptr dst, fhe_dptr src, size_t batch, int stream);
/* DCRTPolyImpl::ExpandCRTBasis (dcrtpoly-impl.h:1311-1358) with basis = Q -> P: src over Q in `format` -> dst over QP
 * [batch][sq+sp][N] in `result_format` (0 EVALUATION, 1 COEFFICIENT).  src is not modified. */
int fhe_expand_crt_basis(fhe_basis* q_to_p, fhe_dptr dst_qp, fhe_dptr src_q, size_t batch, int format, int result_format,
                         int stream);

/* DCRTPolyImpl::ExpandCRTBasisReverseOrder (dcrtpoly-impl.h:1361-1410, the multiparty key-switching form): as
 * fhe_expand_crt_basis, but the new towers come first: dst over PQ [batch][sp+sq][N]. */
int fhe_expand_crt_basis_reverse_order(fhe_basis* q_to_p, fhe_dptr dst_pq, fhe_dptr src_q, size_t batch, int format,
                                       int result_format, int stream);
/* DCRTPolyImpl::FastExpandCRTBasisPloverQ (dcrtpoly-impl.h:1413-1468; BFV multTech HPSPOVERQ / HPSPOVERQLEVELED):
 * src over Q -> dst over QlPl [batch][sql+spl][N] holding x*Pl/Q.  q_to_pl is created with the scheme's tables
 * qhat_inv_modq = m_negRlQHatInvModq[l] ([-Pl*(Q/q_i)^-1]_{q_i}) and qhat_modp = m_qInvModr ([q_i^-1]_{p_j})
 * (bfvrns-cryptoparameters.cpp:501-523); pl_to_ql is the exact switch Pl -> Ql with its tables derived from the moduli
 * (m_RlHatInvModr, m_RlHatModq, m_alphaRlModq, m_rInv: :390-443).  No NTT: the format is whatever src is in (the scheme
 * calls it in COEFFICIENT).  src is not modified. */
int fhe_fast_expand_crt_basis_pl_over_q(fhe_basis* q_to_pl, fhe_basis* pl_to_ql, fhe_dptr dst_qlpl, fhe_dptr src_q,
                                        size_t batch, int stream);

typedef struct fhe_tower_scale fhe_tower_scale;  /* one constant c_i < q_i per tower of `ring` (+ Shoup companions) */
int fhe_tower_scale_create(fhe_ring* ring, const uint64_t* c, fhe_tower_scale** out);
int fhe_tower_scale_destroy(fhe_tower_scale* plan);
/* DCRTPolyImpl::ExpandCRTBasisQlHat (dcrtpoly-impl.h:1514-1534) with c = m_QlHatModq[l] over the ring Q: src over the
 * first size_ql towers [batch][size_ql][N] -> dst over Q [batch][T][N]: towers < size_ql multiplied by c_i, the rest
 * zero.  Format preserved. */
int fhe_expand_crt_basis_ql_hat(fhe_tower_scale* plan, uint32_t size_ql, fhe_dptr dst_q, fhe_dptr src_ql, size_t batch,
                                int stream);
/* DCRTPolyImpl::Times(const std::vector<NativeInteger>&) / TimesNoCheck (dcrtpoly-impl.h:617-638): every tower times its
 * own constant, c_i reduced modulo q_i by the caller (NativeVectorT::ModMulEq reduces an oversized scalar first,
 * mubintvecnat.cpp:370-385).  dst may be src.  Format preserved. */
int fhe_dcrt_times_towers(fhe_tower_scale* plan, fhe_dptr dst, fhe_dptr src, size_t batch, int stream);
/* DCRTPolyImpl::ScaleAndRoundPOverQ (dcrtpoly-impl.h:2052-2066) with c = pInvModq over the ring Q (T towers): src over
 * Q plus one extra tower p [batch][T+1][N] -> dst over Q: ((x_i - (x_p mod q_i)) mod q_i) * c_i.  Format preserved. */
int fhe_scale_and_round_p_over_q(fhe_tower_scale* plan, fhe_dptr dst_q, fhe_dptr src_qp, size_t batch, int stream);

typedef struct fhe_scale_round fhe_scale_round;
/* DCRTPolyImpl::ScaleAndRound(paramsOutput, tOSHatInvModsDivsModo, tOSHatInvModsDivsFrac, modoBarretMu)
 * (dcrtpoly-impl.h:1876-1960): `ring` = the input basis (sI + size_out towers); the output moduli are its last
 * size_out towers when out_last != 0 (the reference's "output modulus is P" case), else its first.  frac [sI] and
 * modo [size_out][sI+1] (row-major) are the scheme's tables (bfvrns-cryptoparameters.cpp:278-311); the 128-bit
 * Barrett constants are built here.  Format is whatever src is in (the schemes call it in COEFFICIENT). */
int fhe_scale_round_create(fhe_ring* ring, uint32_t size_out, int out_last, const double* frac, const uint64_t* modo,
                           fhe_scale_round** out);
int fhe_scale_round_destroy(fhe_scale_round* plan);
int fhe_scale_and_round(fhe_scale_round* plan, fhe_dptr dst, fhe_dptr src, size_t batch, int stream);

typedef struct fhe_rescale fhe_rescale;  /* constants for dropping the last tower of `ring` */
/* qlql [T-1] = [Ql*[Ql^-1]_{ql}/ql]_{q_i} and qlinv [T-1] = [ql^-1]_{q_i}
 * (src/pke/lib/scheme/ckksrns/ckksrns-cryptoparameters.cpp:66-88); NULL -> derived from the moduli.
 * The ring must have >= 2 towers and outlive the plan; one plan per stream. */
int fhe_rescale_create(fhe_ring* ring, const uint64_t* qlql, const uint64_t* qlinv, fhe_rescale** out);
int fhe_rescale_destroy(fhe_rescale* plan);
int fhe_rescale_tables(const fhe_rescale* plan, uint64_t* qlql, uint64_t* qlinv);
/* DCRTPolyImpl::DropLastElementAndScale (dcrtpoly-impl.h:746-769, CKKS rescale): src [batch][T][N] in `format`
 * -> dst [batch][T-1][N].  With format 1 (COEFFICIENT) the towers come back forward-transformed, as the
 * reference leaves them (dcrtpoly-impl.h:765-766). */
int fhe_drop_last_element_and_scale(fhe_rescale* plan, fhe_dptr dst, fhe_dptr src, size_t batch, int format, int stream);
/* DCRTPolyImpl::ModReduce (dcrtpoly-impl.h:792-812, BGV modulus switching) with plaintext modulus t > 0;
 * [-t^-1]_{ql} and [t]_{q_i} (bgvrns-cryptoparameters.cpp:91-99) are derived here.  Format is preserved. */
int fhe_mod_reduce(fhe_rescale* plan, fhe_dptr dst, fhe_dptr src, size_t batch, uint64_t t, int format, int stream);

/* ---- BFV decryption: the scaling step ------------------------------------------------------------ */
/* DCRTPolyImpl::ScaleAndRound(t, tQHatInvModqDivqModt, ..Precon, tQHatInvModqBDivqModt, ..Precon, tQHatInvModqDivqFrac,
 * tQHatInvModqBDivqFrac) -> NativePoly (dcrtpoly-impl.h:1537-1816): src over the ring Q [batch][T][N], COEFFICIENT ->
 * dst [batch][N] = round(t/Q * x) mod t.  The reference picks one of eight branches from the bit lengths of q_0, t and
 * the tower count (t a power of two; residues used whole or split in halves; products reduced modulo t or not); the
 * same choice is made here and every branch follows the reference statement for statement, IEEE doubles in its order
 * of operations.  Tables: the scheme's (bfvrns-cryptoparameters.cpp:447-494: modt / frac, and b_modt / b_frac for the
 * split branch), or NULL to have them derived from the moduli; Shoup companions are built here.  2 <= t < 2^60. */
typedef struct fhe_decrypt_scale fhe_decrypt_scale;
int fhe_decrypt_scale_create(fhe_ring* ring, uint64_t t, const uint64_t* modt, const uint64_t* b_modt, const double* frac,
                             const double* b_frac, fhe_decrypt_scale** out);
int fhe_decrypt_scale_destroy(fhe_decrypt_scale* plan);
/* host copies of the tables in use; *branch = 4 * (t power of two) + 2 * (split) + (products reduced modulo t) */
int fhe_decrypt_scale_tables(const fhe_decrypt_scale* plan, uint64_t* modt, uint64_t* b_modt, double* frac, double* b_frac,
                             int* branch);
int fhe_scale_and_round_native(fhe_decrypt_scale* plan, fhe_dptr dst, fhe_dptr src, size_t batch, int stream);

/* ---- hybrid key switching (CKKS / BGV relinearisation and rotations) --------------------------- */
/* KeySwitchHYBRID (src/pke/lib/keyswitch/keyswitch-hybrid.cpp:324-480) at one level as a device composition.
 * ring_qlp: the size_ql current towers of Q followed by the towers of P; alpha: towers per digit
 * (CryptoParametersRNS::GetNumPerPartQ); digits = ceil(size_ql / alpha), the last one may be shorter.  The digit
 * tables ([(Q_j/q_i)^-1]_{q_i}, [Q_j/q_i]_{p}, rns-cryptoparameters.cpp:289-340) and [P^-1]_{q_i} are derived from
 * the moduli.  The ring must outlive the plan; a plan keeps scratch buffers, so use one per stream.
 *   fhe_keyswitch_precompute = EvalKeySwitchPrecomputeCore (:330-420): c over Ql [batch][size_ql][N], EVALUATION ->
 *       digits [digits][batch][T][N], EVALUATION (per digit: INTT of its towers, ApproxSwitchCRTBasis to the
 *       complement basis, NTT, the digit's own towers copied in).
 *   fhe_keyswitch_inner = EvalFastKeySwitchCoreExt (:442-480): ctilda0 = sum_j digits[j] * evk_b[j],
 *       ctilda1 = sum_j digits[j] * evk_a[j], one fused pass; the key polys are given over the towers of ring_qlp
 *       ([digits][T][N], EVALUATION: towers i < size_ql and size_q.. of the scheme's key, :463-476).
 *   fhe_keyswitch_fast = EvalFastKeySwitchCore (:415-437), what every rotation of a hoisted set calls on the shared
 *       digits: the inner product above, then ApproxModDown of both halves at once; ct0, ct1 over Ql
 *       [batch][size_ql][N]; t = 0 CKKS, t > 0 BGV / BFV.
 *   fhe_keyswitch_hybrid = KeySwitchCore (:324-328) = precompute, then fast.  Between the launches that read c and
 *       the two that write ct0 / ct1 everything touches the plan's buffers and the key only; that part is replayed
 *       from a captured CUDA graph where the launches cost more than the kernels (FHE_B200_KS_GRAPH=0 turns it off). */
typedef struct fhe_keyswitch fhe_keyswitch;
int fhe_keyswitch_create(fhe_ring* ring_qlp, uint32_t size_ql, uint32_t alpha, fhe_keyswitch** out);
int fhe_keyswitch_destroy(fhe_keyswitch* plan);
uint32_t fhe_keyswitch_digits(const fhe_keyswitch* plan);
int fhe_keyswitch_precompute(fhe_keyswitch* plan, fhe_dptr digits, fhe_dptr c, size_t batch, int stream);
int fhe_keyswitch_inner(fhe_keyswitch* plan, fhe_dptr ctilda0, fhe_dptr ctilda1, fhe_dptr digits, fhe_dptr evk_b,
                        fhe_dptr evk_a, size_t batch, int stream);
int fhe_keyswitch_fast(fhe_keyswitch* plan, fhe_dptr ct0, fhe_dptr ct1, fhe_dptr digits, fhe_dptr evk_b, fhe_dptr evk_a,
                       size_t batch, uint64_t t, int stream);
int fhe_keyswitch_hybrid(fhe_keyswitch* plan, fhe_dptr ct0, fhe_dptr ct1, fhe_dptr c, fhe_dptr evk_b, fhe_dptr evk_a,
                         size_t batch, uint64_t t, int stream);
/* acc0 += ct0, acc1 += ct1 of the same key switch - the step every caller of KeySwitchCore in the schemes takes next
 * (base-leveledshe.cpp:210-213: cv[0] += (*ab)[0]; cv[1] += (*ab)[1]) - folded into the composition's last launch.
 * acc0, acc1 over Ql, EVALUATION, distinct from each other and from c. */
int fhe_keyswitch_hybrid_add(fhe_keyswitch* plan, fhe_dptr acc0, fhe_dptr acc1, fhe_dptr c, fhe_dptr evk_b, fhe_dptr evk_a,
                             size_t batch, uint64_t t, int stream);

/* ---- wire format of device-resident polys ---------------------------------------------------- */
/* The reference's binary serialisation of one DCRTPoly — Serial::Serialize(poly, os, SerType::BINARY)
 * (utils/serial.h:99-103): a cereal portable-binary stream of DCRTPolyImpl::save (dcrtpoly.h:349-354) over
 * PolyImpl::save (poly.h:322-327), NativeVectorT::save (mubintvecnat.h:666-675: N raw u64 per tower) and the
 * params (ildcrtparams.h:350-355, elemparams.h:223-231), BigInteger = bigintdyn (MATHBACKEND 4).  Byte-identical
 * to what the reference writes for DCRTPoly(params, format) with params = ILDCRTParams(2N, moduli, roots).
 * fhe_serial_bytes: size of one poly's stream.  fhe_serialize: src [batch][T][N] in HBM -> `batch` consecutive
 * streams at dst_host (headers written by the host, tower payloads copied from HBM straight into their
 * places; complete after fhe_sync on `stream`).  fhe_deserialize: parses `batch` consecutive streams (the
 * reference's own output for any DCRTPoly over native towers, class versions <= 1), checks N, the moduli and
 * the roots against the ring, and copies the payloads to dst; *format receives the polys' common format,
 * *consumed the bytes read.  Nothing is written to dst when a stream is rejected (FHE_EINVAL). */
size_t fhe_serial_bytes(fhe_ring* ring);
/* Host-only (no device): the stream of one poly with its payload holes left untouched — returns the stream
 * size, writes the headers when cap suffices, and the byte offset of every tower's N*8 payload to data_off[T]
 * (psi NULL -> roots written as 0); and the checker fhe_deserialize runs, reporting where the payloads lie. */
size_t fhe_serial_template(uint32_t N, uint32_t T, const uint64_t* q, const uint64_t* psi, int format, void* blob,
                           size_t cap, size_t* data_off);
int fhe_serial_parse(uint32_t N, uint32_t T, const uint64_t* q, const uint64_t* psi, const void* blob, size_t len,
                     size_t* data_off, int* format, size_t* consumed);
int fhe_serialize(fhe_ring* ring, void* dst_host, fhe_dptr src, size_t batch, int format, int stream);
int fhe_deserialize(fhe_ring* ring, fhe_dptr dst, const void* src_host, size_t len, size_t batch, int* format,
                    size_t* consumed, int stream);

/* seeded uniform coefficients in [0, q_tower) (synthetic inputs for tests and benchmarks) */
int fhe_fill_uniform(fhe_ring* ring, fhe_dptr dst, size_t batch, uint64_t seed, int stream);

/* ---- BV key switching: the digit decomposition ---------------------------------------------------- */
/* DCRTPolyImpl::CRTDecompose(baseBits = 0) (dcrtpoly-impl.h:266-288; KeySwitchBV::EvalKeySwitchPrecomputeCore,
 * src/pke/lib/keyswitch/keyswitch-bv.cpp:308-312): src over the ring [batch][T][N] in `format` (0 EVALUATION,
 * 1 COEFFICIENT) -> dst [T][batch][T][N], EVALUATION: digit i holds, in tower k, the COEFFICIENT tower i of src moved to
 * q_k with NativeVector::SwitchModulus (mubintvecnat.cpp:112-135) and transformed, and in tower i the EVALUATION tower
 * itself.  The plan holds the T x T switching constants and scratch (one per stream).  src is not modified. */
typedef struct fhe_crt_decomp fhe_crt_decomp;
int fhe_crt_decompose_create(fhe_ring* ring, fhe_crt_decomp** out);
int fhe_crt_decompose_destroy(fhe_crt_decomp* plan);
int fhe_crt_decompose(fhe_crt_decomp* plan, fhe_dptr dst_digits, fhe_dptr src, size_t batch, int format, int stream);

/* ---- BFV, the BEHZ multiplication technique ------------------------------------------------------- */
/* ring_qbsk: size_q towers of Q followed by Bsk = B u {msk} (at least one tower of B; msk last), with roots of unity,
 * the ring CryptoParametersBFVRNS builds as m_paramsQBsk (src/pke/lib/scheme/bfvrns/bfvrns-cryptoparameters.cpp:
 * 665-711).  Every table of :724-891 is derived here from the moduli and the plaintext modulus t (mtilde = 2^16,
 * gamma = 2^26 are the scheme's constants): fhe_behz_table returns host copies, numbered 0 mtildeQHatInvModq[nq],
 * 1 QHatModbsk[nq][nb], 2 QHatModmtilde[nq], 3 QModbsk[nb], 4 negQInvModmtilde, 5 mtildeInvModbsk[nb], 6 tQHatInvModq[nq],
 * 7 qInvModbsk[nq][nb], 8 tQInvModbsk[nb], 9 BHatInvModb[nb-1], 10 BHatModq[nb-1][nq], 11 BHatModmsk[nb-1],
 * 12 BInvModmsk, 13 BModq[nq], 14 tgamma, 15 negInvqModtgamma[nq], 16 tgammaQHatInvModq[nq].  The plan keeps scratch
 * buffers: one per stream.  The ring must outlive it. */
typedef struct fhe_behz fhe_behz;
int fhe_behz_create(fhe_ring* ring_qbsk, uint32_t size_q, uint64_t t, fhe_behz** out);
int fhe_behz_destroy(fhe_behz* plan);
int fhe_behz_table(const fhe_behz* plan, int which, uint64_t* out, size_t* len);     /* out may be NULL (length only) */
/* DCRTPolyImpl::FastBaseConvqToBskMontgomery (dcrtpoly-impl.h:2069-2209): src over Q [batch][nq][N] in `format`
 * (0 EVALUATION, 1 COEFFICIENT) -> dst over QBsk [batch][nq+nb][N], EVALUATION.  src is not modified. */
int fhe_fast_base_conv_q_to_bsk_montgomery(fhe_behz* plan, fhe_dptr dst_qbsk, fhe_dptr src_q, size_t batch, int format,
                                           int stream);
/* DCRTPolyImpl::FastRNSFloorq (dcrtpoly-impl.h:2212-2271), in place on x over QBsk [batch][nq+nb][N] (the scheme calls
 * it in COEFFICIENT): the towers of Bsk receive floor(t/Q * x).  update_q != 0 also leaves the towers of Q multiplied
 * by [t*(Q/q_i)^-1]_{q_i} as the reference does (:2233-2243; FastBaseConvSK overwrites them, so the scheme's sequence
 * can pass 0 and save the pass). */
int fhe_fast_rns_floor_q(fhe_behz* plan, fhe_dptr x_qbsk, size_t batch, int update_q, int stream);
/* fhe_fast_rns_floor_q for callers that may not allocate (CUDA graph capture, serving loops, harnesses that forbid
 * cudaMalloc).  No counterpart in the reference.  Same arguments and the
 * same result; the conversion's intermediate lives in a static 8 MiB device array instead of the plan's scratch, so
 * batch * (towers of Bsk) * N must not exceed 2^20 words (FHE_EINVAL otherwise), and the call performs no cudaMalloc /
 * cudaFree / arena allocation.  Every GPU that loads the library holds the 8 MiB until exit.  All calls on a device share
 * the array (nothing in it has to start at zero: the first kernel writes every word the second reads; it keeps the last
 * call's contents): calls that overlap on the GPU - different streams, or different plans - race; serialise them.  The
 * first call on a device resolves the array's device address (cudaGetSymbolAddress): under graph capture make one
 * call outside the capture first. */
int fhe_fast_rns_floor_q_noalloc(fhe_behz* plan, fhe_dptr x_qbsk, size_t batch, int update_q, int stream);
/* DCRTPolyImpl::FastBaseConvSK (dcrtpoly-impl.h:2309-2420): src over QBsk (only the towers of Bsk are read, none is
 * modified) -> dst over Q [batch][nq][N]. */
int fhe_fast_base_conv_sk(fhe_behz* plan, fhe_dptr dst_q, fhe_dptr src_qbsk, size_t batch, int stream);
/* DCRTPolyImpl::ScaleAndRound(moduliQ, t, tgamma, tgammaQHatModq, .., negInvqModtgamma, ..) -> NativePoly
 * (dcrtpoly-impl.h:2005-2050, BEHZ decryption): src over Q [batch][nq][N], COEFFICIENT -> dst [batch][N]. */
int fhe_scale_and_round_behz(fhe_behz* plan, fhe_dptr dst, fhe_dptr src_q, size_t batch, int stream);

/* ---- flat vectors with an explicit modulus (the PimData surface) ----------------------------- */
int fhe_vec_op(fhe_ctx* ctx, int op, fhe_dptr dst, fhe_dptr a, fhe_dptr b, size_t n, uint64_t q, int stream);
int fhe_vec_op_scalar(fhe_ctx* ctx, int op, fhe_dptr dst, fhe_dptr a, uint64_t scalar, size_t n,
                      uint64_t q, int stream);

/* mod-ops of the PimData surface (pim/PimData.h:174-233): kind is `enum mod_kernel` of pim/common.h,
 * `mod` the vector's current modulus; p / q as PimData::do_mod_ops passes them (new modulus,
 * exponent, numerator or divisor in p; denominator in q).  Values follow OpenFHE's CPU NativeVector
 * methods (mubintvecnat.cpp:111-195, 370-487).  dst may alias src.  MODINVERSE(EQ) blocks and
 * returns FHE_EINVAL when an element has no inverse (the reference throws math_error). */
int fhe_vec_modop(fhe_ctx* ctx, int kind, fhe_dptr dst, fhe_dptr src, size_t n, uint64_t mod, uint64_t p,
                  uint64_t q, int stream);

/* Order-independent digest of nwords device words: sum_xor[0] = sum mod 2^64, sum_xor[1] = xor.
 * Digests of shards add up (sum) / xor together to the digest of the whole, so sharded results can be
 * compared without gathering them.  Blocking. */
int fhe_checksum(fhe_ctx* ctx, fhe_dptr src, size_t nwords, uint64_t* sum_xor, int stream);

/* ---- legacy descriptor path ------------------------------------------------------------------ */
/* desc points to a `struct mod_add_sub_mult` (pim/common.h:23-30) whose operands are byte offsets
 * from `heap_base` and byte sizes; kernel_id picks add / sub / mult; for FHE_KERNEL_MOD_OPS desc
 * points to a `struct mod_ops` (pim/common.h:48-55).  Blocking, like
 * PimManager::start_kernel (DPU_SYNCHRONOUS, PimManager.h:68). */
int fhe_launch(fhe_ctx* ctx, int kernel_id, const void* desc, fhe_dptr heap_base);

/* ---- end-to-end convenience: HOST operands -------------------------------------------------- */
/* dst = a (op) b with host buffers of batch*towers*N words each.  Streams the batch through the
 * device in chunks: pinned staging, H2D / kernel / D2H overlapped on separate streams. Blocking. */
int fhe_dcrt_op_host(fhe_ring* ring, int op, uint64_t* dst, const uint64_t* a, const uint64_t* b, size_t batch);
/* Several ops on the SAME host operands with one upload: dsts[i] = a (ops[i]) b, i < nops <= 4.
 * This is the PimData usage pattern of the reference (operands copied to the device once in the
 * constructor, results fetched by sync(), pim/PimData.h:16-34) applied to a whole batch. Blocking. */
int fhe_dcrt_multi_op_host(fhe_ring* ring, const int* ops, int nops, uint64_t* const* dsts,
                           const uint64_t* a, const uint64_t* b, size_t batch);
/* The reference's in-place operator sequence on host operands (DCRTPolyImpl::operator+= / -= / *=,
 * dcrtpoly-impl.h:398-416, dcrtpoly.h:142-148, and DCRTPolyImpl::SwitchFormat, dcrtpoly-impl.h:2518-2524): for
 * i < nops <= 8, a (ops[i])= b, or a.SwitchFormat() for FHE_OP_NTT / FHE_OP_INTT, whole batch; the intermediate
 * values never leave the device; one upload of a and b (b is not read, and may be NULL, when no op uses it), one
 * download of a.  Chunks of the batch are pipelined over the stream pool (H2D, kernels, D2H overlap). Blocking. */
int fhe_dcrt_chain_host(fhe_ring* ring, const int* ops, int nops, uint64_t* a, const uint64_t* b, size_t batch);
/* in-place transform of a host buffer (inverse = 0 forward, 1 inverse). Blocking. */
int fhe_ntt_host(fhe_ring* ring, uint64_t* x, size_t batch, int inverse);

/* ---- several devices of one node ----------------------------------------------- */
/* The reference scatters a vector over `nr_dpus` DPUs (PimManager::copy_to_pim, src/core/pim/host/PimManager.cpp:15-24:
 * size / nr_dpus words each) and launches the same kernel on all of them (PimManager.h:68).  Here a group is `ndev`
 * GPUs driven by one process: one context per device, the batch split in contiguous blocks (sizes differing by at most
 * one polynomial, the first ranks take the remainder), the ring's tables replicated on every device, no collective
 * inside the path.  NCCL (libnccl.so.2, loaded at run time) is used only by the gathers below, on this library's own
 * streams.  Every array argument below has one entry per device of the group, in rank order. */
typedef struct fhe_group fhe_group;
typedef struct fhe_group_ring fhe_group_ring;
/* devices NULL: 0..ndev-1; ndev 0: every visible device.  opts as fhe_init. */
int fhe_group_create(int ndev, const int* devices, const char* opts, fhe_group** out);
int fhe_group_destroy(fhe_group* group);
int fhe_group_size(const fhe_group* group);
fhe_ctx* fhe_group_ctx(fhe_group* group, int rank);
int fhe_group_sync(fhe_group* group);                       /* every device */
/* the block of a batch that `rank` owns: items [*first, *first + *count) */
int fhe_group_partition(const fhe_group* group, size_t batch, int rank, size_t* first, size_t* count);
/* fhe_launch on every device at once (PimManager::start_kernel with several devices): the descriptor's sizes are per
 * device, as PimData computes them (pim/PimData.h:281-290), heap_bases[rank] is that device's heap.  Blocking. */
int fhe_group_launch(fhe_group* group, int kernel_id, const void* desc, const fhe_dptr* heap_bases);
int fhe_group_ring_create(fhe_group* group, uint32_t N, uint32_t towers, const uint64_t* q, const uint64_t* psi,
                          fhe_group_ring** out);
int fhe_group_ring_destroy(fhe_group_ring* ring);
fhe_ring* fhe_group_ring_at(fhe_group_ring* ring, int rank); /* the ring on one device, for every single-device call */
/* one buffer per device, sized for its block of `batch` polynomials */
int fhe_group_alloc(fhe_group_ring* ring, size_t batch, fhe_dptr* out);
int fhe_group_free(fhe_group_ring* ring, const fhe_dptr* bufs);
/* host [batch][T][N] <-> the devices' blocks (asynchronous with pinned host memory; fhe_group_sync completes them) */
int fhe_group_scatter(fhe_group_ring* ring, const fhe_dptr* dst, const uint64_t* host, size_t batch, int stream);
int fhe_group_gather(fhe_group_ring* ring, uint64_t* host, const fhe_dptr* src, size_t batch, int stream);
/* fhe_dcrt_op / fhe_ntt_forward|inverse on every device's block, launched back to back (the devices run concurrently) */
int fhe_group_dcrt_op(fhe_group_ring* ring, int op, const fhe_dptr* dst, const fhe_dptr* a, const fhe_dptr* b,
                      size_t batch, int stream);
int fhe_group_ntt(fhe_group_ring* ring, const fhe_dptr* x, size_t batch, int inverse, int stream);
/* north_star's "NCCL over NVLink only to gather results": dst[rank] (batch polynomials each) receives every device's
 * block, HBM to HBM (ncclAllGather, or grouped ncclSend / ncclRecv when the blocks differ in size).  Asynchronous. */
int fhe_group_all_gather(fhe_group_ring* ring, const fhe_dptr* dst, const fhe_dptr* src, size_t batch, int stream);
/* fhe_dcrt_chain_host over all devices: each streams its block of the host batch through its own pipeline. Blocking. */
int fhe_group_chain_host(fhe_group_ring* ring, const int* ops, int nops, uint64_t* a, const uint64_t* b, size_t batch);

/* One rank of a NCCL communicator on a context's device, for one-process-per-GPU callers (torchrun): rank 0 calls
 * fhe_comm_unique_id (128 bytes) and hands the id to the other ranks by any means; every rank then calls
 * fhe_comm_create.  fhe_comm_all_gather: dst (nranks * words) receives every rank's `words` words, on `stream`. */
typedef struct fhe_comm fhe_comm;
int fhe_comm_unique_id(void* id, size_t cap);
int fhe_comm_create(fhe_ctx* ctx, int nranks, int rank, const void* id, fhe_comm** out);
int fhe_comm_destroy(fhe_comm* comm);
int fhe_comm_all_gather(fhe_comm* comm, fhe_dptr dst, fhe_dptr src, size_t words, int stream);

/* ---- diagnostics ---------------------------------------------------------------------------- */
/* integer-pipe probe: launches blocks x threads, each thread doing iters x 8 lazy Shoup products
 * (mulhi64 + 2 mullo64 = the multiply work of one NTT butterfly); returns the kernel time so the
 * caller can derive the Shoup-product rate the NTT roofline is quoted against. */
int fhe_imad_probe(fhe_ctx* ctx, int blocks, int threads, int iters, float* elapsed_ms);

#ifdef __cplusplus
}
#endif
#endif /* FHE_B200_H */
