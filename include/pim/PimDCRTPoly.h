/*
 * pim/PimDCRTPoly.h — device-resident batch of DCRTPolys with the operator surface of the
 * reference's DCRTPolyImpl on this path.  New header (the fork has no counterpart: its hook sits
 * one level lower, NativeVectorT::ModAddEq, mubintvecnat.cpp:250-259); it is what
 * DCRTPolyImpl::{Plus,Minus,Times,+=,-=,*=,SwitchFormat} dispatch to when a PimManager handle is
 * attached, one CUDA launch for all towers of all polynomials of the batch.
 *
 * Mirrors, with the same checks and the same exception kinds (OPENFHE_THROW(math_error /
 * not_implemented_error) become pim_math_error / pim_not_implemented_error):
 *   Plus / Minus / Times            lattice/hal/default/dcrtpoly.h:164-200, dcrtpoly-impl.h:398-407
 *   operator+= -= *= (no checks)    dcrtpoly-impl.h:410-416, 438-444, dcrtpoly.h:142-148
 *   SwitchFormat / SetFormat        dcrtpoly-impl.h:2518-2524, lattice/ilelement.h:448-451
 *   AutomorphismTransform           dcrtpoly-impl.h:348-369, poly-impl.h:312-380
 *   ApproxSwitchCRTBasis / ApproxModUp / ApproxModDown      dcrtpoly-impl.h:1034-1175
 *   DropLastElementAndScale / ModReduce                     dcrtpoly-impl.h:746-812
 * Layout in HBM: [batch][tower][N] uint64.  Header only; link with libfhe_b200.so.
 *
 * Several GPUs (PimManager with `gpus=k`): a PimRing then holds the ring's tables on every device and a PimDCRTPoly
 * holds one block of the batch per device (contiguous blocks, fhe_group_partition).  The hot-path methods — construction
 * from / sync to host data, copy, Plus / Minus / Times, += -= *=, SwitchFormat / SetFormat — run on all devices at once,
 * with no exchange between them; AllGather() is the optional NCCL gather of a result.  The methods beside
 * the path (automorphism, basis conversion, rescale, wire format) need a single-device manager and say so.
 */
#ifndef _PIM_DCRT_POLY_
#define _PIM_DCRT_POLY_

#include "PimManager.h"

#include <cstdint>
#include <memory>
#include <stdexcept>
#include <ostream>
#include <string>
#include <vector>

struct pim_math_error : std::runtime_error { using std::runtime_error::runtime_error; };
struct pim_not_implemented_error : std::runtime_error { using std::runtime_error::runtime_error; };

enum PimFormat { PIM_EVALUATION = 0, PIM_COEFFICIENT = 1 };   /* lbcrypto::Format values */

/* ILDCRTParams on the device: ring dimension, moduli, roots, twiddle tables (fhe_ring). */
class PimRing {
public:
  PimRing(PimManager *pim, uint32_t N, const std::vector<uint64_t> &q, const std::vector<uint64_t> &psi)
      : pim_(pim), n_(N), q_(q), psi_(psi) {
    if (!psi.empty() && psi.size() != q.size()) throw pim_math_error("PimRing: one root of unity per modulus");
    PimManager::check(fhe_group_ring_create(pim->group(), N, (uint32_t)q.size(), q_.data(),
                                            psi_.empty() ? nullptr : psi_.data(), &gring_), "PimRing");
    ring_ = fhe_group_ring_at(gring_, 0);
  }
  /* towers [first, first+count) of `parent`, sharing its device tables (the ILDCRTParams of a sub-basis) */
  static std::shared_ptr<PimRing> View(const std::shared_ptr<PimRing> &parent, uint32_t first, uint32_t count) {
    std::shared_ptr<PimRing> v(new PimRing(parent, first, count));
    return v;
  }
  ~PimRing() {
    if (gring_) fhe_group_ring_destroy(gring_);      /* owns the per-device rings */
    else fhe_ring_destroy(ring_);                    /* a view: only the handle */
  }
  PimRing(const PimRing &) = delete;
  PimRing &operator=(const PimRing &) = delete;

  /* the benchmark's recipe: poly-benchmark-16k.cpp:82-101 */
  static std::shared_ptr<PimRing> Generate(PimManager *pim, uint32_t nbits, uint32_t N, uint32_t towers) {
    std::vector<uint64_t> q(towers), psi(towers);
    PimManager::check(fhe_gen_params(nbits, N, towers, q.data(), psi.data()), "PimRing::Generate");
    return std::make_shared<PimRing>(pim, N, q, psi);
  }

  uint32_t GetRingDimension() const { return n_; }
  size_t GetNumTowers() const { return q_.size(); }
  const std::vector<uint64_t> &GetModuli() const { return q_; }
  const std::vector<uint64_t> &GetRoots() const { return psi_; }
  fhe_ring *handle() const { return ring_; }                      /* the ring on the first device */
  fhe_group_ring *group_handle() const { return gring_; }        /* every device's (null for a view) */
  uint32_t devices() const { return pim_->devices(); }
  PimManager *manager() const { return pim_; }

private:
  PimRing(const std::shared_ptr<PimRing> &parent, uint32_t first, uint32_t count)
      : pim_(parent->pim_), n_(parent->n_), parent_(parent) {
    if (pim_->devices() > 1) throw pim_not_implemented_error("PimRing::View needs a single-device PimManager");
    PimManager::check(fhe_ring_view(parent->ring_, first, count, &ring_), "PimRing::View");
    q_.assign(parent->q_.begin() + first, parent->q_.begin() + first + count);
    if (!parent->psi_.empty()) psi_.assign(parent->psi_.begin() + first, parent->psi_.begin() + first + count);
  }
  PimManager *pim_;
  uint32_t n_;
  std::vector<uint64_t> q_, psi_;
  fhe_ring *ring_ = nullptr;
  fhe_group_ring *gring_ = nullptr;
  std::shared_ptr<PimRing> parent_;   /* keeps the owner of the tables alive */
};

/* Tables of DCRTPolyImpl::ApproxSwitchCRTBasis from one basis to another (dcrtpoly-impl.h:1034-1082): what
 * CryptoParametersRNS precomputes per basis pair (rns-cryptoparameters.cpp:190-229, 273-283).  Pass the
 * scheme's own QHatInvModq / QHatModp, or leave them empty to have them derived from the moduli. */
class PimBasis {
public:
  PimBasis(std::shared_ptr<PimRing> from, std::shared_ptr<PimRing> to, const std::vector<uint64_t> &QHatInvModq = {},
           const std::vector<uint64_t> &QHatModp = {})
      : from_(std::move(from)), to_(std::move(to)) {
    PimManager::check(fhe_basis_create(from_->handle(), to_->handle(), QHatInvModq.empty() ? nullptr : QHatInvModq.data(),
                                       QHatModp.empty() ? nullptr : QHatModp.data(), &basis_), "PimBasis");
  }
  ~PimBasis() { fhe_basis_destroy(basis_); }
  PimBasis(const PimBasis &) = delete;
  PimBasis &operator=(const PimBasis &) = delete;
  fhe_basis *handle() const { return basis_; }
  const std::shared_ptr<PimRing> &From() const { return from_; }
  const std::shared_ptr<PimRing> &To() const { return to_; }

private:
  std::shared_ptr<PimRing> from_, to_;
  fhe_basis *basis_ = nullptr;
};

/* Constants of DropLastElementAndScale / ModReduce for a ring's last tower (ckksrns-cryptoparameters.cpp:66-88) */
class PimRescale {
public:
  explicit PimRescale(std::shared_ptr<PimRing> ring, const std::vector<uint64_t> &QlQlInvModqlDivqlModq = {},
                      const std::vector<uint64_t> &qlInvModq = {})
      : ring_(std::move(ring)), rest_(PimRing::View(ring_, 0, (uint32_t)ring_->GetNumTowers() - 1)) {
    PimManager::check(fhe_rescale_create(ring_->handle(), QlQlInvModqlDivqlModq.empty() ? nullptr : QlQlInvModqlDivqlModq.data(),
                                         qlInvModq.empty() ? nullptr : qlInvModq.data(), &plan_), "PimRescale");
  }
  ~PimRescale() { fhe_rescale_destroy(plan_); }
  PimRescale(const PimRescale &) = delete;
  PimRescale &operator=(const PimRescale &) = delete;
  fhe_rescale *handle() const { return plan_; }
  const std::shared_ptr<PimRing> &Ring() const { return ring_; }
  const std::shared_ptr<PimRing> &Rest() const { return rest_; }   /* params after DropLastElement */

private:
  std::shared_ptr<PimRing> ring_, rest_;
  fhe_rescale *plan_ = nullptr;
};

/* Tables of DCRTPolyImpl::ScaleAndRound(paramsOutput, ...) (dcrtpoly-impl.h:1876-1960): `ring` is the input basis, the
 * output basis its last (outLast) or first sizeOut towers; frac / modo are the scheme's tOSHatInvModsDivsFrac /
 * tOSHatInvModsDivsModo (row-major [sizeOut][sizeIn+1], bfvrns-cryptoparameters.cpp:278-311). */
class PimScaleRound {
public:
  PimScaleRound(std::shared_ptr<PimRing> ring, uint32_t sizeOut, bool outLast, const std::vector<double> &frac,
                const std::vector<uint64_t> &modo)
      : ring_(std::move(ring)),
        out_(PimRing::View(ring_, outLast ? (uint32_t)ring_->GetNumTowers() - sizeOut : 0, sizeOut)) {
    if (frac.size() != ring_->GetNumTowers() - sizeOut || modo.size() != (size_t)sizeOut * (frac.size() + 1))
      throw pim_math_error("PimScaleRound: table sizes do not match the bases");
    PimManager::check(fhe_scale_round_create(ring_->handle(), sizeOut, outLast ? 1 : 0, frac.data(), modo.data(), &plan_),
                      "PimScaleRound");
  }
  ~PimScaleRound() { fhe_scale_round_destroy(plan_); }
  PimScaleRound(const PimScaleRound &) = delete;
  PimScaleRound &operator=(const PimScaleRound &) = delete;
  fhe_scale_round *handle() const { return plan_; }
  const std::shared_ptr<PimRing> &Out() const { return out_; }

private:
  std::shared_ptr<PimRing> ring_, out_;
  fhe_scale_round *plan_ = nullptr;
};

/* One constant per tower: m_QlHatModq[l] for ExpandCRTBasisQlHat, pInvModq for ScaleAndRoundPOverQ
 * (dcrtpoly-impl.h:1514-1534, 2052-2066; BFV multTech HPSPOVERQ / HPSPOVERQLEVELED). */
class PimTowerScale {
public:
  PimTowerScale(std::shared_ptr<PimRing> ring, const std::vector<uint64_t> &c) : ring_(std::move(ring)) {
    if (c.size() != ring_->GetNumTowers()) throw pim_math_error("PimTowerScale: one constant per tower");
    PimManager::check(fhe_tower_scale_create(ring_->handle(), c.data(), &plan_), "PimTowerScale");
  }
  ~PimTowerScale() { fhe_tower_scale_destroy(plan_); }
  PimTowerScale(const PimTowerScale &) = delete;
  PimTowerScale &operator=(const PimTowerScale &) = delete;
  fhe_tower_scale *handle() const { return plan_; }
  const std::shared_ptr<PimRing> &Ring() const { return ring_; }

private:
  std::shared_ptr<PimRing> ring_;
  fhe_tower_scale *plan_ = nullptr;
};

/* Hybrid key switching over a ring Ql u P (KeySwitchHYBRID::KeySwitchCore, src/pke/lib/keyswitch/keyswitch-hybrid.cpp:
 * 324-480): digit tables over the (possibly truncated) partitions of Ql and the ModDown tables, derived from the moduli. */
class PimKeySwitch {
public:
  PimKeySwitch(std::shared_ptr<PimRing> ring_qlp, uint32_t size_ql, uint32_t alpha) : ring_(std::move(ring_qlp)) {
    PimManager::check(fhe_keyswitch_create(ring_->handle(), size_ql, alpha, &plan_), "PimKeySwitch");
  }
  ~PimKeySwitch() { fhe_keyswitch_destroy(plan_); }
  PimKeySwitch(const PimKeySwitch &) = delete;
  PimKeySwitch &operator=(const PimKeySwitch &) = delete;
  fhe_keyswitch *handle() const { return plan_; }
  uint32_t Digits() const { return fhe_keyswitch_digits(plan_); }
  const std::shared_ptr<PimRing> &Ring() const { return ring_; }

private:
  std::shared_ptr<PimRing> ring_;
  fhe_keyswitch *plan_ = nullptr;
};

/* DCRTPolyImpl::CRTDecompose(0) over a ring (dcrtpoly-impl.h:266-288): the digits of KeySwitchBV */
class PimCrtDecomp {
public:
  explicit PimCrtDecomp(std::shared_ptr<PimRing> ring) : ring_(std::move(ring)) {
    PimManager::check(fhe_crt_decompose_create(ring_->handle(), &plan_), "PimCrtDecomp");
  }
  ~PimCrtDecomp() { fhe_crt_decompose_destroy(plan_); }
  PimCrtDecomp(const PimCrtDecomp &) = delete;
  PimCrtDecomp &operator=(const PimCrtDecomp &) = delete;
  fhe_crt_decomp *handle() const { return plan_; }
  const std::shared_ptr<PimRing> &Ring() const { return ring_; }

private:
  std::shared_ptr<PimRing> ring_;
  fhe_crt_decomp *plan_ = nullptr;
};

/* The BEHZ set over a ring Q u Bsk (dcrtpoly-impl.h:2005-2050, 2069-2420; BFV multTech BEHZ): every table of
 * bfvrns-cryptoparameters.cpp:724-891 derived from the moduli and the plaintext modulus t. */
class PimBehz {
public:
  PimBehz(std::shared_ptr<PimRing> ring_qbsk, uint32_t size_q, uint64_t t) : ring_(std::move(ring_qbsk)), size_q_(size_q), t_(t) {
    PimManager::check(fhe_behz_create(ring_->handle(), size_q, t, &plan_), "PimBehz");
  }
  ~PimBehz() { fhe_behz_destroy(plan_); }
  PimBehz(const PimBehz &) = delete;
  PimBehz &operator=(const PimBehz &) = delete;
  fhe_behz *handle() const { return plan_; }
  const std::shared_ptr<PimRing> &Ring() const { return ring_; }
  uint32_t SizeQ() const { return size_q_; }
  uint64_t PlaintextModulus() const { return t_; }
  std::vector<uint64_t> Table(int which) const {       /* numbering of fhe_b200.h */
    size_t n = 0;
    PimManager::check(fhe_behz_table(plan_, which, nullptr, &n), "PimBehz::Table");
    std::vector<uint64_t> v(n);
    PimManager::check(fhe_behz_table(plan_, which, v.data(), nullptr), "PimBehz::Table");
    return v;
  }

private:
  std::shared_ptr<PimRing> ring_;
  uint32_t size_q_;
  uint64_t t_;
  fhe_behz *plan_ = nullptr;
};

class PimDCRTPoly {
public:
  PimDCRTPoly(std::shared_ptr<PimRing> ring, size_t batch, PimFormat format, const uint64_t *host = nullptr,
              int stream = 0)
      : ring_(std::move(ring)), batch_(batch), format_(format), stream_(stream) {
    allocate();
    if (host) {
      if (multi()) PimManager::check(fhe_group_scatter(ring_->group_handle(), dv_.data(), host, batch_, stream_), "PimDCRTPoly upload");
      else PimManager::check(fhe_h2d(ctx(), d_, host, bytes(), stream_), "PimDCRTPoly upload");
      sync_streams("PimDCRTPoly upload");
    }
  }
  PimDCRTPoly(const PimDCRTPoly &o) : PimDCRTPoly(o.ring_, o.batch_, o.format_, nullptr, o.stream_) {
    for (uint32_t i = 0; i < devices(); i++)
      PimManager::check(fhe_d2d(ctx(i), ptr(i), o.ptr(i), block_bytes(i), stream_), "PimDCRTPoly copy");
  }
  PimDCRTPoly(PimDCRTPoly &&o) noexcept
      : ring_(std::move(o.ring_)), batch_(o.batch_), format_(o.format_), stream_(o.stream_), d_(o.d_), dv_(std::move(o.dv_)) {
    o.d_ = 0;
    o.dv_.clear();
  }
  PimDCRTPoly &operator=(PimDCRTPoly o) {
    std::swap(ring_, o.ring_); std::swap(batch_, o.batch_); std::swap(format_, o.format_);
    std::swap(stream_, o.stream_); std::swap(d_, o.d_); std::swap(dv_, o.dv_);
    return *this;
  }
  ~PimDCRTPoly() {
    if (!ring_) return;
    if (!dv_.empty()) fhe_group_free(ring_->group_handle(), dv_.data());
    else if (d_) fhe_free(ctx(), d_);
  }

  PimFormat GetFormat() const { return format_; }
  size_t GetBatch() const { return batch_; }
  size_t words() const { return batch_ * ring_->GetNumTowers() * ring_->GetRingDimension(); }
  size_t bytes() const { return words() * sizeof(uint64_t); }
  fhe_dptr device_ptr() const { return d_; }                       /* the block on the first device */
  fhe_dptr device_ptr(uint32_t rank) const { return ptr(rank); }
  uint32_t devices() const { return ring_->devices(); }
  /* polynomials [first, first + count) of the batch live on device `rank` */
  void GetBlock(uint32_t rank, size_t *first, size_t *count) const {
    PimManager::check(fhe_group_partition(ring_->manager()->group(), batch_, (int)rank, first, count), "GetBlock");
  }
  const std::shared_ptr<PimRing> &GetParams() const { return ring_; }

  /* device -> host, blocking (the PimData::sync of this handle) */
  void sync(uint64_t *out) const {
    if (multi()) PimManager::check(fhe_group_gather(ring_->group_handle(), out, dv_.data(), batch_, stream_), "PimDCRTPoly::sync");
    else PimManager::check(fhe_d2h(ctx(), out, d_, bytes(), stream_), "PimDCRTPoly::sync");
    sync_streams("PimDCRTPoly::sync");
  }
  std::vector<uint64_t> GetValues() const { std::vector<uint64_t> v(words()); sync(v.data()); return v; }

  /* north_star's optional gather: every device receives the whole batch, HBM to HBM over NVLink (NCCL on the
   * library's own streams).  Returns one device pointer per device (batch polynomials each); the caller frees them
   * with fhe_free on context(rank).  Complete after manager()->sync(). */
  std::vector<fhe_dptr> AllGather() const {
    std::vector<fhe_dptr> out(devices(), 0);
    for (uint32_t i = 0; i < devices(); i++) PimManager::check(fhe_alloc(ctx(i), bytes() ? bytes() : 8, &out[i]), "AllGather");
    std::vector<fhe_dptr> src(devices());
    for (uint32_t i = 0; i < devices(); i++) src[i] = ptr(i);
    PimManager::check(fhe_group_all_gather(ring_->group_handle(), out.data(), src.data(), batch_, stream_), "AllGather");
    return out;
  }

  /* Wire format: `batch` consecutive streams, each what Serial::Serialize(poly, os, SerType::BINARY) writes for the
   * DCRTPoly it stands for (utils/serial.h:99-103, dcrtpoly.h:349-354).  Blocking. */
  std::string Serialize() const {
    single("Serialize");
    std::string blob(fhe_serial_bytes(ring_->handle()) * batch_, '\0');
    PimManager::check(fhe_serialize(ring_->handle(), &blob[0], d_, batch_, (int)format_, stream_), "Serialize");
    PimManager::check(fhe_sync(ctx(), stream_), "Serialize");
    return blob;
  }
  void Serialize(std::ostream &os) const { std::string b = Serialize(); os.write(b.data(), (std::streamsize)b.size()); }
  /* Serial::Deserialize (utils/serial.h:111-115) of `batch` consecutive streams into a device-resident batch; throws
   * when a stream is malformed or was written for another ring.  *consumed = bytes read. */
  static PimDCRTPoly Deserialize(std::shared_ptr<PimRing> ring, const void *blob, size_t len, size_t batch,
                                 size_t *consumed = nullptr, int stream = 0) {
    if (ring->devices() > 1) throw pim_not_implemented_error("Deserialize needs a single-device PimManager");
    PimDCRTPoly p(std::move(ring), batch, PIM_EVALUATION, nullptr, stream);
    int fmt = 0;
    PimManager::check(fhe_deserialize(p.ring_->handle(), p.d_, blob, len, batch, &fmt, consumed, stream), "Deserialize");
    PimManager::check(fhe_sync(p.ctx(), stream), "Deserialize");
    p.format_ = fmt == 0 ? PIM_EVALUATION : PIM_COEFFICIENT;
    return p;
  }

  PimDCRTPoly Plus(const PimDCRTPoly &rhs) const { check(rhs, "add", false); return binary(rhs, FHE_OP_ADD); }
  PimDCRTPoly Minus(const PimDCRTPoly &rhs) const { check(rhs, "subtract", false); return binary(rhs, FHE_OP_SUB); }
  PimDCRTPoly Times(const PimDCRTPoly &rhs) const { check(rhs, "multiply", true); return binary(rhs, FHE_OP_MUL); }
  PimDCRTPoly operator+(const PimDCRTPoly &rhs) const { return Plus(rhs); }
  PimDCRTPoly operator-(const PimDCRTPoly &rhs) const { return Minus(rhs); }
  PimDCRTPoly operator*(const PimDCRTPoly &rhs) const { return Times(rhs); }
  PimDCRTPoly &operator+=(const PimDCRTPoly &rhs) { return inplace(rhs, FHE_OP_ADD); }
  PimDCRTPoly &operator-=(const PimDCRTPoly &rhs) { return inplace(rhs, FHE_OP_SUB); }
  PimDCRTPoly &operator*=(const PimDCRTPoly &rhs) { return inplace(rhs, FHE_OP_MUL); }

  void SwitchFormat() {
    const int inverse = format_ == PIM_COEFFICIENT ? 0 : 1;
    if (multi()) PimManager::check(fhe_group_ntt(ring_->group_handle(), dv_.data(), batch_, inverse, stream_), "SwitchFormat");
    else if (inverse) PimManager::check(fhe_ntt_inverse(ring_->handle(), d_, batch_, stream_), "SwitchFormat");
    else PimManager::check(fhe_ntt_forward(ring_->handle(), d_, batch_, stream_), "SwitchFormat");
    format_ = inverse ? PIM_COEFFICIENT : PIM_EVALUATION;
  }
  /* dcrtpoly-impl.h:348-358 -> poly-impl.h:312-365; k even throws math_error there */
  PimDCRTPoly AutomorphismTransform(uint32_t k) const {
    single("AutomorphismTransform");
    if (k % 2 == 0) throw pim_math_error("Automorphism index not odd\n");
    PimDCRTPoly out(ring_, batch_, format_, nullptr, stream_);
    PimManager::check(fhe_automorphism(ring_->handle(), out.d_, d_, batch_, k, (int)format_, stream_), "AutomorphismTransform");
    return out;
  }
  /* dcrtpoly-impl.h:360-369 -> poly-impl.h:368-380; `precomp` = device copy of PrecomputeAutoMap's table */
  PimDCRTPoly AutomorphismTransform(uint32_t k, fhe_dptr precomp) const {
    single("AutomorphismTransform");
    if (format_ != PIM_EVALUATION)
      throw pim_not_implemented_error("Automorphism Poly Format not EVALUATION or not power-of-two");
    if (k % 2 == 0) throw pim_math_error("Automorphism index not odd\n");
    PimDCRTPoly out(ring_, batch_, format_, nullptr, stream_);
    PimManager::check(fhe_automorphism_map(ring_->handle(), out.d_, d_, batch_, precomp, stream_), "AutomorphismTransform");
    return out;
  }
  void SetFormat(PimFormat f) { if (format_ != f) SwitchFormat(); }

  /* dcrtpoly-impl.h:1034-1082: this over basis.From() -> result over basis.To(), same format flag */
  PimDCRTPoly ApproxSwitchCRTBasis(const PimBasis &basis) const {
    single("SwitchCRTBasis");
    single("ApproxSwitchCRTBasis");
    PimDCRTPoly out(basis.To(), batch_, format_, nullptr, stream_);
    PimManager::check(fhe_approx_switch_crt_basis(basis.handle(), out.d_, d_, batch_, stream_), "ApproxSwitchCRTBasis");
    return out;
  }
  /* dcrtpoly-impl.h:1085-1133 (the reference works in place; here the extended polynomial is returned): this over Q,
   * qToP from Q to P, ringQP the ring of Q followed by P.  Result in EVALUATION. */
  PimDCRTPoly ApproxModUp(const PimBasis &qToP, const std::shared_ptr<PimRing> &ringQP) const {
    single("ApproxModUp");
    PimDCRTPoly out(ringQP, batch_, PIM_EVALUATION, nullptr, stream_);
    PimManager::check(fhe_approx_mod_up(qToP.handle(), out.d_, d_, batch_, (int)format_, stream_), "ApproxModUp");
    return out;
  }
  /* dcrtpoly-impl.h:1136-1175: this over QP in EVALUATION, pToQ from P to Q; t > 0 selects the BGV form */
  PimDCRTPoly ApproxModDown(const PimBasis &pToQ, uint64_t t = 0) const {
    single("ApproxModDown");
    if (format_ != PIM_EVALUATION) throw pim_not_implemented_error("ApproxModDown expects Format::EVALUATION");
    PimDCRTPoly out(pToQ.To(), batch_, PIM_EVALUATION, nullptr, stream_);
    PimManager::check(fhe_approx_mod_down(pToQ.handle(), out.d_, d_, batch_, t, stream_), "ApproxModDown");
    return out;
  }
  /* dcrtpoly-impl.h:1178-1230: the exact switch (BFV) */
  PimDCRTPoly SwitchCRTBasis(const PimBasis &basis) const {
    PimDCRTPoly out(basis.To(), batch_, format_, nullptr, stream_);
    PimManager::check(fhe_switch_crt_basis(basis.handle(), out.d_, d_, batch_, stream_), "SwitchCRTBasis");
    return out;
  }
  /* dcrtpoly-impl.h:1311-1358 (in place in the reference) */
  PimDCRTPoly ExpandCRTBasis(const PimBasis &qToP, const std::shared_ptr<PimRing> &ringQP, PimFormat resultFormat) const {
    single("ExpandCRTBasis");
    PimDCRTPoly out(ringQP, batch_, resultFormat, nullptr, stream_);
    PimManager::check(fhe_expand_crt_basis(qToP.handle(), out.d_, d_, batch_, (int)format_, (int)resultFormat, stream_),
                      "ExpandCRTBasis");
    return out;
  }
  /* dcrtpoly-impl.h:1361-1410 (in place in the reference): ringPQ holds the P towers first */
  PimDCRTPoly ExpandCRTBasisReverseOrder(const PimBasis &qToP, const std::shared_ptr<PimRing> &ringPQ,
                                         PimFormat resultFormat) const {
    single("ExpandCRTBasisReverseOrder");
    PimDCRTPoly out(ringPQ, batch_, resultFormat, nullptr, stream_);
    PimManager::check(fhe_expand_crt_basis_reverse_order(qToP.handle(), out.d_, d_, batch_, (int)format_, (int)resultFormat,
                                                         stream_), "ExpandCRTBasisReverseOrder");
    return out;
  }
  /* dcrtpoly-impl.h:1413-1468 (in place in the reference): qToPl carries mPlQHatInvModq / qInvModp, plToQl is the exact
   * switch back; ringQlPl = [Ql | Pl] */
  PimDCRTPoly FastExpandCRTBasisPloverQ(const PimBasis &qToPl, const PimBasis &plToQl,
                                        const std::shared_ptr<PimRing> &ringQlPl) const {
    single("FastExpandCRTBasisPloverQ");
    PimDCRTPoly out(ringQlPl, batch_, format_, nullptr, stream_);
    PimManager::check(fhe_fast_expand_crt_basis_pl_over_q(qToPl.handle(), plToQl.handle(), out.d_, d_, batch_, stream_),
                      "FastExpandCRTBasisPloverQ");
    return out;
  }
  /* dcrtpoly-impl.h:1514-1534 (in place in the reference): this over the first towers of plan.Ring() */
  PimDCRTPoly ExpandCRTBasisQlHat(const PimTowerScale &plan) const {
    single("ExpandCRTBasisQlHat");
    PimDCRTPoly out(plan.Ring(), batch_, format_, nullptr, stream_);
    PimManager::check(fhe_expand_crt_basis_ql_hat(plan.handle(), (uint32_t)ring_->GetNumTowers(), out.d_, d_, batch_, stream_),
                      "ExpandCRTBasisQlHat");
    return out;
  }
  /* dcrtpoly-impl.h:617-627: Times(const std::vector<NativeInteger>&), one constant per tower */
  PimDCRTPoly Times(const PimTowerScale &plan) const {
    single("Times");
    if (ring_->GetNumTowers() != plan.Ring()->GetNumTowers()) throw pim_math_error("tower size mismatch; cannot multiply");
    PimDCRTPoly out(ring_, batch_, format_, nullptr, stream_);
    PimManager::check(fhe_dcrt_times_towers(plan.handle(), out.d_, d_, batch_, stream_), "Times");
    return out;
  }
  /* dcrtpoly-impl.h:2052-2066 (in place in the reference): this over plan.Ring() plus one tower */
  PimDCRTPoly ScaleAndRoundPOverQ(const PimTowerScale &plan) const {
    single("ScaleAndRoundPOverQ");
    if (ring_->GetNumTowers() != plan.Ring()->GetNumTowers() + 1)
      throw pim_math_error("ScaleAndRoundPOverQ: the input must have one tower more than the output basis");
    PimDCRTPoly out(plan.Ring(), batch_, format_, nullptr, stream_);
    PimManager::check(fhe_scale_and_round_p_over_q(plan.handle(), out.d_, d_, batch_, stream_), "ScaleAndRoundPOverQ");
    return out;
  }
  /* dcrtpoly-impl.h:1876-1960 */
  PimDCRTPoly ScaleAndRound(const PimScaleRound &plan) const {
    single("ScaleAndRound");
    PimDCRTPoly out(plan.Out(), batch_, format_, nullptr, stream_);
    PimManager::check(fhe_scale_and_round(plan.handle(), out.d_, d_, batch_, stream_), "ScaleAndRound");
    return out;
  }
  /* dcrtpoly-impl.h:746-769 (in place in the reference).  As there, a COEFFICIENT input comes back with forward-
   * transformed towers while the format flag stays COEFFICIENT (dcrtpoly-impl.h:765-766). */
  PimDCRTPoly DropLastElementAndScale(const PimRescale &plan) const {
    single("DropLastElementAndScale");
    PimDCRTPoly out(plan.Rest(), batch_, format_, nullptr, stream_);
    PimManager::check(fhe_drop_last_element_and_scale(plan.handle(), out.d_, d_, batch_, (int)format_, stream_),
                      "DropLastElementAndScale");
    return out;
  }
  /* dcrtpoly-impl.h:792-812 */
  PimDCRTPoly ModReduce(const PimRescale &plan, uint64_t t) const {
    single("ModReduce");
    PimDCRTPoly out(plan.Rest(), batch_, format_, nullptr, stream_);
    PimManager::check(fhe_mod_reduce(plan.handle(), out.d_, d_, batch_, t, (int)format_, stream_), "ModReduce");
    return out;
  }

private:
  fhe_ctx *ctx(uint32_t rank = 0) const { return ring_->manager()->context(rank); }
  bool multi() const { return !dv_.empty(); }
  fhe_dptr ptr(uint32_t rank) const { return multi() ? dv_[rank] : d_; }
  size_t block_bytes(uint32_t rank) const {
    if (!multi()) return bytes();
    size_t first, count;
    GetBlock(rank, &first, &count);
    return count * ring_->GetNumTowers() * ring_->GetRingDimension() * sizeof(uint64_t);
  }
  void allocate() {
    if (ring_->devices() > 1) {
      if (!ring_->group_handle()) throw pim_not_implemented_error("PimDCRTPoly over a ring view needs a single-device PimManager");
      dv_.assign(ring_->devices(), 0);
      PimManager::check(fhe_group_alloc(ring_->group_handle(), batch_, dv_.data()), "PimDCRTPoly");
      d_ = dv_[0];
    } else {
      PimManager::check(fhe_alloc(ctx(), bytes() ? bytes() : 8, &d_), "PimDCRTPoly");
    }
  }
  void sync_streams(const char *what) const {
    for (uint32_t i = 0; i < devices(); i++) PimManager::check(fhe_sync(ctx(i), stream_), what);
  }
  void single(const char *what) const {
    if (multi()) throw pim_not_implemented_error(std::string(what) + " needs a single-device PimManager (the batch is split over " +
                                                 std::to_string(devices()) + " devices)");
  }

  void check(const PimDCRTPoly &rhs, const char *verb, bool need_eval) const {
    if (ring_->GetRingDimension() != rhs.ring_->GetRingDimension()) throw pim_math_error("RingDimension missmatch");
    if (need_eval) {
      if (format_ != PIM_EVALUATION || rhs.format_ != PIM_EVALUATION)
        throw pim_not_implemented_error("operator* for DCRTPolyImpl supported only in Format::EVALUATION");
    } else if (format_ != rhs.format_) {
      throw pim_not_implemented_error("Format missmatch");
    }
    if (ring_->GetNumTowers() != rhs.ring_->GetNumTowers())
      throw pim_math_error(std::string("tower size mismatch; cannot ") + verb);
    if (ring_->GetModuli()[0] != rhs.ring_->GetModuli()[0]) throw pim_math_error("Modulus missmatch");
    if (batch_ != rhs.batch_) throw pim_math_error("batch size mismatch");
  }
  PimDCRTPoly binary(const PimDCRTPoly &rhs, int op) const {
    PimDCRTPoly out(ring_, batch_, format_, nullptr, stream_);
    run_op(op, out, rhs);
    return out;
  }
  PimDCRTPoly &inplace(const PimDCRTPoly &rhs, int op) {
    run_op(op, *this, rhs);
    return *this;
  }
  void run_op(int op, const PimDCRTPoly &dst, const PimDCRTPoly &rhs) const {
    if (multi())
      PimManager::check(fhe_group_dcrt_op(ring_->group_handle(), op, dst.dv_.data(), dv_.data(), rhs.dv_.data(), batch_, stream_),
                        "DCRTPoly op");
    else
      PimManager::check(fhe_dcrt_op(ring_->handle(), op, dst.d_, d_, rhs.d_, batch_, stream_), "DCRTPoly op");
  }

  std::shared_ptr<PimRing> ring_;
  size_t batch_;
  PimFormat format_;
  int stream_;
  fhe_dptr d_ = 0;                 /* single device: the batch; several: the first device's block (= dv_[0]) */
  std::vector<fhe_dptr> dv_;       /* several devices: one block per device */
};

#endif /* _PIM_DCRT_POLY_ */
