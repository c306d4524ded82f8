/*
 * pim/PimHook.h — the body of the live intercept: what DCRTPolyImpl::operator+= / -= / *= and
 * DCRTPolyImpl::SwitchFormat call when a PimManager handle is attached (SetPim).
 *
 * The fork only sketched this call (`pim->Run_On_Pim(this, b)` in a comment,
 * src/core/lib/math/hal/intnat/mubintvecnat.cpp:250-259) one level lower, per NativeVector; here the
 * intercept sits at DCRTPolyImpl level so that all towers of a polynomial go to the GPU in ONE launch
 * (dcrtpoly-impl.h:410-416, 438-444, 2518-2524; dcrtpoly.h:142-148).  The patch that inserts the calls
 * is oracle/hook_patch.py (applied to a temporary copy of the reference headers).
 *
 * The templates are duck-typed on OpenFHE's PolyImpl<NativeVector> (GetValues, GetModulus,
 * GetRootOfUnity, GetRingDimension, SetValues) so this header needs no OpenFHE include.  This is the
 * copy-per-call form of the hook (operands live in host memory between calls, like the fork's);
 * PimDCRTPoly.h is the resident form.  Failures throw: there is no silent CPU fallback.
 */
#ifndef _PIM_HOOK_
#define _PIM_HOOK_

#include "PimDCRTPoly.h"

#include <map>
#include <memory>
#include <mutex>
#include <tuple>
#include <vector>

namespace pimhook {

/* device rings are cached per (N, moduli, roots): fhe_ring_create builds the twiddle tables once, like
 * the reference's static table cache keyed by modulus (transformnat-impl.h:595-598) */
inline std::shared_ptr<PimRing> ring_for(PimManager *pim, uint32_t n, const std::vector<uint64_t> &q,
                                         const std::vector<uint64_t> &psi) {
  static std::mutex mu;
  static std::map<std::tuple<uint32_t, std::vector<uint64_t>, std::vector<uint64_t>>, std::shared_ptr<PimRing>> cache;
  std::lock_guard<std::mutex> lock(mu);
  auto key = std::make_tuple(n, q, psi);
  auto it = cache.find(key);
  if (it != cache.end()) return it->second;
  auto ring = std::make_shared<PimRing>(pim, n, q, psi);
  cache.emplace(std::move(key), ring);
  return ring;
}

template <class Towers>
inline std::shared_ptr<PimRing> ring_of(PimManager *pim, const Towers &towers) {
  std::vector<uint64_t> q, psi;
  for (const auto &t : towers) {
    q.push_back(t.GetModulus().ConvertToInt());
    psi.push_back(t.GetRootOfUnity().ConvertToInt());
  }
  return ring_for(pim, (uint32_t)towers[0].GetRingDimension(), q, psi);
}

template <class Vec>
inline const uint64_t *words(const Vec &v) {
  static_assert(sizeof(v[0]) == sizeof(uint64_t), "NativeInteger is expected to be one 64-bit word (ubintnat.h:141)");
  return reinterpret_cast<const uint64_t *>(&v[0]);
}

struct DeviceBuf {   /* arena block, released on scope exit */
  fhe_ctx *ctx;
  fhe_dptr p = 0;
  DeviceBuf(fhe_ctx *c, size_t bytes) : ctx(c) { PimManager::check(fhe_alloc(c, bytes, &p), "pim hook: device buffer"); }
  ~DeviceBuf() { fhe_free(ctx, p); }
};

/* a (op)= b over all towers: op is FHE_OP_ADD / FHE_OP_SUB / FHE_OP_MUL.  Returns true when done on the GPU. */
template <class Towers>
inline bool dcrt_eq(PimManager *pim, Towers &a, const Towers &b, int op) {
  if (!pim || a.empty() || a.size() != b.size()) return false;
  const size_t T = a.size(), N = a[0].GetRingDimension(), bytes = N * sizeof(uint64_t);
  auto ring = ring_of(pim, a);
  fhe_ctx *ctx = pim->context();
  DeviceBuf da(ctx, T * bytes), db(ctx, T * bytes);
  for (size_t t = 0; t < T; t++) {
    PimManager::check(fhe_h2d(ctx, da.p + t * bytes, words(a[t].GetValues()), bytes, 0), "pim hook: upload");
    PimManager::check(fhe_h2d(ctx, db.p + t * bytes, words(b[t].GetValues()), bytes, 0), "pim hook: upload");
  }
  PimManager::check(fhe_dcrt_op(ring->handle(), op, da.p, da.p, db.p, 1, 0), "pim hook: dcrt op");
  for (size_t t = 0; t < T; t++)
    PimManager::check(fhe_d2h(ctx, const_cast<uint64_t *>(words(a[t].GetValues())), da.p + t * bytes, bytes, 0),
                      "pim hook: download");
  PimManager::check(fhe_sync(ctx, 0), "pim hook: sync");
  return true;
}

/* forward (to_evaluation) or inverse transform of all towers; every tower is left in `new_format`. */
template <class Towers, class Format>
inline bool dcrt_switch_format(PimManager *pim, Towers &a, bool to_evaluation, Format new_format) {
  if (!pim || a.empty()) return false;
  const size_t T = a.size(), N = a[0].GetRingDimension(), bytes = N * sizeof(uint64_t);
  auto ring = ring_of(pim, a);
  fhe_ctx *ctx = pim->context();
  DeviceBuf d(ctx, T * bytes);
  for (size_t t = 0; t < T; t++)
    PimManager::check(fhe_h2d(ctx, d.p + t * bytes, words(a[t].GetValues()), bytes, 0), "pim hook: upload");
  PimManager::check(to_evaluation ? fhe_ntt_forward(ring->handle(), d.p, 1, 0) : fhe_ntt_inverse(ring->handle(), d.p, 1, 0),
                    "pim hook: transform");
  PimManager::check(fhe_sync(ctx, 0), "pim hook: sync");
  for (size_t t = 0; t < T; t++) {
    auto v = a[t].GetValues();                      /* copy: same length and modulus */
    PimManager::check(fhe_d2h(ctx, const_cast<uint64_t *>(words(v)), d.p + t * bytes, bytes, 0), "pim hook: download");
    PimManager::check(fhe_sync(ctx, 0), "pim hook: sync");
    a[t].SetValues(std::move(v), new_format);       /* sets the tower's format without transforming on the CPU */
  }
  return true;
}

}  // namespace pimhook

#endif /* _PIM_HOOK_ */
