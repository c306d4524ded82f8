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
 * GetRootOfUnity, GetRingDimension, SetValues, OverrideFormat) so this header needs no OpenFHE include.  Two forms:
 * copy-per-call (dcrt_eq / dcrt_switch_format: operands live in host memory between calls, like the fork's; needs
 * no change to the class layout) and resident (Resident / resident_eq / resident_switch_format below: the towers
 * stay in HBM between hooked operators).  Failures throw: there is no silent CPU fallback.
 */
#ifndef _PIM_HOOK_
#define _PIM_HOOK_

#include "PimDCRTPoly.h"

#include <algorithm>
#include <atomic>
#include <cstdio>
#include <cstring>
#include <cstdlib>
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

/* Streams.  The reference's callers are OpenMP loops (dcrtpoly-impl.h:412; the schemes' loops over ciphertexts), so the
 * hook is entered from many host threads at once.  Every host thread enqueues on ITS OWN stream of the context's pool
 * (round robin over the pool), with its own pinned staging block, so the uploads, kernels and downloads of different
 * threads overlap instead of queueing behind one another on stream 0.  A device copy remembers the stream that wrote
 * it last; a thread that meets a copy written on another stream waits for that stream first (operands handed from
 * one thread to another are rare: a ciphertext is normally processed by one thread from start to finish).
 * FHE_B200_HOOK_ONE_STREAM=1 restores the single-stream behaviour. */
inline int hook_stream(fhe_ctx *ctx) {
  static const bool one = [] { const char *e = std::getenv("FHE_B200_HOOK_ONE_STREAM"); return e && *e && *e != '0'; }();
  if (one) return 0;
  static std::atomic<int> next{0};
  thread_local int mine = -1;
  if (mine < 0) mine = next.fetch_add(1);
  const int n = fhe_num_streams(ctx);
  return n > 0 ? mine % n : 0;
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
  const int S = hook_stream(ctx);
  DeviceBuf da(ctx, T * bytes), db(ctx, T * bytes);
  for (size_t t = 0; t < T; t++) {
    PimManager::check(fhe_h2d(ctx, da.p + t * bytes, words(a[t].GetValues()), bytes, S), "pim hook: upload");
    PimManager::check(fhe_h2d(ctx, db.p + t * bytes, words(b[t].GetValues()), bytes, S), "pim hook: upload");
  }
  PimManager::check(fhe_dcrt_op(ring->handle(), op, da.p, da.p, db.p, 1, S), "pim hook: dcrt op");
  for (size_t t = 0; t < T; t++)
    PimManager::check(fhe_d2h(ctx, const_cast<uint64_t *>(words(a[t].GetValues())), da.p + t * bytes, bytes, S),
                      "pim hook: download");
  PimManager::check(fhe_sync(ctx, S), "pim hook: sync");
  return true;
}

/* forward (to_evaluation) or inverse transform of all towers; every tower is left in `new_format`. */
template <class Towers, class Format>
inline bool dcrt_switch_format(PimManager *pim, Towers &a, bool to_evaluation, Format new_format) {
  if (!pim || a.empty()) return false;
  const size_t T = a.size(), N = a[0].GetRingDimension(), bytes = N * sizeof(uint64_t);
  auto ring = ring_of(pim, a);
  fhe_ctx *ctx = pim->context();
  const int S = hook_stream(ctx);
  DeviceBuf d(ctx, T * bytes);
  for (size_t t = 0; t < T; t++)
    PimManager::check(fhe_h2d(ctx, d.p + t * bytes, words(a[t].GetValues()), bytes, S), "pim hook: upload");
  PimManager::check(to_evaluation ? fhe_ntt_forward(ring->handle(), d.p, 1, S) : fhe_ntt_inverse(ring->handle(), d.p, 1, S),
                    "pim hook: transform");
  PimManager::check(fhe_sync(ctx, S), "pim hook: sync");
  for (size_t t = 0; t < T; t++) {
    auto v = a[t].GetValues();                      /* copy: same length and modulus */
    PimManager::check(fhe_d2h(ctx, const_cast<uint64_t *>(words(v)), d.p + t * bytes, bytes, S), "pim hook: download");
    PimManager::check(fhe_sync(ctx, S), "pim hook: sync");
    a[t].SetValues(std::move(v), new_format);       /* sets the tower's format without transforming on the CPU */
  }
  return true;
}

/* ---- resident form ---------------------------------------------------------------------------------------
 * A DCRTPolyImpl that has been through a hooked operator keeps its towers in HBM ([T][N] words) and only the
 * device copy is updated by further hooked operators; the host towers are refreshed lazily, the first time
 * anything else reads them.  oracle/hook_patch.py arranges that by renaming the member to m_vectors_ and routing
 * every other use of `m_vectors` through DCRTPolyImpl::PimHostVectors() (const: download if the device copy is
 * newer; non-const: download, then drop the device copy because the caller may write).  This is the
 * "operands stay resident between ops; host copies only on sync()" intent of pim/PimData.h:16-34 applied at
 * DCRTPoly level. */
struct Resident {
  std::shared_ptr<PimRing> ring;
  fhe_ctx *ctx = nullptr;
  fhe_dptr d = 0;
  size_t towers = 0, n = 0;
  bool device_newer = false;          /* guarded by mu */
  int stream = 0;                     /* the stream whose work wrote `d` last */
  std::mutex mu;
  std::shared_ptr<void> owner;        /* set: `d` is a view into a block that several device copies share (CRTDecompose) */
  ~Resident() { if (d && !owner) fhe_free(ctx, d); }
};

inline long &resident_uploads() { static long n = 0; return n; }     /* counters for the tests */
inline long &resident_downloads() { static long n = 0; return n; }

/* one pinned staging block per process, so that a whole DCRTPoly crosses PCIe in ONE copy at pinned-memory speed
 * (the towers themselves are separately allocated pageable vectors, dcrtpoly.h:421) */
struct Staging {
  std::mutex mu;
  fhe_ctx *ctx = nullptr;
  void *p = nullptr;
  size_t bytes = 0;
  uint64_t *need(fhe_ctx *c, size_t n) {          /* call with mu held */
    if (n > bytes || c != ctx) {
      if (p) fhe_host_free(ctx, p);
      p = nullptr; bytes = 0; ctx = c;
      PimManager::check(fhe_host_alloc(c, n, &p), "pim hook: pinned staging");
      bytes = n;
    }
    return static_cast<uint64_t *>(p);
  }
};
inline Staging &staging() { thread_local Staging s; return s; }      /* one per host thread: see hook_stream */

/* before `r` is read or written on stream S: wait for the stream that wrote it last, if that is another one */
inline void join_stream(const std::shared_ptr<Resident> &r, int S) {
  if (r && r->stream != S) PimManager::check(fhe_sync(r->ctx, r->stream), "pim hook: stream hand-over");
}

inline void trace_site(int kind);

/* makes `towers` resident (uploading them if they are not yet) and returns the device copy */
template <class Towers>
inline std::shared_ptr<Resident> ensure_device(PimManager *pim, const Towers &towers, std::shared_ptr<Resident> &res) {
  const int S = hook_stream(pim->context());
  auto cur = std::atomic_load(&res);
  if (cur && cur->towers == towers.size() && cur->n == towers[0].GetRingDimension()) { join_stream(cur, S); return cur; }
  /* first use of this operand: several threads may share it (a key polynomial inside an OpenMP loop), one uploads;
   * the lock is striped by the operand's address so that unrelated operands upload concurrently */
  static std::mutex create_mu[64];
  std::lock_guard<std::mutex> create(create_mu[(reinterpret_cast<uintptr_t>(&res) >> 4) & 63]);
  cur = std::atomic_load(&res);
  if (cur && cur->towers == towers.size() && cur->n == towers[0].GetRingDimension()) { join_stream(cur, S); return cur; }
  auto r = std::make_shared<Resident>();
  r->ring = ring_of(pim, towers);
  r->ctx = pim->context();
  r->stream = S;
  r->towers = towers.size();
  r->n = towers[0].GetRingDimension();
  const size_t bytes = r->n * sizeof(uint64_t);
  PimManager::check(fhe_alloc(r->ctx, r->towers * bytes, &r->d), "pim hook: device buffer");
  {
    Staging &st = staging();
    std::lock_guard<std::mutex> lock(st.mu);
    uint64_t *h = st.need(r->ctx, r->towers * bytes);
    for (size_t t = 0; t < r->towers; t++) memcpy(h + t * r->n, words(towers[t].GetValues()), bytes);
    PimManager::check(fhe_h2d(r->ctx, r->d, h, r->towers * bytes, S), "pim hook: upload");
    PimManager::check(fhe_sync(r->ctx, S), "pim hook: sync");   /* the staging block is reused after this */
  }
  resident_uploads()++;
  trace_site(1);
  std::atomic_store(&res, r);
  return r;
}

/* FHE_B200_HOOK_TRACE=1 (debugging aid): which callers force a download.  Counts the return addresses a few frames above
 * sync_host and prints the most frequent at exit as offsets into their module (resolve with addr2line -f -C -e <module>). */
#if defined(__GLIBC__)
#include <dlfcn.h>
#include <execinfo.h>
#endif
inline void trace_site(int kind) {      /* kind 0: download, 1: upload */
#if defined(__GLIBC__)
  static const bool on = [] { const char *e = std::getenv("FHE_B200_HOOK_TRACE"); return e && *e && *e != '0'; }();
  if (!on) return;
  static std::mutex mu;
  static std::map<std::vector<void *>, long> hist[2];
  static const bool reg = [] {
    std::atexit([] {
      for (int k = 0; k < 2; k++) {
        std::vector<std::pair<long, std::vector<void *>>> v;
        for (auto &kv : hist[k]) v.push_back({kv.second, kv.first});
        std::sort(v.begin(), v.end(), [](const auto &a, const auto &b) { return a.first > b.first; });
        for (size_t i = 0; i < v.size() && i < 16; i++) {
          std::fprintf(stderr, "%s site %5ld x:", k ? "upload" : "download", v[i].first);
          for (void *a : v[i].second) {
            Dl_info di;
            if (dladdr(a, &di) && di.dli_fbase) std::fprintf(stderr, " %#lx", (unsigned long)((char *)a - (char *)di.dli_fbase));
          }
          std::fprintf(stderr, "\n");
        }
      }
    });
    return true;
  }();
  (void)reg;
  void *bt[8];
  const int n = backtrace(bt, 8);
  std::vector<void *> key;
  for (int i = 2; i < n && i < 6; i++) key.push_back(bt[i]);
  std::lock_guard<std::mutex> lock(mu);
  hist[kind][key]++;
#endif
}

/* host towers <- device copy if that is newer; every tower is put in `format` (the DCRTPoly's own) */
template <class Towers, class Format>
inline void sync_host(Towers &towers, const std::shared_ptr<Resident> &res, Format format) {
  if (!res) return;
  std::lock_guard<std::mutex> lock(res->mu);
  if (!res->device_newer) return;
  const size_t bytes = res->n * sizeof(uint64_t);
  const int S = hook_stream(res->ctx);
  join_stream(res, S);
  {
    Staging &st = staging();
    std::lock_guard<std::mutex> slock(st.mu);
    uint64_t *h = st.need(res->ctx, res->towers * bytes);
    PimManager::check(fhe_d2h(res->ctx, h, res->d, res->towers * bytes, S), "pim hook: download");
    PimManager::check(fhe_sync(res->ctx, S), "pim hook: sync");
    for (size_t t = 0; t < res->towers; t++) {       /* in place: same length and modulus */
      if (towers[t].IsEmpty()) towers[t].SetValuesToZero();     /* a device result: host storage appears on first read */
      memcpy(const_cast<uint64_t *>(words(towers[t].GetValues())), h + t * res->n, bytes);
      towers[t].OverrideFormat(format);               /* poly.h:179-181: sets the format without transforming */
    }
  }
  resident_downloads()++;
  res->device_newer = false;
  trace_site(0);
}

/* FHE_B200_HOOK_SYNC=1: every hooked operator waits for its kernels (per-operation latency instead of the default
 * enqueue-and-return, where only a host read waits) — for measurements */
inline void finish(fhe_ctx *ctx) {
  static const bool on = [] { const char *e = std::getenv("FHE_B200_HOOK_SYNC"); return e && *e && *e != '0'; }();
  if (on) PimManager::check(fhe_sync(ctx, hook_stream(ctx)), "pim hook: sync");
}

/* a second device copy of the same values (copy construction / assignment of a resident DCRTPoly): HBM to HBM, so a
 * copy of a device result costs no download and a copy of an uploaded operand no second upload */
inline std::shared_ptr<Resident> clone_resident(const std::shared_ptr<Resident> &src) {
  if (!src) return nullptr;
  auto r = std::make_shared<Resident>();
  std::lock_guard<std::mutex> lock(src->mu);
  const int S = hook_stream(src->ctx);
  join_stream(src, S);
  r->stream = S;
  r->ring = src->ring;
  r->ctx = src->ctx;
  r->towers = src->towers;
  r->n = src->n;
  const size_t bytes = r->towers * r->n * sizeof(uint64_t);
  PimManager::check(fhe_alloc(r->ctx, bytes, &r->d), "pim hook: device buffer");
  PimManager::check(fhe_d2d(r->ctx, r->d, src->d, bytes, S), "pim hook: device copy");
  r->device_newer = src->device_newer;
  return r;
}

/* dst = src for an attached DCRTPoly: src is uploaded on its first copy (and stays resident, so later copies and
 * operators find it in HBM), dst becomes a device-only poly — towers without host storage plus an HBM-to-HBM copy.
 * Returns false when there is nothing to upload (empty towers): the caller copies on the host as the reference does. */
template <class Towers>
inline bool copy_resident(PimManager *pim, Towers &dst, std::shared_ptr<Resident> &rdst, const Towers &src,
                          std::shared_ptr<Resident> &rsrc) {
  if (!pim || src.empty()) return false;
  auto from = std::atomic_load(&rsrc);
  if (!from) {
    for (const auto &t : src)
      if (t.IsEmpty()) return false;
    from = ensure_device(pim, src, rsrc);
  }
  auto c = clone_resident(from);
  c->device_newer = true;
  Towers shell;
  shell.reserve(src.size());
  for (const auto &t : src) shell.emplace_back(t.GetParams(), t.GetFormat(), false);
  dst = std::move(shell);
  rdst = c;
  return true;
}

/* Process-wide opt-in: with FHE_B200_ATTACH=1 in the environment every DCRTPoly is born with the backend handle
 * attached (the fork attaches it per object with SetPim, dcrtpoly.h:410-412), so unmodified callers — the reference's
 * own benchmark/src/poly-benchmark-16k.cpp — run their +, *, SwitchFormat on the device. */
inline std::shared_ptr<PimManager> default_handle() {
  static std::shared_ptr<PimManager> h = [] {
    const char *e = std::getenv("FHE_B200_ATTACH");
    if (!e || !*e || *e == '0') return std::shared_ptr<PimManager>();
    if (const char *r = std::getenv("FHE_B200_ATTACH_REPORT"); r && *r && *r != '0')      /* for unmodified programs: proof of use */
      std::atexit([] { std::fprintf(stderr, "fhe_b200: %ld kernel launches\n", fhe_launch_count()); });
    return std::shared_ptr<PimManager>(PimManager::getPim(16), [](PimManager *) {});      /* 16 streams: see hook_stream */
  }();
  return h;
}

/* a (op)= b on the device copies; nothing crosses PCIe when both are already resident */
template <class Towers>
inline bool resident_eq(PimManager *pim, const Towers &a, std::shared_ptr<Resident> &ra, const Towers &b,
                        std::shared_ptr<Resident> &rb, int op) {
  if (!pim || a.empty() || a.size() != b.size()) return false;
  auto da = ensure_device(pim, a, ra);
  auto db = (&a == &b) ? da : ensure_device(pim, b, rb);
  const int S = hook_stream(da->ctx);
  PimManager::check(fhe_dcrt_op(da->ring->handle(), op, da->d, da->d, db->d, 1, S), "pim hook: dcrt op");
  finish(da->ctx);
  std::lock_guard<std::mutex> lock(da->mu);
  da->stream = S;
  da->device_newer = true;
  return true;
}

/* out = a (op) b, out-of-place (Plus / Minus / Times, dcrtpoly.h:164-200, dcrtpoly-impl.h:398-407): `out` holds
 * allocated host towers of the right shape; only its device copy is written and marked newer. */
template <class Towers>
inline bool resident_binary(PimManager *pim, const Towers &a, std::shared_ptr<Resident> &ra, const Towers &b,
                            std::shared_ptr<Resident> &rb, const Towers &out, std::shared_ptr<Resident> &rout, int op) {
  if (!pim || a.empty() || a.size() != b.size() || out.size() != a.size()) return false;
  auto da = ensure_device(pim, a, ra);
  auto db = (&a == &b) ? da : ensure_device(pim, b, rb);
  const int S = hook_stream(da->ctx);
  auto r = std::make_shared<Resident>();
  r->ring = da->ring;
  r->ctx = da->ctx;
  r->stream = S;
  r->towers = da->towers;
  r->n = da->n;
  PimManager::check(fhe_alloc(r->ctx, r->towers * r->n * sizeof(uint64_t), &r->d), "pim hook: device buffer");
  PimManager::check(fhe_dcrt_op(da->ring->handle(), op, r->d, da->d, db->d, 1, S), "pim hook: dcrt op");
  finish(da->ctx);
  r->device_newer = true;
  rout = r;
  return true;
}

struct DeviceBlock {          /* an arena block shared by several device copies (views) */
  fhe_ctx *ctx;
  fhe_dptr p = 0;
  ~DeviceBlock() { if (p) fhe_free(ctx, p); }
};

/* a device copy of the given shape with no contents yet; the caller launches into it on stream S */
inline std::shared_ptr<Resident> fresh_resident(const std::shared_ptr<PimRing> &ring, fhe_ctx *ctx, int S, size_t towers, size_t n) {
  auto r = std::make_shared<Resident>();
  r->ring = ring;
  r->ctx = ctx;
  r->stream = S;
  r->towers = towers;
  r->n = n;
  PimManager::check(fhe_alloc(ctx, towers * n * sizeof(uint64_t), &r->d), "pim hook: device buffer");
  r->device_newer = true;
  return r;
}

/* two such copies as adjacent views of ONE block (one allocation, and a composition that writes both - the two halves of
 * a key switch - can do it in one launch: fhe_keyswitch_hybrid / fhe_keyswitch_fast check for adjacency) */
inline void fresh_pair(const std::shared_ptr<PimRing> &ring, fhe_ctx *ctx, int S, size_t towers, size_t n,
                       std::shared_ptr<Resident> &a, std::shared_ptr<Resident> &b) {
  auto blk = std::make_shared<DeviceBlock>();
  blk->ctx = ctx;
  const size_t bytes = towers * n * sizeof(uint64_t);
  PimManager::check(fhe_alloc(ctx, 2 * bytes, &blk->p), "pim hook: device buffer");
  std::shared_ptr<Resident> *out[2] = {&a, &b};
  for (int i = 0; i < 2; i++) {
    auto r = std::make_shared<Resident>();
    r->ring = ring;
    r->ctx = ctx;
    r->stream = S;
    r->towers = towers;
    r->n = n;
    r->d = blk->p + i * bytes;
    r->owner = blk;
    r->device_newer = true;
    *out[i] = r;
  }
}

/* LeveledSHEBase::EvalMultCore, the 2 x 2 case (src/pke/lib/schemebase/base-leveledshe.cpp:667-672, and the same
 * products in LeveledSHEBFVRNS::EvalMult, bfvrns-leveledshe.cpp:324-336): (a0*b0, a0*b1 + a1*b0, a1*b1) as ONE launch
 * (fhe_ct_tensor_product) into three fresh device copies, instead of two polynomial copies, four Times and a +=. */
template <class Towers>
inline bool resident_tensor_product(PimManager *pim, const Towers &a0, std::shared_ptr<Resident> &ra0, const Towers &a1,
                                    std::shared_ptr<Resident> &ra1, const Towers &b0, std::shared_ptr<Resident> &rb0, const Towers &b1,
                                    std::shared_ptr<Resident> &rb1, std::shared_ptr<Resident> (&out)[3]) {
  if (!pim || a0.empty() || a1.size() != a0.size() || b0.size() != a0.size() || b1.size() != a0.size()) return false;
  auto d0 = ensure_device(pim, a0, ra0), d1 = ensure_device(pim, a1, ra1);
  auto e0 = (&b0 == &a0) ? d0 : ensure_device(pim, b0, rb0), e1 = (&b1 == &a1) ? d1 : ensure_device(pim, b1, rb1);
  if (d1->ring != d0->ring || e0->ring != d0->ring || e1->ring != d0->ring) return false;      /* same moduli, same ring object */
  fhe_ctx *ctx = d0->ctx;
  const int S = hook_stream(ctx);
  std::shared_ptr<Resident> r[3];
  for (auto &x : r) x = fresh_resident(d0->ring, ctx, S, d0->towers, d0->n);
  const fhe_dptr c[3] = {r[0]->d, r[1]->d, r[2]->d}, a[2] = {d0->d, d1->d}, b[2] = {e0->d, e1->d};
  PimManager::check(fhe_ct_tensor_product(d0->ring->handle(), c, a, b, 1, S), "pim hook: tensor product");
  finish(ctx);
  for (int i = 0; i < 3; i++) out[i] = r[i];
  return true;
}

template <class Towers>
inline bool resident_switch_format(PimManager *pim, const Towers &a, std::shared_ptr<Resident> &ra, bool to_evaluation) {
  if (!pim || a.empty()) return false;
  auto da = ensure_device(pim, a, ra);
  const int S = hook_stream(da->ctx);
  PimManager::check(to_evaluation ? fhe_ntt_forward(da->ring->handle(), da->d, 1, S) : fhe_ntt_inverse(da->ring->handle(), da->d, 1, S),
                    "pim hook: transform");
  finish(da->ctx);
  std::lock_guard<std::mutex> lock(da->mu);
  da->stream = S;
  da->device_newer = true;
  return true;
}

/* ---- the neighbouring DCRTPoly methods on resident operands -----------------------
 * Plans (basis tables, rescale constants) are cached per pair of rings, like the schemes' per-parameter-set tables;
 * a plan owns scratch buffers, so a call holds its mutex while it enqueues work (everything runs on stream 0). */
struct PlanLock {
  std::mutex mu;
};
template <class Plan>
struct CachedPlan {
  std::shared_ptr<Plan> plan;
  std::shared_ptr<PlanLock> lock = std::make_shared<PlanLock>();
};

/* a plan owns scratch buffers, so every stream gets its own (fhe_b200.h: "use one per stream") */
inline CachedPlan<PimBasis> basis_for(const std::shared_ptr<PimRing> &from, const std::shared_ptr<PimRing> &to) {
  static std::mutex mu;
  static std::map<std::tuple<PimRing *, PimRing *, int>, CachedPlan<PimBasis>> cache;
  std::lock_guard<std::mutex> lock(mu);
  auto key = std::make_tuple(from.get(), to.get(), hook_stream(from->manager()->context()));
  auto it = cache.find(key);
  if (it != cache.end()) return it->second;
  CachedPlan<PimBasis> c;
  c.plan = std::make_shared<PimBasis>(from, to);        /* tables derived from the moduli = the schemes' values */
  cache.emplace(key, c);
  return c;
}

inline CachedPlan<PimRescale> rescale_for(const std::shared_ptr<PimRing> &ring) {
  static std::mutex mu;
  static std::map<std::pair<PimRing *, int>, CachedPlan<PimRescale>> cache;
  std::lock_guard<std::mutex> lock(mu);
  auto key = std::make_pair(ring.get(), hook_stream(ring->manager()->context()));
  auto it = cache.find(key);
  if (it != cache.end()) return it->second;
  CachedPlan<PimRescale> c;
  c.plan = std::make_shared<PimRescale>(ring);
  cache.emplace(key, c);
  return c;
}

/* a fresh device copy shaped like `towers` (which only have to carry the right moduli, roots and lengths) */
template <class Towers>
inline std::shared_ptr<Resident> new_resident(PimManager *pim, const Towers &towers) {
  auto r = std::make_shared<Resident>();
  r->ring = ring_of(pim, towers);
  r->ctx = pim->context();
  r->stream = hook_stream(r->ctx);
  r->towers = towers.size();
  r->n = towers[0].GetRingDimension();
  PimManager::check(fhe_alloc(r->ctx, r->towers * r->n * sizeof(uint64_t), &r->d), "pim hook: device buffer");
  r->device_newer = true;
  return r;
}

/* DCRTPolyImpl::DropLastElement(s) (dcrtpoly-impl.h:719-741) on a resident polynomial: towers are contiguous in HBM, so the
 * polynomial over the first `keep` towers is a prefix of the same block - a view, no kernel, no copy.  `rest` = the host
 * towers that remain (moduli and roots). */
template <class Towers>
inline std::shared_ptr<Resident> resident_prefix(PimManager *pim, const std::shared_ptr<Resident> &full, const Towers &rest) {
  if (!full || rest.empty() || rest.size() > full->towers) return nullptr;
  auto r = std::make_shared<Resident>();
  std::lock_guard<std::mutex> lock(full->mu);
  r->ring = ring_of(pim, rest);
  r->ctx = full->ctx;
  r->stream = full->stream;
  r->towers = rest.size();
  r->n = full->n;
  r->d = full->d;
  r->owner = full;                  /* keeps the block alive; the view itself frees nothing */
  r->device_newer = full->device_newer;
  return r;
}

/* DCRTPolyImpl::Plus / Minus(const std::vector<Integer>&) (dcrtpoly-impl.h:556-563, 577-584): out = a (+|-) one constant per
 * tower, every coefficient.  (Plus: EVALUATION format only - the caller checks.) */
template <class Towers, class IntVec>
inline bool resident_tower_scalars(PimManager *pim, int op, const Towers &a, std::shared_ptr<Resident> &ra, const Towers &out,
                                   std::shared_ptr<Resident> &rout, const IntVec &c) {
  if (!pim || a.empty() || a.size() > 64 || out.size() != a.size() || c.size() < a.size() || a[0].GetRingDimension() < 2) return false;
  const int S = hook_stream(pim->context());
  auto da = ensure_device(pim, a, ra);
  auto r = new_resident(pim, out);
  uint64_t k[64];
  for (size_t i = 0; i < a.size(); i++) k[i] = c[i].ConvertToInt();
  PimManager::check(fhe_dcrt_op_scalars(da->ring->handle(), op, r->d, da->d, k, 1, S), "pim hook: tower scalars");
  finish(r->ctx);
  rout = r;
  return true;
}

/* DCRTPolyImpl::AutomorphismTransform (dcrtpoly-impl.h:348-369): out = sigma_k(a); format 0 EVALUATION, 1 COEFFICIENT */
template <class Towers>
inline bool resident_automorphism(PimManager *pim, const Towers &a, std::shared_ptr<Resident> &ra, const Towers &out,
                                  std::shared_ptr<Resident> &rout, uint32_t k, int format) {
  if (!pim || a.empty() || out.size() != a.size()) return false;
  const int S = hook_stream(pim->context());
  auto da = ensure_device(pim, a, ra);
  auto r = new_resident(pim, out);
  PimManager::check(fhe_automorphism(da->ring->handle(), r->d, da->d, 1, k, format, S), "pim hook: automorphism");
  finish(da->ctx);
  rout = r;
  return true;
}

/* DCRTPolyImpl::DropLastElementAndScale (dcrtpoly-impl.h:746-769), EVALUATION input: `rest` = the towers that
 * remain (the caller has already shrunk its host tower vector); the device copy is replaced by the rescaled one. */
template <class Towers>
inline bool resident_rescale(PimManager *pim, const std::shared_ptr<Resident> &full, const Towers &rest,
                             std::shared_ptr<Resident> &res) {
  const int S = hook_stream(pim->context());
  join_stream(full, S);
  auto plan = rescale_for(full->ring);
  auto r = new_resident(pim, rest);
  {
    std::lock_guard<std::mutex> lock(plan.lock->mu);
    PimManager::check(fhe_drop_last_element_and_scale(plan.plan->handle(), r->d, full->d, 1, 0, S), "pim hook: rescale");
  }
  res = r;      /* the old device copy is released when `full` goes out of scope; stream order keeps it valid until then */
  return true;
}

/* the scheme's QHatInvModq (the table the caller passed in) against the one the cached plan derived from the moduli:
 * a caller with other constants (another level's tables, a partial basis) stays on the CPU path */
template <class IntVec>
inline bool same_hat_inv(const CachedPlan<PimBasis> &plan, const IntVec &QHatInvModq, size_t sq, size_t sp) {
  if (QHatInvModq.size() < sq) return false;
  std::vector<uint64_t> hi(sq), hm(sq * sp);
  if (fhe_basis_tables(plan.plan->handle(), hi.data(), hm.data()) != 0) return false;
  for (size_t i = 0; i < sq; i++)
    if (hi[i] != QHatInvModq[i].ConvertToInt()) return false;
  return true;
}

/* DCRTPolyImpl::ApproxSwitchCRTBasis (dcrtpoly-impl.h:1034-1082): a over its first `sq` towers -> out over its own moduli */
template <class Towers, class IntVec>
inline bool resident_approx_switch(PimManager *pim, const Towers &a, std::shared_ptr<Resident> &ra, size_t sq, const Towers &out,
                                   std::shared_ptr<Resident> &rout, const IntVec &QHatInvModq) {
  if (!pim || a.empty() || out.empty() || sq != a.size()) return false;      /* partial source bases stay on the CPU path */
  const int S = hook_stream(pim->context());
  auto da = ensure_device(pim, a, ra);
  auto plan = basis_for(da->ring, ring_of(pim, out));
  if (!same_hat_inv(plan, QHatInvModq, sq, out.size())) return false;
  auto r = new_resident(pim, out);
  {
    std::lock_guard<std::mutex> lock(plan.lock->mu);
    PimManager::check(fhe_approx_switch_crt_basis(plan.plan->handle(), r->d, da->d, 1, S), "pim hook: basis switch");
  }
  rout = r;
  return true;
}

/* DCRTPolyImpl::ApproxModDown (dcrtpoly-impl.h:1136-1175), EVALUATION input over QP: out over the first `sq` towers */
template <class Towers, class IntVec>
inline bool resident_mod_down(PimManager *pim, const Towers &a, std::shared_ptr<Resident> &ra, size_t sq, uint64_t t,
                              const Towers &out, std::shared_ptr<Resident> &rout, const IntVec &PHatInvModp) {
  if (!pim || a.empty() || sq == 0 || sq >= a.size() || out.size() != sq) return false;
  const int S = hook_stream(pim->context());
  auto da = ensure_device(pim, a, ra);
  static std::mutex mu;
  static std::map<std::pair<PimRing *, size_t>, std::pair<std::shared_ptr<PimRing>, std::shared_ptr<PimRing>>> views;
  std::shared_ptr<PimRing> rq, rp;
  {
    std::lock_guard<std::mutex> lock(mu);
    auto key = std::make_pair(da->ring.get(), sq);
    auto it = views.find(key);
    if (it == views.end())
      it = views.emplace(key, std::make_pair(PimRing::View(da->ring, 0, (uint32_t)sq),
                                             PimRing::View(da->ring, (uint32_t)sq, (uint32_t)(a.size() - sq)))).first;
    rq = it->second.first;
    rp = it->second.second;
  }
  auto plan = basis_for(rp, rq);
  if (!same_hat_inv(plan, PHatInvModp, a.size() - sq, sq)) return false;
  auto r = std::make_shared<Resident>();
  r->ring = rq;
  r->ctx = pim->context();
  r->stream = S;
  r->towers = sq;
  r->n = da->n;
  PimManager::check(fhe_alloc(r->ctx, r->towers * r->n * sizeof(uint64_t), &r->d), "pim hook: device buffer");
  r->device_newer = true;
  {
    std::lock_guard<std::mutex> lock(plan.lock->mu);
    PimManager::check(fhe_approx_mod_down(plan.plan->handle(), r->d, da->d, 1, t, S), "pim hook: mod down");
  }
  rout = r;
  return true;
}

/* DCRTPolyImpl::ApproxModUp (dcrtpoly-impl.h:1085-1133): a over Q (either format) -> out over QP, EVALUATION.
 * `out` = host shells of the QP towers (moduli and roots only). */
template <class Towers, class IntVec>
inline bool resident_mod_up(PimManager *pim, const Towers &a, std::shared_ptr<Resident> &ra, const Towers &out,
                            std::shared_ptr<Resident> &rout, int format, const IntVec &QHatInvModq) {
  if (!pim || a.empty() || out.size() <= a.size()) return false;
  const int S = hook_stream(pim->context());
  auto da = ensure_device(pim, a, ra);
  auto r = new_resident(pim, out);
  const size_t sq = a.size(), sp = out.size() - sq;
  static std::mutex mu;
  static std::map<std::pair<PimRing *, size_t>, std::shared_ptr<PimRing>> views;      /* the P towers of a QP ring */
  std::shared_ptr<PimRing> rp;
  {
    std::lock_guard<std::mutex> lock(mu);
    auto key = std::make_pair(r->ring.get(), sq);
    auto it = views.find(key);
    if (it == views.end()) it = views.emplace(key, PimRing::View(r->ring, (uint32_t)sq, (uint32_t)sp)).first;
    rp = it->second;
  }
  auto plan = basis_for(da->ring, rp);
  if (!same_hat_inv(plan, QHatInvModq, sq, sp)) return false;
  {
    std::lock_guard<std::mutex> lock(plan.lock->mu);
    PimManager::check(fhe_approx_mod_up(plan.plan->handle(), r->d, da->d, 1, format, S), "pim hook: mod up");
  }
  finish(r->ctx);
  rout = r;
  return true;
}

/* DCRTPolyImpl::ModReduce (dcrtpoly-impl.h:792-812, BGV modulus switching): like resident_rescale with the plaintext
 * modulus t; the format is preserved */
template <class Towers>
inline bool resident_mod_reduce(PimManager *pim, const std::shared_ptr<Resident> &full, const Towers &rest,
                                std::shared_ptr<Resident> &res, uint64_t t, int format) {
  const int S = hook_stream(pim->context());
  join_stream(full, S);
  auto plan = rescale_for(full->ring);
  auto r = new_resident(pim, rest);
  {
    std::lock_guard<std::mutex> lock(plan.lock->mu);
    PimManager::check(fhe_mod_reduce(plan.plan->handle(), r->d, full->d, 1, t, format, S), "pim hook: mod reduce");
  }
  finish(r->ctx);
  res = r;
  return true;
}

/* DCRTPolyImpl::SwitchCRTBasis (dcrtpoly-impl.h:1178-1230, BFV, exact): a over Q -> out over P, same format */
template <class Towers, class IntVec>
inline bool resident_switch_exact(PimManager *pim, const Towers &a, std::shared_ptr<Resident> &ra, const Towers &out,
                                  std::shared_ptr<Resident> &rout, const IntVec &QHatInvModq) {
  if (!pim || a.empty() || out.empty()) return false;
  const int S = hook_stream(pim->context());
  auto da = ensure_device(pim, a, ra);
  auto r = new_resident(pim, out);
  auto plan = basis_for(da->ring, r->ring);
  if (!same_hat_inv(plan, QHatInvModq, a.size(), out.size())) return false;
  {
    std::lock_guard<std::mutex> lock(plan.lock->mu);
    PimManager::check(fhe_switch_crt_basis(plan.plan->handle(), r->d, da->d, 1, S), "pim hook: exact basis switch");
  }
  finish(r->ctx);
  rout = r;
  return true;
}

/* DCRTPolyImpl::ExpandCRTBasis (dcrtpoly-impl.h:1311-1358, BFV): a over Q (either format) -> out over QP in
 * result_format; `out` = host shells of the QP towers */
template <class Towers, class IntVec>
inline bool resident_expand(PimManager *pim, const Towers &a, std::shared_ptr<Resident> &ra, const Towers &out,
                            std::shared_ptr<Resident> &rout, int format, int result_format, const IntVec &QHatInvModq,
                            bool reverse = false) {      /* reverse: ExpandCRTBasisReverseOrder (:1361-1410), P towers first */
  if (!pim || a.empty() || out.size() <= a.size()) return false;
  const int S = hook_stream(pim->context());
  auto da = ensure_device(pim, a, ra);
  auto r = new_resident(pim, out);
  const size_t sq = a.size(), sp = out.size() - sq;
  static std::mutex mu;
  static std::map<std::tuple<PimRing *, size_t, bool>, std::shared_ptr<PimRing>> views;
  std::shared_ptr<PimRing> rp;
  {
    std::lock_guard<std::mutex> lock(mu);
    auto key = std::make_tuple(r->ring.get(), sq, reverse);
    auto it = views.find(key);
    if (it == views.end()) it = views.emplace(key, PimRing::View(r->ring, reverse ? 0u : (uint32_t)sq, (uint32_t)sp)).first;
    rp = it->second;
  }
  auto plan = basis_for(da->ring, rp);
  if (!same_hat_inv(plan, QHatInvModq, sq, sp)) return false;
  {
    std::lock_guard<std::mutex> lock(plan.lock->mu);
    if (reverse)
      PimManager::check(fhe_expand_crt_basis_reverse_order(plan.plan->handle(), r->d, da->d, 1, format, result_format, S),
                        "pim hook: expand basis (reverse order)");
    else
      PimManager::check(fhe_expand_crt_basis(plan.plan->handle(), r->d, da->d, 1, format, result_format, S), "pim hook: expand basis");
  }
  finish(r->ctx);
  rout = r;
  return true;
}

/* DCRTPolyImpl::FastExpandCRTBasisPloverQ (dcrtpoly-impl.h:1413-1468, BFV HPSPOVERQ / HPSPOVERQLEVELED): a over Q -> out
 * over [Ql | Pl] (`out` = host shells of those towers), format unchanged.  The first basis carries the scheme's own
 * mPlQHatInvModq / qInvModp (cached by the table's address and checked against a copy of its values); the exact switch
 * back uses the tables derived from the moduli, checked against the caller's PlHatInvModp. */
template <class Towers, class IntVec, class Rows>
inline bool resident_fast_expand_pl_over_q(PimManager *pim, const Towers &a, std::shared_ptr<Resident> &ra, const Towers &out,
                                           std::shared_ptr<Resident> &rout, const IntVec &mPlQHatInvModq, const Rows &qInvModp,
                                           const IntVec &PlHatInvModp) {
  const size_t sq = a.size(), spl = out.size() / 2, sql = out.size() - spl;
  if (!pim || a.empty() || spl == 0 || sql != spl || mPlQHatInvModq.size() < sq || qInvModp.size() < sq) return false;
  const int S = hook_stream(pim->context());
  auto da = ensure_device(pim, a, ra);
  auto r = new_resident(pim, out);
  static std::mutex mu;
  static std::map<PimRing *, std::pair<std::shared_ptr<PimRing>, std::shared_ptr<PimRing>>> views;
  struct Entry {
    CachedPlan<PimBasis> plan;
    std::vector<uint64_t> neg;
  };
  static std::map<std::tuple<PimRing *, PimRing *, const void *, int>, Entry> cache;
  std::shared_ptr<PimRing> rql, rpl;
  Entry e;
  {
    std::lock_guard<std::mutex> lock(mu);
    auto it = views.find(r->ring.get());
    if (it == views.end())
      it = views.emplace(r->ring.get(), std::make_pair(PimRing::View(r->ring, 0, (uint32_t)sql),
                                                       PimRing::View(r->ring, (uint32_t)sql, (uint32_t)spl))).first;
    rql = it->second.first;
    rpl = it->second.second;
    std::vector<uint64_t> neg(sq);
    for (size_t i = 0; i < sq; i++) neg[i] = mPlQHatInvModq[i].ConvertToInt();
    auto key = std::make_tuple(da->ring.get(), rpl.get(), (const void *)&mPlQHatInvModq[0], S);
    auto ce = cache.find(key);
    if (ce == cache.end() || ce->second.neg != neg) {
      std::vector<uint64_t> flat;
      flat.reserve(sq * spl);
      for (size_t i = 0; i < sq; i++) {
        if (qInvModp[i].size() < spl) return false;
        for (size_t j = 0; j < spl; j++) flat.push_back(qInvModp[i][j].ConvertToInt());
      }
      Entry fresh;
      fresh.plan.plan = std::make_shared<PimBasis>(da->ring, rpl, neg, flat);
      fresh.neg = neg;
      cache[key] = fresh;
      ce = cache.find(key);
    }
    e = ce->second;
  }
  auto back = basis_for(rpl, rql);
  if (!same_hat_inv(back, PlHatInvModp, spl, sql)) return false;
  {
    std::lock_guard<std::mutex> l1(e.plan.lock->mu);
    std::lock_guard<std::mutex> l2(back.lock->mu);
    PimManager::check(fhe_fast_expand_crt_basis_pl_over_q(e.plan.plan->handle(), back.plan->handle(), r->d, da->d, 1, S),
                      "pim hook: fast expand P over Q");
  }
  finish(r->ctx);
  rout = r;
  return true;
}

/* DCRTPolyImpl::ExpandCRTBasisQlHat (dcrtpoly-impl.h:1514-1534; mode 0: a over the first towers of `out`'s basis) and
 * DCRTPolyImpl::ScaleAndRoundPOverQ (dcrtpoly-impl.h:2052-2066; mode 1: a over `out`'s basis plus one tower).  The plan is
 * built from the caller's constants, cached by their address and checked against a copy of the values. */
template <class Towers, class IntVec>
inline bool resident_tower_scale(PimManager *pim, int mode, const Towers &a, std::shared_ptr<Resident> &ra, const Towers &out,
                                 std::shared_ptr<Resident> &rout, const IntVec &c) {
  /* mode 2: Times(const std::vector<NativeInteger>&) (dcrtpoly-impl.h:617-627) = mode 0 over all towers with the
   * constants reduced first, as NativeVectorT::ModMulEq does */
  const bool times = mode == 2;
  if (times) mode = 0;
  const size_t T = out.size(), sl = mode == 0 ? a.size() : T;
  if (!pim || a.empty() || T == 0 || c.size() < sl) return false;
  if (mode == 0 ? a.size() > T : a.size() != T + 1) return false;
  if (times && a.size() != T) return false;
  const int S = hook_stream(pim->context());
  auto da = ensure_device(pim, a, ra);
  auto r = new_resident(pim, out);
  struct Entry {
    CachedPlan<PimTowerScale> plan;
    std::vector<uint64_t> c;
  };
  static std::mutex mu;
  static std::map<std::tuple<PimRing *, const void *, size_t>, Entry> cache;
  Entry e;
  {
    std::lock_guard<std::mutex> lock(mu);
    std::vector<uint64_t> k(T, 0);
    for (size_t i = 0; i < sl; i++) {
      k[i] = c[i].ConvertToInt();
      if (times) k[i] %= out[i].GetModulus().ConvertToInt();
    }
    auto key = std::make_tuple(r->ring.get(), (const void *)&c[0], sl);
    auto it = cache.find(key);
    if (it == cache.end() || it->second.c != k) {
      Entry fresh;
      fresh.plan.plan = std::make_shared<PimTowerScale>(r->ring, k);
      fresh.c = k;
      cache[key] = fresh;
      it = cache.find(key);
    }
    e = it->second;
  }
  {
    std::lock_guard<std::mutex> lock(e.plan.lock->mu);
    if (mode == 0)
      PimManager::check(fhe_expand_crt_basis_ql_hat(e.plan.plan->handle(), (uint32_t)sl, r->d, da->d, 1, S),
                        "pim hook: expand basis (QlHat)");
    else
      PimManager::check(fhe_scale_and_round_p_over_q(e.plan.plan->handle(), r->d, da->d, 1, S), "pim hook: scale and round P over Q");
  }
  rout = r;
  return true;
}

/* DCRTPolyImpl::ScaleAndRound(paramsOutput, ...) (dcrtpoly-impl.h:1876-1960, BFV): a over QP -> out over the output
 * basis (its first towers, or its last when out_last).  The plan is built from the scheme's own tables and cached by
 * their address; a hit is checked against a copy of the fractions, so a table reallocated at the same address is caught. */
template <class Towers, class Rows>
inline bool resident_scale_round(PimManager *pim, const Towers &a, std::shared_ptr<Resident> &ra, const Towers &out,
                                 std::shared_ptr<Resident> &rout, bool out_last, const std::vector<double> &frac, const Rows &modo) {
  const size_t so = out.size();
  if (!pim || a.empty() || so == 0 || so >= a.size() || frac.size() != a.size() - so || modo.size() != so) return false;
  const int S = hook_stream(pim->context());
  auto da = ensure_device(pim, a, ra);
  struct Entry {
    std::shared_ptr<PimScaleRound> plan;
    std::shared_ptr<PlanLock> lock = std::make_shared<PlanLock>();
    std::vector<double> frac;
  };
  static std::mutex mu;
  static std::map<std::tuple<PimRing *, size_t, bool, const double *, int>, Entry> cache;
  Entry e;
  {
    std::lock_guard<std::mutex> lock(mu);
    auto key = std::make_tuple(da->ring.get(), so, out_last, frac.data(), S);
    auto it = cache.find(key);
    if (it == cache.end() || it->second.frac != frac) {
      std::vector<uint64_t> flat;
      flat.reserve(so * (frac.size() + 1));
      for (size_t j = 0; j < so; j++) {
        if (modo[j].size() < frac.size() + 1) return false;
        for (size_t i = 0; i <= frac.size(); i++) flat.push_back(modo[j][i].ConvertToInt());
      }
      Entry fresh;
      fresh.plan = std::make_shared<PimScaleRound>(da->ring, (uint32_t)so, out_last, frac, flat);
      fresh.frac = frac;
      cache[key] = fresh;
      it = cache.find(key);
    }
    e = it->second;
  }
  auto r = std::make_shared<Resident>();
  r->ring = e.plan->Out();
  r->ctx = pim->context();
  r->stream = S;
  r->towers = so;
  r->n = da->n;
  PimManager::check(fhe_alloc(r->ctx, r->towers * r->n * sizeof(uint64_t), &r->d), "pim hook: device buffer");
  r->device_newer = true;
  {
    std::lock_guard<std::mutex> lock(e.lock->mu);
    PimManager::check(fhe_scale_and_round(e.plan->handle(), r->d, da->d, 1, S), "pim hook: scale and round");
  }
  finish(r->ctx);
  rout = r;
  return true;
}

/* ---- BFV, BEHZ (dcrtpoly-impl.h:2069-2420) ------------------------------------------------------------------
 * A plan over the ring Q u Bsk derives every table from the moduli and t.  FastBaseConvqToBskMontgomery and
 * FastBaseConvSK do not receive t (their tables do not depend on it): they take any plan already cached for the ring,
 * or build one with t = 2.  Each hook compares one of the caller's tables with the plan's own and leaves the call to
 * the CPU statements when they differ, or when the device plan refuses the shape. */
inline CachedPlan<PimBehz> behz_for(const std::shared_ptr<PimRing> &qbsk, uint32_t nq, uint64_t t) {
  static std::mutex mu;
  static std::map<std::tuple<PimRing *, uint32_t, int>, std::vector<CachedPlan<PimBehz>>> cache;
  std::lock_guard<std::mutex> lock(mu);
  auto &plans = cache[std::make_tuple(qbsk.get(), nq, hook_stream(qbsk->manager()->context()))];
  for (auto &p : plans)
    if (t == 0 || p.plan->PlaintextModulus() == t) return p;
  CachedPlan<PimBehz> c;
  try {
    c.plan = std::make_shared<PimBehz>(qbsk, nq, t ? t : 2);
  } catch (const std::exception &) {
    return c;      /* no plan: the CPU path runs */
  }
  plans.push_back(c);
  return c;
}

template <class IntVec>
inline bool same_behz_table(const CachedPlan<PimBehz> &plan, int which, const IntVec &theirs) {
  if (!plan.plan) return false;
  const std::vector<uint64_t> mine = plan.plan->Table(which);
  if (theirs.size() != mine.size()) return false;
  for (size_t i = 0; i < mine.size(); i++)
    if (mine[i] != theirs[i].ConvertToInt()) return false;
  return true;
}

/* DCRTPolyImpl::FastBaseConvqToBskMontgomery (dcrtpoly-impl.h:2069-2209): a over Q (either format) -> out over QBsk,
 * EVALUATION; `out` = host shells of the QBsk towers */
template <class Towers, class IntVec>
inline bool resident_behz_q_to_bsk(PimManager *pim, const Towers &a, std::shared_ptr<Resident> &ra, const Towers &out,
                                   std::shared_ptr<Resident> &rout, int format, const IntVec &mtildeQHatInvModq) {
  if (!pim || a.empty() || out.size() < a.size() + 2 || a[0].GetRingDimension() < 2) return false;
  const int S = hook_stream(pim->context());
  auto plan = behz_for(ring_of(pim, out), (uint32_t)a.size(), 0);
  if (!same_behz_table(plan, 0, mtildeQHatInvModq)) return false;
  auto da = ensure_device(pim, a, ra);
  auto r = new_resident(pim, out);
  {
    std::lock_guard<std::mutex> lock(plan.lock->mu);
    PimManager::check(fhe_fast_base_conv_q_to_bsk_montgomery(plan.plan->handle(), r->d, da->d, 1, format, S),
                      "pim hook: FastBaseConvqToBskMontgomery");
  }
  finish(r->ctx);
  rout = r;
  return true;
}

/* DCRTPolyImpl::FastRNSFloorq (dcrtpoly-impl.h:2212-2271), in place on a over QBsk (nq = moduliQ.size()) */
template <class Towers, class IntVec>
inline bool resident_behz_floor(PimManager *pim, const Towers &a, std::shared_ptr<Resident> &ra, size_t nq, uint64_t t,
                                const IntVec &tQHatInvModq) {
  if (!pim || a.empty() || nq == 0 || a.size() < nq + 2 || a[0].GetRingDimension() < 2) return false;
  const int S = hook_stream(pim->context());
  auto plan = behz_for(ring_of(pim, a), (uint32_t)nq, t);
  if (!same_behz_table(plan, 6, tQHatInvModq)) return false;
  auto da = ensure_device(pim, a, ra);
  {
    std::lock_guard<std::mutex> lock(plan.lock->mu);
    PimManager::check(fhe_fast_rns_floor_q(plan.plan->handle(), da->d, 1, 1, S), "pim hook: FastRNSFloorq");
  }
  finish(da->ctx);
  std::lock_guard<std::mutex> lock(da->mu);
  da->stream = S;
  da->device_newer = true;
  return true;
}

/* DCRTPolyImpl::FastBaseConvSK (dcrtpoly-impl.h:2309-2420): a over QBsk -> a new device copy over Q (its first nq
 * towers); the caller then shrinks its host tower vector as the reference does (:2402-2408) */
template <class Towers, class IntVec>
inline bool resident_behz_conv_sk(PimManager *pim, const Towers &a, std::shared_ptr<Resident> &ra, size_t nq,
                                  std::shared_ptr<Resident> &rout, const IntVec &BHatInvModb) {
  if (!pim || a.empty() || nq == 0 || a.size() < nq + 2 || a[0].GetRingDimension() < 2) return false;
  const int S = hook_stream(pim->context());
  auto full = ring_of(pim, a);
  auto plan = behz_for(full, (uint32_t)nq, 0);
  if (!same_behz_table(plan, 9, BHatInvModb)) return false;
  auto da = ensure_device(pim, a, ra);
  static std::mutex mu;
  static std::map<std::pair<PimRing *, size_t>, std::shared_ptr<PimRing>> views;      /* the Q towers of a QBsk ring */
  std::shared_ptr<PimRing> rq;
  {
    std::lock_guard<std::mutex> lock(mu);
    auto key = std::make_pair(full.get(), nq);
    auto it = views.find(key);
    if (it == views.end()) it = views.emplace(key, PimRing::View(full, 0, (uint32_t)nq)).first;
    rq = it->second;
  }
  auto r = std::make_shared<Resident>();
  r->ring = rq;
  r->ctx = pim->context();
  r->stream = S;
  r->towers = nq;
  r->n = da->n;
  PimManager::check(fhe_alloc(r->ctx, r->towers * r->n * sizeof(uint64_t), &r->d), "pim hook: device buffer");
  r->device_newer = true;
  {
    std::lock_guard<std::mutex> lock(plan.lock->mu);
    PimManager::check(fhe_fast_base_conv_sk(plan.plan->handle(), r->d, da->d, 1, S), "pim hook: FastBaseConvSK");
  }
  finish(r->ctx);
  rout = r;
  return true;
}

/* ---- ciphertext level: KeySwitchHYBRID::KeySwitchCore (src/pke/lib/keyswitch/keyswitch-hybrid.cpp:324-480) ------------
 * The reference builds the digits tower by tower on the host (SetElementAtIndex), runs the inner product with the key on
 * NativePolys and only the basis conversions on DCRTPolys, so with the per-polynomial hooks alone a key switch bounces
 * between host and device.  Here the whole of EvalKeySwitchPrecomputeCore + EvalFastKeySwitchCoreExt + the two ApproxModDown
 * is ONE device composition (fhe_keyswitch_hybrid) on the resident ciphertext element and a device copy of the key:
 * the key's towers that the level uses (the first sizeQl and the sizeP towers of P, :462-478) are gathered once per
 * (key, level) into [digits][Ql u P][N] blocks, HBM to HBM, and kept. */
inline CachedPlan<PimKeySwitch> keyswitch_for(const std::shared_ptr<PimRing> &qlp, uint32_t size_ql, uint32_t alpha) {
  static std::mutex mu;
  static std::map<std::tuple<PimRing *, uint32_t, uint32_t, int>, CachedPlan<PimKeySwitch>> cache;
  std::lock_guard<std::mutex> lock(mu);
  auto key = std::make_tuple(qlp.get(), size_ql, alpha, hook_stream(qlp->manager()->context()));
  auto it = cache.find(key);
  if (it != cache.end()) return it->second;
  CachedPlan<PimKeySwitch> c;
  c.plan = std::make_shared<PimKeySwitch>(qlp, size_ql, alpha);
  cache.emplace(key, c);
  return c;
}

template <class Towers>
struct KeyRef {
  const Towers &towers;
  std::shared_ptr<Resident> &res;
};

struct KeyBlock {          /* evk_b, evk_a as fhe_keyswitch_inner reads them: [digits][sizeQl + sizeP][N] each */
  fhe_ctx *ctx = nullptr;
  fhe_dptr b = 0, a = 0, from = 0;
  size_t bytes = 0;
  ~KeyBlock() {
    if (b) fhe_free(ctx, b);
    if (a) fhe_free(ctx, a);
  }
};

/* the ring Ql u P of a key switch at the level of `c` (its sizeQl towers), P = the key's last sizeP towers; false when the
 * key's chain of moduli is not c's */
template <class Towers>
inline bool keyswitch_ring(PimManager *pim, const Towers &c, size_t sizeQl, const Towers &key0, size_t sizeP,
                           std::shared_ptr<PimRing> &ring) {
  if (key0.size() < sizeQl + sizeP) return false;
  const size_t sizeQ = key0.size() - sizeP;
  std::vector<uint64_t> q, psi;
  for (size_t i = 0; i < sizeQl; i++) {
    if (key0[i].GetModulus() != c[i].GetModulus()) return false;
    q.push_back(c[i].GetModulus().ConvertToInt());
    psi.push_back(c[i].GetRootOfUnity().ConvertToInt());
  }
  for (size_t i = sizeQ; i < sizeQ + sizeP; i++) {
    q.push_back(key0[i].GetModulus().ConvertToInt());
    psi.push_back(key0[i].GetRootOfUnity().ConvertToInt());
  }
  ring = ring_for(pim, (uint32_t)c[0].GetRingDimension(), q, psi);
  return true;
}

/* the key's device block for a level: built on the first use (HBM to HBM gathers on stream S), then shared */
template <class Towers, class KeyAt>
inline std::shared_ptr<KeyBlock> key_block_for(PimManager *pim, KeyAt key_at, size_t np, size_t sizeQl, size_t sizeP, size_t n,
                                               const void *key_id, int S) {
  static std::mutex mu;
  static std::map<std::tuple<const void *, size_t, fhe_ctx *>, std::shared_ptr<KeyBlock>> keys;
  static size_t held = 0;
  fhe_ctx *ctx = pim->context();
  const size_t T = sizeQl + sizeP;
  std::lock_guard<std::mutex> lock(mu);
  KeyRef<Towers> k0 = key_at(0, 0);
  const size_t sizeQ = k0.towers.size() - sizeP;
  auto d0 = ensure_device(pim, k0.towers, k0.res);
  auto id = std::make_tuple(key_id, sizeQl, ctx);
  auto it = keys.find(id);
  if (it != keys.end() && it->second->from == d0->d) return it->second;
  const size_t tower = n * sizeof(uint64_t), bytes = np * T * tower;
  if (held + 2 * bytes > ((size_t)16 << 30)) { keys.clear(); held = 0; }      /* bounded: blocks in use live on through their shared_ptr */
  auto kb = std::make_shared<KeyBlock>();
  kb->ctx = ctx;
  kb->from = d0->d;
  kb->bytes = 2 * bytes;
  PimManager::check(fhe_alloc(ctx, bytes, &kb->b), "pim hook: key block");
  PimManager::check(fhe_alloc(ctx, bytes, &kb->a), "pim hook: key block");
  for (int which = 0; which < 2; which++)
    for (size_t j = 0; j < np; j++) {
      KeyRef<Towers> kr = key_at(which, j);
      if (kr.towers.size() != sizeQ + sizeP) return nullptr;
      auto dk = ensure_device(pim, kr.towers, kr.res);
      const fhe_dptr dst = (which ? kb->a : kb->b) + j * T * tower;
      PimManager::check(fhe_d2d(ctx, dst, dk->d, sizeQl * tower, S), "pim hook: key gather");
      PimManager::check(fhe_d2d(ctx, dst + sizeQl * tower, dk->d + sizeQ * tower, sizeP * tower, S), "pim hook: key gather");
    }
  PimManager::check(fhe_sync(ctx, S), "pim hook: key gather");      /* other threads use the block on their own streams */
  keys[id] = kb;
  held += 2 * bytes;
  return kb;
}


/* acc0 / acc1 given: the result is ADDED to those polynomials (cv[0] += (*ab)[0]; cv[1] += (*ab)[1] of the schemes'
 * relinearisation, base-leveledshe.cpp:210-213) by the composition's last launch; r0 / r1 are then not written */
template <class Towers, class KeyAt>
inline bool resident_keyswitch_hybrid(PimManager *pim, const Towers &c, std::shared_ptr<Resident> &rc, size_t nkeys, KeyAt key_at,
                                      size_t sizeP, size_t alpha, uint64_t t, const void *key_id, std::shared_ptr<Resident> &r0,
                                      std::shared_ptr<Resident> &r1, const Towers *acc0 = nullptr,
                                      std::shared_ptr<Resident> *racc0 = nullptr, const Towers *acc1 = nullptr,
                                      std::shared_ptr<Resident> *racc1 = nullptr) {
  if (!pim || c.empty() || sizeP == 0 || alpha == 0 || nkeys == 0) return false;
  const size_t sizeQl = c.size(), np = (sizeQl + alpha - 1) / alpha, n = c[0].GetRingDimension();
  if (np > nkeys || n < 2) return false;
  std::shared_ptr<PimRing> ring;
  if (!keyswitch_ring(pim, c, sizeQl, key_at(0, 0).towers, sizeP, ring)) return false;
  fhe_ctx *ctx = pim->context();
  const int S = hook_stream(ctx);
  auto plan = keyswitch_for(ring, (uint32_t)sizeQl, (uint32_t)alpha);
  auto kb = key_block_for<Towers>(pim, key_at, np, sizeQl, sizeP, n, key_id, S);
  if (!kb) return false;
  auto da = ensure_device(pim, c, rc);
  if (acc0) {
    if (!acc1 || acc0->size() != sizeQl || acc1->size() != sizeQl || acc0 == &c || acc1 == &c || acc0 == acc1) return false;
    auto d0 = ensure_device(pim, *acc0, *racc0), d1 = ensure_device(pim, *acc1, *racc1);
    if (d0->ring != da->ring || d1->ring != da->ring) return false;
    {
      std::lock_guard<std::mutex> lock(plan.lock->mu);
      PimManager::check(fhe_keyswitch_hybrid_add(plan.plan->handle(), d0->d, d1->d, da->d, kb->b, kb->a, 1, t, S),
                        "pim hook: hybrid key switch, accumulated");
    }
    finish(ctx);
    for (auto &d : {d0, d1}) {
      std::lock_guard<std::mutex> lock(d->mu);
      d->stream = S;
      d->device_newer = true;
    }
    return true;
  }
  std::shared_ptr<Resident> o0, o1;
  fresh_pair(da->ring, ctx, S, sizeQl, n, o0, o1);
  {
    std::lock_guard<std::mutex> lock(plan.lock->mu);
    PimManager::check(fhe_keyswitch_hybrid(plan.plan->handle(), o0->d, o1->d, da->d, kb->b, kb->a, 1, t, S), "pim hook: hybrid key switch");
  }
  finish(ctx);
  r0 = o0;
  r1 = o1;
  return true;
}

/* KeySwitchHYBRID::EvalKeySwitchPrecomputeCore (keyswitch-hybrid.cpp:330-420, the hoisted half of a rotation): c over Ql ->
 * one device copy per digit over Ql u P, views into one [digits][Ql u P][N] block.  p_q / p_psi = the moduli and roots of P. */
template <class Towers>
inline bool resident_keyswitch_digits(PimManager *pim, const Towers &c, std::shared_ptr<Resident> &rc, const std::vector<uint64_t> &p_q,
                                      const std::vector<uint64_t> &p_psi, size_t alpha, size_t max_digits,
                                      std::vector<std::shared_ptr<Resident>> &digits) {
  if (!pim || c.empty() || p_q.empty() || p_q.size() != p_psi.size() || alpha == 0) return false;
  const size_t sizeQl = c.size(), sizeP = p_q.size(), T = sizeQl + sizeP, n = c[0].GetRingDimension();
  const size_t np = (sizeQl + alpha - 1) / alpha;
  if (n < 2 || np > max_digits) return false;
  std::vector<uint64_t> q, psi;
  for (size_t i = 0; i < sizeQl; i++) { q.push_back(c[i].GetModulus().ConvertToInt()); psi.push_back(c[i].GetRootOfUnity().ConvertToInt()); }
  q.insert(q.end(), p_q.begin(), p_q.end());
  psi.insert(psi.end(), p_psi.begin(), p_psi.end());
  auto ring = ring_for(pim, (uint32_t)n, q, psi);
  fhe_ctx *ctx = pim->context();
  const int S = hook_stream(ctx);
  auto plan = keyswitch_for(ring, (uint32_t)sizeQl, (uint32_t)alpha);
  auto da = ensure_device(pim, c, rc);
  const size_t poly = T * n * sizeof(uint64_t);
  auto blk = std::make_shared<DeviceBlock>();
  blk->ctx = ctx;
  PimManager::check(fhe_alloc(ctx, np * poly, &blk->p), "pim hook: device buffer");
  {
    std::lock_guard<std::mutex> lock(plan.lock->mu);
    PimManager::check(fhe_keyswitch_precompute(plan.plan->handle(), blk->p, da->d, 1, S), "pim hook: key-switch digits");
  }
  finish(ctx);
  digits.clear();
  for (size_t j = 0; j < np; j++) {
    auto r = std::make_shared<Resident>();
    r->ring = ring;
    r->ctx = ctx;
    r->stream = S;
    r->towers = T;
    r->n = n;
    r->d = blk->p + j * poly;
    r->owner = blk;
    r->device_newer = true;
    digits.push_back(r);
  }
  return true;
}

/* KeySwitchHYBRID::EvalFastKeySwitchCoreExt (keyswitch-hybrid.cpp:442-480): the inner product of the digits with the key,
 * both halves, over Ql u P.  digit_at(j) = KeyRef of digit j; digits that are consecutive views of one block (what
 * resident_keyswitch_digits returns) are read in place, others are gathered first. */
template <class Towers, class DigitAt, class KeyAt>
inline bool resident_keyswitch_inner(PimManager *pim, size_t np, DigitAt digit_at, size_t nkeys, KeyAt key_at, size_t sizeP, size_t alpha,
                                     const void *key_id, std::shared_ptr<Resident> &r0, std::shared_ptr<Resident> &r1,
                                     bool down = false, uint64_t t = 0) {
  /* down: EvalFastKeySwitchCore (keyswitch-hybrid.cpp:415-437) - the inner product AND ApproxModDown of both halves as one
   * call (fhe_keyswitch_fast); the results are over Ql */
  if (!pim || np == 0 || np > nkeys || sizeP == 0 || alpha == 0) return false;
  KeyRef<Towers> d0 = digit_at(0);
  const size_t T = d0.towers.size();
  if (T <= sizeP) return false;
  const size_t sizeQl = T - sizeP, n = d0.towers[0].GetRingDimension();
  if (n < 2 || np != (sizeQl + alpha - 1) / alpha) return false;
  std::shared_ptr<PimRing> ring;
  if (!keyswitch_ring(pim, d0.towers, sizeQl, key_at(0, 0).towers, sizeP, ring)) return false;
  {   /* the digits' P towers must be the key's */
    const Towers &k0 = key_at(0, 0).towers;
    const size_t sizeQ = k0.size() - sizeP;
    for (size_t i = 0; i < sizeP; i++)
      if (d0.towers[sizeQl + i].GetModulus() != k0[sizeQ + i].GetModulus()) return false;
  }
  fhe_ctx *ctx = pim->context();
  const int S = hook_stream(ctx);
  auto plan = keyswitch_for(ring, (uint32_t)sizeQl, (uint32_t)alpha);
  auto kb = key_block_for<Towers>(pim, key_at, np, sizeQl, sizeP, n, key_id, S);
  if (!kb) return false;
  const size_t poly = T * n * sizeof(uint64_t);
  std::vector<std::shared_ptr<Resident>> dd;
  bool contiguous = true;
  for (size_t j = 0; j < np; j++) {
    KeyRef<Towers> dj = digit_at(j);
    if (dj.towers.size() != T) return false;
    dd.push_back(ensure_device(pim, dj.towers, dj.res));
    if (dd[j]->d != dd[0]->d + j * poly) contiguous = false;
  }
  std::shared_ptr<DeviceBlock> gathered;
  fhe_dptr base = dd[0]->d;
  if (!contiguous) {
    gathered = std::make_shared<DeviceBlock>();
    gathered->ctx = ctx;
    PimManager::check(fhe_alloc(ctx, np * poly, &gathered->p), "pim hook: device buffer");
    for (size_t j = 0; j < np; j++) PimManager::check(fhe_d2d(ctx, gathered->p + j * poly, dd[j]->d, poly, S), "pim hook: digit gather");
    base = gathered->p;
  }
  std::shared_ptr<Resident> o0, o1;
  if (down) {
    std::vector<uint64_t> q, psi;
    for (size_t i = 0; i < sizeQl; i++) {
      q.push_back(d0.towers[i].GetModulus().ConvertToInt());
      psi.push_back(d0.towers[i].GetRootOfUnity().ConvertToInt());
    }
    auto ql = ring_for(pim, (uint32_t)n, q, psi);
    fresh_pair(ql, ctx, S, sizeQl, n, o0, o1);
    std::lock_guard<std::mutex> lock(plan.lock->mu);
    PimManager::check(fhe_keyswitch_fast(plan.plan->handle(), o0->d, o1->d, base, kb->b, kb->a, 1, t, S), "pim hook: fast key switch");
  } else {
    o0 = fresh_resident(ring, ctx, S, T, n);
    o1 = fresh_resident(ring, ctx, S, T, n);
    PimManager::check(fhe_keyswitch_inner(plan.plan->handle(), o0->d, o1->d, base, kb->b, kb->a, 1, S), "pim hook: key-switch inner product");
  }
  finish(ctx);
  r0 = o0;
  r1 = o1;
  return true;
}

/* DCRTPolyImpl::CRTDecompose(baseBits = 0) (dcrtpoly-impl.h:266-288; KeySwitchBV's digits): a -> one device copy per
 * digit, views into one [T][T][N] block (the transform of all digits is one launch sequence) */
template <class Towers>
inline bool resident_crt_decompose(PimManager *pim, const Towers &a, std::shared_ptr<Resident> &ra, int format,
                                   std::vector<std::shared_ptr<Resident>> &digits) {
  if (!pim || a.empty() || a[0].GetRingDimension() < 2) return false;
  const int S = hook_stream(pim->context());
  auto da = ensure_device(pim, a, ra);
  static std::mutex mu;
  static std::map<std::pair<PimRing *, int>, CachedPlan<PimCrtDecomp>> cache;
  CachedPlan<PimCrtDecomp> plan;
  {
    std::lock_guard<std::mutex> lock(mu);
    auto key = std::make_pair(da->ring.get(), S);
    auto it = cache.find(key);
    if (it == cache.end()) {
      CachedPlan<PimCrtDecomp> c;
      c.plan = std::make_shared<PimCrtDecomp>(da->ring);
      it = cache.emplace(key, c).first;
    }
    plan = it->second;
  }
  struct Block {
    fhe_ctx *ctx;
    fhe_dptr p = 0;
    ~Block() { if (p) fhe_free(ctx, p); }
  };
  const size_t T = a.size(), n = da->n, poly = T * n * sizeof(uint64_t);
  auto blk = std::make_shared<Block>();
  blk->ctx = da->ctx;
  PimManager::check(fhe_alloc(da->ctx, T * poly, &blk->p), "pim hook: device buffer");
  {
    std::lock_guard<std::mutex> lock(plan.lock->mu);
    PimManager::check(fhe_crt_decompose(plan.plan->handle(), blk->p, da->d, 1, format, S), "pim hook: CRT decomposition");
  }
  finish(da->ctx);
  digits.clear();
  for (size_t i = 0; i < T; i++) {
    auto r = std::make_shared<Resident>();
    r->ring = da->ring;
    r->ctx = da->ctx;
    r->stream = S;
    r->towers = T;
    r->n = n;
    r->d = blk->p + i * poly;
    r->owner = blk;
    r->device_newer = true;
    digits.push_back(r);
  }
  return true;
}

}  // namespace pimhook

#endif /* _PIM_HOOK_ */
