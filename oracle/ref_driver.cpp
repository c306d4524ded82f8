// ref_driver.cpp — thin extern "C" driver over the REFERENCE's own classes (TEST INFRASTRUCTURE ONLY).
//
// Compiled by oracle/Makefile together with the reference's sources where they lie under
// /root/reference into oracle/_ref/libopenfhe_ref.so.  Nothing here restates arithmetic: every
// result comes from the reference's DCRTPoly::operator+= / -= / *= (dcrtpoly-impl.h:398-416,
// dcrtpoly.h:142-148), DCRTPoly::SwitchFormat (dcrtpoly-impl.h:2518-2524) and the parameter
// recipe of benchmark/src/poly-benchmark-16k.cpp:82-101.  Used to (1) pin oracle/fhe_oracle.c
// and (2) serve as the "reference" CPU baseline in bench.py.
#include "lattice/lat-hal.h"
#include "math/math-hal.h"
#include "math/nbtheory.h"
#include "utils/parallel.h"
#include "utils/serial.h"

#include <chrono>
#include <functional>
#include <cstdint>
#include <cstring>
#include <memory>
#include <string>
#include <vector>
#include <omp.h>

using namespace lbcrypto;

struct ref_ctx {
    uint32_t N, T;
    std::shared_ptr<ILDCRTParams<BigInteger>> params;
};
struct ref_polys {
    ref_ctx* ctx;
    std::vector<DCRTPoly> v;
};

static thread_local std::string g_ref_err;

// The fork's PimManager / DpuMemory bookkeeping (every NativeVector constructs a PimData, mubintvecnat.h:128) is not
// thread-safe (SURVEY §8(b) "Threading"), and these methods copy towers inside OpenMP loops.  Running the call on
// thread 0 of a two-thread team makes the reference's inner `omp parallel for` regions nested, hence serial.
template <class F>
static void serial_omp(F f) {
#pragma omp parallel num_threads(2)
    {
        if (omp_get_thread_num() == 0) f();
    }
}

extern "C" {

const char* ref_last_error(void) { return g_ref_err.c_str(); }

// Live intercept (only in the hooked build, oracle/Makefile `hooked`): attach / detach the backend handle
// the way src/pke/examples/simple-integers.cpp:120-130 sketches, DCRTPolyImpl::SetPim (dcrtpoly.h:410-417).
// Returns 0, 1 on failure (message in ref_last_error), -1 when this build has no hook.
int ref_polys_set_pim(ref_polys* p, int on) {
#ifdef FHE_HOOKED
    try {
        std::shared_ptr<PimManager> h;
        if (on) h = std::shared_ptr<PimManager>(PimManager::getPim(16), [](PimManager*) {});      // 16 streams, one per host thread (pim/PimHook.h)
        for (auto& poly : p->v) poly.SetPim(h);
        return 0;
    }
    catch (const std::exception& e) {
        g_ref_err = e.what();
        return 1;
    }
#else
    (void)p; (void)on;
    return -1;
#endif
}

// resident-form counters of the hooked build: uploads / downloads of whole DCRTPolys so far (-1 without a hook)
void ref_hook_counters(long* uploads, long* downloads) {
#ifdef FHE_HOOKED
    *uploads = pimhook::resident_uploads();
    *downloads = pimhook::resident_downloads();
#else
    *uploads = *downloads = -1;
#endif
}

// poly-benchmark-16k.cpp:82-101
void ref_gen_params(uint32_t nbits, uint32_t N, uint32_t T, uint64_t* q, uint64_t* psi) {
    uint32_t m = 2 * N;
    NativeInteger first = FirstPrime<NativeInteger>(nbits, m);
    NativeInteger prev  = first;
    for (uint32_t i = 0; i < T; i++) {
        prev   = PreviousPrime<NativeInteger>(prev, m);
        q[i]   = prev.ConvertToInt();
        psi[i] = RootOfUnity<NativeInteger>(m, prev).ConvertToInt();
    }
}

uint64_t ref_root_of_unity(uint32_t m, uint64_t q) {
    return RootOfUnity<NativeInteger>(m, NativeInteger(q)).ConvertToInt();
}
uint64_t ref_first_prime(uint32_t nbits, uint32_t m) { return FirstPrime<NativeInteger>(nbits, m).ConvertToInt(); }
uint64_t ref_next_prime(uint64_t q, uint32_t m) { return NextPrime<NativeInteger>(NativeInteger(q), m).ConvertToInt(); }
uint64_t ref_previous_prime(uint64_t q, uint32_t m) { return PreviousPrime<NativeInteger>(NativeInteger(q), m).ConvertToInt(); }

ref_ctx* ref_ctx_create(uint32_t N, uint32_t T, const uint64_t* q, const uint64_t* psi) {
    auto* c = new ref_ctx{N, T, nullptr};
    std::vector<NativeInteger> moduli(T), roots(T);
    for (uint32_t i = 0; i < T; i++) {
        moduli[i] = NativeInteger(q[i]);
        roots[i]  = NativeInteger(psi[i]);
    }
    ChineseRemainderTransformFTT<NativeVector>().PreCompute(roots, 2 * N, moduli);
    c->params = std::make_shared<ILDCRTParams<BigInteger>>(2 * N, moduli, roots);
    return c;
}
void ref_ctx_destroy(ref_ctx* c) { delete c; }

// data layout [B][T][N]; format: 0 = EVALUATION, 1 = COEFFICIENT (lbcrypto::Format)
ref_polys* ref_polys_load(ref_ctx* c, const uint64_t* data, size_t B, int format) {
    auto* p  = new ref_polys{c, {}};
    Format f = format == 0 ? Format::EVALUATION : Format::COEFFICIENT;
    p->v.reserve(B);
    for (size_t b = 0; b < B; b++) {
        DCRTPoly poly(c->params, f, true);
        for (uint32_t t = 0; t < c->T; t++) {
            auto tp = c->params->GetParams()[t];
            NativeVector vec(c->N, tp->GetModulus());
            const uint64_t* src = data + (b * c->T + t) * (size_t)c->N;
            for (uint32_t j = 0; j < c->N; j++)
                vec[j] = NativeInteger(src[j]);
            NativePoly np(tp, f, false);
            np.SetValues(std::move(vec), f);
            poly.SetElementAtIndex(t, std::move(np));
        }
        p->v.push_back(std::move(poly));
    }
    return p;
}
void ref_polys_store(const ref_polys* p, uint64_t* out) {
    ref_ctx* c = p->ctx;
    for (size_t b = 0; b < p->v.size(); b++)
        for (uint32_t t = 0; t < c->T; t++) {
            const auto& vals = p->v[b].GetElementAtIndex(t).GetValues();
            uint64_t* dst    = out + (b * c->T + t) * (size_t)c->N;
            for (uint32_t j = 0; j < c->N; j++)
                dst[j] = vals[j].ConvertToInt();
        }
}
// the same with an omp-for over the batch (static schedule: poly i is read by the thread that ran ref_polys_op /
// ref_polys_switch_format on it with the same thread count, so in the hooked build its download rides that thread's stream)
void ref_polys_store_mt(const ref_polys* p, uint64_t* out, int threads) {
    ref_ctx* c = p->ctx;
    if (threads <= 1) { ref_polys_store(p, out); return; }
#pragma omp parallel for num_threads(threads) schedule(static)
    for (long b = 0; b < (long)p->v.size(); b++)
        for (uint32_t t = 0; t < c->T; t++) {
            const auto& vals = p->v[b].GetElementAtIndex(t).GetValues();
            uint64_t* dst    = out + ((size_t)b * c->T + t) * (size_t)c->N;
            for (uint32_t j = 0; j < c->N; j++)
                dst[j] = vals[j].ConvertToInt();
        }
}
void ref_polys_free(ref_polys* p) { delete p; }
int ref_polys_format(const ref_polys* p) { return p->v.empty() ? -1 : (p->v[0].GetFormat() == Format::EVALUATION ? 0 : 1); }

// a (op)= b for every poly of the batch.  op: 0 += , 1 -= , 2 *= .
// outer_threads <= 1: reference-faithful (serial over the batch, OpenFHE's own omp-for over
// towers inside each call).  outer_threads > 1: "all cores" — omp-for over the batch, the
// reference's inner tower loop then runs on the calling thread (nested parallelism is off).
// Returns elapsed seconds (steady_clock) of the op loop only.
double ref_polys_op(ref_polys* a, const ref_polys* b, int op, int outer_threads) {
    size_t B = a->v.size();
    auto t0  = std::chrono::steady_clock::now();
    if (outer_threads <= 1) {
        try {
            for (size_t i = 0; i < B; i++) {
                if (op == 0) a->v[i] += b->v[i];
                else if (op == 1) a->v[i] -= b->v[i];
                else a->v[i] *= b->v[i];
            }
        }
        catch (const std::exception& e) {     // the hooked build throws when the backend fails (no CPU fallback)
            g_ref_err = e.what();
            return -1.0;
        }
    }
    else {
#pragma omp parallel for num_threads(outer_threads) schedule(static)
        for (long i = 0; i < (long)B; i++) {
            if (op == 0) a->v[i] += b->v[i];
            else if (op == 1) a->v[i] -= b->v[i];
            else a->v[i] *= b->v[i];
        }
    }
    return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
}

// DCRTPoly::SwitchFormat on every poly of the batch (forward NTT if COEFFICIENT, inverse if EVALUATION).
double ref_polys_switch_format(ref_polys* a, int outer_threads) {
    size_t B = a->v.size();
    auto t0  = std::chrono::steady_clock::now();
    if (outer_threads <= 1) {
        try {
            for (size_t i = 0; i < B; i++) a->v[i].SwitchFormat();
        }
        catch (const std::exception& e) {
            g_ref_err = e.what();
            return -1.0;
        }
    }
    else {
#pragma omp parallel for num_threads(outer_threads) schedule(static)
        for (long i = 0; i < (long)B; i++) a->v[i].SwitchFormat();
    }
    return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
}

// out-of-place forms, for parity of Plus/Minus/Times (dcrtpoly.h:164-200)
ref_polys* ref_polys_binary(const ref_polys* a, const ref_polys* b, int op) {
    auto* r = new ref_polys{a->ctx, {}};
    for (size_t i = 0; i < a->v.size(); i++) {
        if (op == 0) r->v.push_back(a->v[i].Plus(b->v[i]));
        else if (op == 1) r->v.push_back(a->v[i].Minus(b->v[i]));
        else if (op == 2) r->v.push_back(a->v[i].Times(b->v[i]));
        else if (op == 3) r->v.push_back(a->v[i]);                       // copy construction
        else { DCRTPoly c; c = a->v[i]; r->v.push_back(std::move(c)); }  // copy assignment
    }
    return r;
}

// DCRTPoly::AutomorphismTransform (dcrtpoly-impl.h:348-369 -> poly-impl.h:312-380) on every poly of the batch.
// precomp != 0: the two-argument form with the table of PrecomputeAutoMap (nbtheory2.cpp:274-286), which is what
// the schemes call (EVALUATION only).  Returns a new batch, or NULL if the reference threw.
ref_polys* ref_polys_automorphism(const ref_polys* a, uint32_t k, int precomp) {
    auto* r = new ref_polys{a->ctx, {}};
    try {
        std::vector<uint32_t> map(a->ctx->N);
        if (precomp) PrecomputeAutoMap(a->ctx->N, k, &map);
        for (size_t i = 0; i < a->v.size(); i++)
            r->v.push_back(precomp ? a->v[i].AutomorphismTransform(k, map) : a->v[i].AutomorphismTransform(k));
        return r;
    }
    catch (const std::exception& e) {
        g_ref_err = e.what();
        delete r;
        return nullptr;
    }
}
void ref_precompute_automap(uint32_t n, uint32_t k, uint32_t* out) {
    std::vector<uint32_t> map(n);
    PrecomputeAutoMap(n, k, &map);
    std::memcpy(out, map.data(), n * sizeof(uint32_t));
}

// ---- RNS basis conversion ---------------------------------------------------------
// Tables are built the way the schemes' parameter code builds them, with the reference's BigInteger:
//   [(Q/q_i)^-1]_{q_i}, [Q/q_i]_{p_j}                 src/pke/lib/schemerns/rns-cryptoparameters.cpp:199-213
//   mu_j = floor(2^128 / p_j) as DoubleNativeInt      rns-cryptoparameters.cpp:273-283
//   [P^-1]_{q_i}                                       rns-cryptoparameters.cpp:190-197
//   rescale: [Ql*[Ql^-1]_{ql}/ql]_{qi}, [ql^-1]_{qi}  src/pke/lib/scheme/ckksrns/ckksrns-cryptoparameters.cpp:66-88
struct ref_switch_tables {
    std::vector<NativeInteger> hatInv, hatInvPrecon;
    std::vector<std::vector<NativeInteger>> hatModp;
    std::vector<DoubleNativeInt> mu;
};
static std::vector<NativeInteger> moduli_of(const ref_ctx* c) {
    std::vector<NativeInteger> m;
    for (auto& tp : c->params->GetParams()) m.push_back(tp->GetModulus());
    return m;
}
// test knob: a caller whose [ (Q/q_0)^-1 ]_{q_0} differs from the value the moduli imply (another level's table, say);
// the device hooks must notice and leave such a call on the CPU path
static int g_table_tweak = 0;
void ref_set_table_tweak(int on) { g_table_tweak = on; }
static ref_switch_tables make_switch_tables(const std::vector<NativeInteger>& mq, const std::vector<NativeInteger>& mp) {
    ref_switch_tables t;
    BigInteger Q(1);
    for (auto& q : mq) Q *= BigInteger(q);
    size_t sq = mq.size(), sp = mp.size();
    t.hatInv.resize(sq); t.hatInvPrecon.resize(sq); t.hatModp.resize(sq);
    for (size_t i = 0; i < sq; i++) {
        BigInteger hat    = Q / BigInteger(mq[i]);
        t.hatInv[i]       = hat.ModInverse(BigInteger(mq[i])).ConvertToInt();
        if (g_table_tweak && i == 0) t.hatInv[i] = t.hatInv[i].ModAdd(NativeInteger(1), mq[i]);
        t.hatInvPrecon[i] = t.hatInv[i].PrepModMulConst(mq[i]);
        t.hatModp[i].resize(sp);
        for (size_t j = 0; j < sp; j++) t.hatModp[i][j] = hat.Mod(BigInteger(mp[j])).ConvertToInt();
    }
    const BigInteger B128("340282366920938463463374607431768211456"), T64("18446744073709551616");
    t.mu.resize(sp);
    for (size_t j = 0; j < sp; j++) {
        BigInteger mu = B128 / BigInteger(mp[j]);
        uint64_t val[2] = {(mu % T64).ConvertToInt(), mu.RShift(64).ConvertToInt()};
        memcpy(&t.mu[j], val, sizeof(DoubleNativeInt));
    }
    return t;
}
// copies of the tables, for comparison with the backend's own: hat_inv [sq], hat_modp [sq][sp]
void ref_switch_tables_get(const ref_ctx* from, const ref_ctx* to, uint64_t* hat_inv, uint64_t* hat_modp) {
    auto t = make_switch_tables(moduli_of(from), moduli_of(to));
    for (size_t i = 0; i < t.hatInv.size(); i++) {
        hat_inv[i] = t.hatInv[i].ConvertToInt();
        for (size_t j = 0; j < t.hatModp[i].size(); j++) hat_modp[i * t.hatModp[i].size() + j] = t.hatModp[i][j].ConvertToInt();
    }
}

// DCRTPoly::ApproxSwitchCRTBasis (dcrtpoly-impl.h:1034-1082): batch over basis `from` -> new batch over `to`
ref_polys* ref_polys_approx_switch(const ref_polys* a, ref_ctx* to) {
    auto t  = make_switch_tables(moduli_of(a->ctx), moduli_of(to));
    auto* r = new ref_polys{to, {}};
    serial_omp([&] {
        for (auto& poly : a->v)
            r->v.push_back(poly.ApproxSwitchCRTBasis(a->ctx->params, to->params, t.hatInv, t.hatInvPrecon, t.hatModp, t.mu));
    });
    return r;
}

// DCRTPoly::ApproxModUp (dcrtpoly-impl.h:1085-1133), in place: the batch moves from basis Q to basis QP
int ref_polys_approx_mod_up(ref_polys* a, ref_ctx* P, ref_ctx* QP) {
    int rc = 0;
    std::string err;
    serial_omp([&] {
        try {
            auto t = make_switch_tables(moduli_of(a->ctx), moduli_of(P));
            for (auto& poly : a->v)
                poly.ApproxModUp(a->ctx->params, P->params, QP->params, t.hatInv, t.hatInvPrecon, t.hatModp, t.mu);
            a->ctx = QP;
        }
        catch (const std::exception& e) { err = e.what(); rc = 1; }
    });
    if (rc) g_ref_err = err;
    return rc;
}

// DCRTPoly::ApproxModDown (dcrtpoly-impl.h:1136-1175): batch over QP (EVALUATION) -> new batch over Q.
// t = 0: CKKS form; t > 0: BGV form with [t^-1]_{p_j} (bgvrns-cryptoparameters.cpp:83-88).
ref_polys* ref_polys_approx_mod_down(const ref_polys* a, ref_ctx* Q, ref_ctx* P, uint64_t t) {
    ref_polys* res = nullptr;
    std::string err;
    serial_omp([&] {
    try {
        auto mq = moduli_of(Q), mp = moduli_of(P);
        auto tb = make_switch_tables(mp, mq);
        BigInteger modP(1);
        for (auto& pj : mp) modP *= BigInteger(pj);
        std::vector<NativeInteger> PInv(mq.size()), PInvPrecon(mq.size()), tInv(mp.size()), tInvPrecon(mp.size()), tModqPrecon(mq.size());
        for (size_t i = 0; i < mq.size(); i++) {
            PInv[i]       = modP.ModInverse(BigInteger(mq[i])).ConvertToInt();
            PInvPrecon[i] = PInv[i].PrepModMulConst(mq[i]);
            if (t) tModqPrecon[i] = NativeInteger(t).PrepModMulConst(mq[i]);
        }
        if (t)
            for (size_t j = 0; j < mp.size(); j++) {
                tInv[j]       = NativeInteger(t).ModInverse(mp[j]);
                tInvPrecon[j] = tInv[j].PrepModMulConst(mp[j]);
            }
        auto* r = new ref_polys{Q, {}};
        for (auto& poly : a->v)
            r->v.push_back(poly.ApproxModDown(Q->params, P->params, PInv, PInvPrecon, tb.hatInv, tb.hatInvPrecon, tb.hatModp,
                                              tb.mu, tInv, tInvPrecon, NativeInteger(t), tModqPrecon));
        res = r;
    }
    catch (const std::exception& e) { err = e.what(); }
    });
    if (!res) g_ref_err = err;
    return res;
}

// Rescale tables of level l = T-1 (ckksrns-cryptoparameters.cpp:66-88, k = 0)
static void make_rescale_tables(const std::vector<NativeInteger>& mq, std::vector<NativeInteger>& c, std::vector<NativeInteger>& cPrecon,
                                std::vector<NativeInteger>& qlInv, std::vector<NativeInteger>& qlInvPrecon) {
    size_t l = mq.size() - 1;
    BigInteger Ql(1);
    for (size_t i = 0; i < l; i++) Ql *= BigInteger(mq[i]);
    BigInteger QlInvModql = Ql.ModInverse(BigInteger(mq[l]));
    BigInteger result     = (QlInvModql * Ql) / BigInteger(mq[l]);
    c.resize(l); cPrecon.resize(l); qlInv.resize(l); qlInvPrecon.resize(l);
    for (size_t i = 0; i < l; i++) {
        c[i]           = result.Mod(BigInteger(mq[i])).ConvertToInt();
        cPrecon[i]     = c[i].PrepModMulConst(mq[i]);
        qlInv[i]       = mq[l].ModInverse(mq[i]);
        qlInvPrecon[i] = qlInv[i].PrepModMulConst(mq[i]);
    }
}
void ref_rescale_tables_get(const ref_ctx* c, uint64_t* qlql, uint64_t* qlinv) {
    std::vector<NativeInteger> a, ap, b, bp;
    make_rescale_tables(moduli_of(c), a, ap, b, bp);
    for (size_t i = 0; i < a.size(); i++) { qlql[i] = a[i].ConvertToInt(); qlinv[i] = b[i].ConvertToInt(); }
}

// DCRTPoly::DropLastElementAndScale (dcrtpoly-impl.h:746-769, CKKS rescale; t = 0) or DCRTPoly::ModReduce
// (dcrtpoly-impl.h:792-812, BGV modulus switching; t > 0), in place: the batch loses its last tower; `rest` is a
// context over the first T-1 towers (used by ref_polys_store afterwards).
int ref_polys_rescale(ref_polys* a, ref_ctx* rest, uint64_t t) {
    int rc = 0;
    std::string err;
    serial_omp([&] {
    try {
        auto mq = moduli_of(a->ctx);
        std::vector<NativeInteger> c, cPrecon, qlInv, qlInvPrecon;
        make_rescale_tables(mq, c, cPrecon, qlInv, qlInvPrecon);
        size_t l = mq.size() - 1;
        for (auto& poly : a->v) {
            if (t == 0) {
                poly.DropLastElementAndScale(c, cPrecon, qlInv, qlInvPrecon);
            } else {                                   // bgvrns-cryptoparameters.cpp:91-99, bgvrns-leveledshe.cpp:56-58
                NativeInteger tt(t);
                std::vector<NativeInteger> tModqPrecon(mq.size());
                for (size_t i = 0; i < mq.size(); i++) tModqPrecon[i] = tt.PrepModMulConst(mq[i]);
                NativeInteger negtInv = mq[l] - tt.ModInverse(mq[l]);
                poly.ModReduce(tt, tModqPrecon, negtInv, negtInv.PrepModMulConst(mq[l]), qlInv, qlInvPrecon);
            }
        }
        a->ctx = rest;
    }
    catch (const std::exception& e) { err = e.what(); rc = 1; }
    });
    if (rc) g_ref_err = err;
    return rc;
}

// ---- BFV (HPS) set: SwitchCRTBasis / ExpandCRTBasis / ScaleAndRound ---------------------------------------
// Tables as src/pke/lib/scheme/bfvrns/bfvrns-cryptoparameters.cpp builds them for multTech == HPS:
//   qInv[i] = 1. / double(q_i)                                   :270-273
//   alphaQModp[a][j] = [a * Q]_{p_j}, a = 0..sizeQ               :343-350
//   QHatModp[j][i] = [Q/q_i]_{p_j} (note the [j][i] order)       (m_QlHatModr)
struct ref_exact_tables {
    ref_switch_tables sw;
    std::vector<std::vector<NativeInteger>> hatModpT, alphaQModp;
    std::vector<double> qInv;
};
static ref_exact_tables make_exact_tables(const std::vector<NativeInteger>& mq, const std::vector<NativeInteger>& mp) {
    ref_exact_tables t;
    t.sw = make_switch_tables(mq, mp);
    size_t sq = mq.size(), sp = mp.size();
    t.hatModpT.assign(sp, std::vector<NativeInteger>(sq));
    for (size_t i = 0; i < sq; i++)
        for (size_t j = 0; j < sp; j++) t.hatModpT[j][i] = t.sw.hatModp[i][j];
    BigInteger Q(1);
    for (auto& q : mq) Q *= BigInteger(q);
    t.alphaQModp.resize(sq + 1);
    for (size_t j = 0; j < sp; j++) {
        NativeInteger QModpj = Q.Mod(BigInteger(mp[j])).ConvertToInt();
        for (size_t a = 0; a < sq + 1; a++) t.alphaQModp[a].push_back(QModpj.ModMul(NativeInteger(a), mp[j]));
    }
    t.qInv.resize(sq);
    for (size_t i = 0; i < sq; i++) t.qInv[i] = 1. / static_cast<double>(mq[i].ConvertToInt());
    return t;
}

// DCRTPoly::SwitchCRTBasis (dcrtpoly-impl.h:1178-1230): batch over Q -> new batch over P
ref_polys* ref_polys_switch_crt_basis(const ref_polys* a, ref_ctx* to) {
    auto t  = make_exact_tables(moduli_of(a->ctx), moduli_of(to));
    auto* r = new ref_polys{to, {}};
    serial_omp([&] {
        for (auto& poly : a->v)
            r->v.push_back(poly.SwitchCRTBasis(to->params, t.sw.hatInv, t.sw.hatInvPrecon, t.hatModpT, t.alphaQModp, t.sw.mu, t.qInv));
    });
    return r;
}

// DCRTPoly::ExpandCRTBasis (dcrtpoly-impl.h:1311-1358), in place: Q -> QP; result_format 0 EVALUATION, 1 COEFFICIENT
int ref_polys_expand_crt_basis(ref_polys* a, ref_ctx* P, ref_ctx* QP, int result_format) {
    int rc = 0;
    std::string err;
    serial_omp([&] {
        try {
            auto t = make_exact_tables(moduli_of(a->ctx), moduli_of(P));
            for (auto& poly : a->v)
                poly.ExpandCRTBasis(QP->params, P->params, t.sw.hatInv, t.sw.hatInvPrecon, t.hatModpT, t.alphaQModp, t.sw.mu, t.qInv,
                                    result_format == 0 ? Format::EVALUATION : Format::COEFFICIENT);
            a->ctx = QP;
        }
        catch (const std::exception& e) { err = e.what(); rc = 1; }
    });
    if (rc) g_ref_err = err;
    return rc;
}

// DCRTPoly::ExpandCRTBasisReverseOrder (dcrtpoly-impl.h:1361-1410), in place: Q -> PQ (the P towers first)
int ref_polys_expand_crt_basis_reverse(ref_polys* a, ref_ctx* P, ref_ctx* PQ, int result_format) {
    int rc = 0;
    std::string err;
    serial_omp([&] {
        try {
            auto t = make_exact_tables(moduli_of(a->ctx), moduli_of(P));
            for (auto& poly : a->v)
                poly.ExpandCRTBasisReverseOrder(PQ->params, P->params, t.sw.hatInv, t.sw.hatInvPrecon, t.hatModpT, t.alphaQModp, t.sw.mu,
                                                t.qInv, result_format == 0 ? Format::EVALUATION : Format::COEFFICIENT);
            a->ctx = PQ;
        }
        catch (const std::exception& e) { err = e.what(); rc = 1; }
    });
    if (rc) g_ref_err = err;
    return rc;
}

// ---- BFV HPSPOVERQ / HPSPOVERQLEVELED set -----------------------------------------------------------------
// Tables of FastExpandCRTBasisPloverQ as src/pke/lib/scheme/bfvrns/bfvrns-cryptoparameters.cpp:501-523 builds them:
//   mNegRlQHatInvModq[i] = q_i - [Rl * (Q/q_i)^-1]_{q_i},  qInvModr[i][j] = [q_i^-1]_{r_j}
static void make_pl_over_q_tables(const std::vector<NativeInteger>& mq, const std::vector<NativeInteger>& mpl,
                                  std::vector<NativeInteger>& neg, std::vector<NativeInteger>& negPrecon,
                                  std::vector<std::vector<NativeInteger>>& qInvModp) {
    BigInteger Q(1), Pl(1);
    for (auto& q : mq) Q *= BigInteger(q);
    for (auto& r : mpl) Pl *= BigInteger(r);
    neg.resize(mq.size()); negPrecon.resize(mq.size()); qInvModp.resize(mq.size());
    for (size_t i = 0; i < mq.size(); i++) {
        BigInteger QHati        = Q / BigInteger(mq[i]);
        BigInteger QHatInvModqi = QHati.ModInverse(BigInteger(mq[i]));
        neg[i]                  = Pl.ModMul(QHatInvModqi, BigInteger(mq[i])).ConvertToInt();
        neg[i]                  = mq[i].Sub(neg[i]);
        negPrecon[i]            = neg[i].PrepModMulConst(mq[i]);
        qInvModp[i].resize(mpl.size());
        for (size_t j = 0; j < mpl.size(); j++) qInvModp[i][j] = mq[i].ModInverse(mpl[j]);
    }
}
void ref_pl_over_q_tables_get(const ref_ctx* Q, const ref_ctx* Pl, uint64_t* neg, uint64_t* qinv_modp) {
    std::vector<NativeInteger> n, np;
    std::vector<std::vector<NativeInteger>> qi;
    make_pl_over_q_tables(moduli_of(Q), moduli_of(Pl), n, np, qi);
    for (size_t i = 0; i < n.size(); i++) {
        neg[i] = n[i].ConvertToInt();
        for (size_t j = 0; j < qi[i].size(); j++) qinv_modp[i * qi[i].size() + j] = qi[i][j].ConvertToInt();
    }
}
// DCRTPoly::FastExpandCRTBasisPloverQ (dcrtpoly-impl.h:1413-1468), in place: Q -> QlPl
int ref_polys_fast_expand_pl_over_q(ref_polys* a, ref_ctx* Ql, ref_ctx* Pl, ref_ctx* QlPl) {
    int rc = 0;
    std::string err;
    serial_omp([&] {
        try {
            auto mq = moduli_of(a->ctx), mpl = moduli_of(Pl), mql = moduli_of(Ql);
            std::vector<NativeInteger> neg, negPrecon;
            std::vector<std::vector<NativeInteger>> qInvModp;
            make_pl_over_q_tables(mq, mpl, neg, negPrecon, qInvModp);
            auto modpMu = make_switch_tables(mq, mpl).mu;
            auto back   = make_exact_tables(mpl, mql);      // Pl -> Ql: bfvrns-cryptoparameters.cpp:390-443
            DCRTPoly::CRTBasisExtensionPrecomputations pre(QlPl->params, Pl->params, Ql->params, neg, negPrecon, qInvModp, modpMu,
                                                           back.sw.hatInv, back.sw.hatInvPrecon, back.hatModpT, back.alphaQModp,
                                                           back.sw.mu, back.qInv);
            for (auto& poly : a->v) poly.FastExpandCRTBasisPloverQ(pre);
            a->ctx = QlPl;
        }
        catch (const std::exception& e) { err = e.what(); rc = 1; }
    });
    if (rc) g_ref_err = err;
    return rc;
}
// DCRTPoly::ExpandCRTBasisQlHat (dcrtpoly-impl.h:1514-1534), in place: Ql (the first towers of Q) -> Q; c[i] < q_i
int ref_polys_expand_ql_hat(ref_polys* a, ref_ctx* Q, const uint64_t* c) {
    int rc = 0;
    std::string err;
    serial_omp([&] {
        try {
            auto mq = moduli_of(Q);
            std::vector<NativeInteger> k(a->ctx->T), kp(a->ctx->T);
            for (uint32_t i = 0; i < a->ctx->T; i++) { k[i] = NativeInteger(c[i]); kp[i] = k[i].PrepModMulConst(mq[i]); }
            for (auto& poly : a->v) poly.ExpandCRTBasisQlHat(Q->params, k, kp, Q->T);
            a->ctx = Q;
        }
        catch (const std::exception& e) { err = e.what(); rc = 1; }
    });
    if (rc) g_ref_err = err;
    return rc;
}
// DCRTPoly::Times(const std::vector<NativeInteger>&) (dcrtpoly-impl.h:617-627): new batch, one constant per tower
ref_polys* ref_polys_times_towers(const ref_polys* a, const uint64_t* c) {
    auto* r = new ref_polys{a->ctx, {}};
    std::vector<NativeInteger> k(a->ctx->T);
    for (uint32_t i = 0; i < a->ctx->T; i++) k[i] = NativeInteger(c[i]);
    serial_omp([&] {
        for (auto& poly : a->v) r->v.push_back(poly.Times(k));
    });
    return r;
}
// DCRTPoly::ScaleAndRoundPOverQ (dcrtpoly-impl.h:2052-2066), in place: Q plus one tower -> Q
int ref_polys_scale_round_p_over_q(ref_polys* a, ref_ctx* Q, const uint64_t* pinv_modq) {
    int rc = 0;
    std::string err;
    serial_omp([&] {
        try {
            std::vector<NativeInteger> k(Q->T);
            for (uint32_t i = 0; i < Q->T; i++) k[i] = NativeInteger(pinv_modq[i]);
            for (auto& poly : a->v) poly.ScaleAndRoundPOverQ(Q->params, k);
            a->ctx = Q;
        }
        catch (const std::exception& e) { err = e.what(); rc = 1; }
    });
    if (rc) g_ref_err = err;
    return rc;
}

// ScaleAndRound tables for the HPS multiplication, t * R / (Q*R) from QR down to R (bfvrns-cryptoparameters.cpp:278-311):
// frac [sizeQ], modr [sizeR][sizeQ+1] (row-major)
static void make_scale_tables(const std::vector<NativeInteger>& mq, const std::vector<NativeInteger>& mr, uint64_t tval,
                              std::vector<double>& frac, std::vector<std::vector<NativeInteger>>& modr) {
    BigInteger R(1), QR(1), t(tval);
    for (auto& r : mr) R *= BigInteger(r);
    for (auto& q : mq) QR *= BigInteger(q);
    QR *= R;
    frac.resize(mq.size());
    for (size_t i = 0; i < mq.size(); i++) {
        BigInteger qi(mq[i].ConvertToInt());
        frac[i] = static_cast<double>(((QR.DividedBy(qi)).ModInverse(qi) * R * t).Mod(qi).ConvertToInt()) /
                  static_cast<double>(qi.ConvertToInt());
    }
    modr.assign(mr.size(), {});
    for (size_t j = 0; j < mr.size(); j++) {
        BigInteger rj(mr[j].ConvertToInt());
        for (size_t i = 0; i < mq.size(); i++) {
            BigInteger qi(mq[i].ConvertToInt());
            BigInteger v = (t * R * ((QR.DividedBy(qi)).ModInverse(qi))) / qi;
            modr[j].push_back(v.Mod(rj).ConvertToInt());
        }
        BigInteger v = (t * R * ((QR.DividedBy(rj)).ModInverse(rj))) / rj;
        modr[j].push_back(v.Mod(rj).ConvertToInt());
    }
}
void ref_scale_tables_get(const ref_ctx* Q, const ref_ctx* R, uint64_t t, double* frac, uint64_t* modr) {
    std::vector<double> f;
    std::vector<std::vector<NativeInteger>> m;
    make_scale_tables(moduli_of(Q), moduli_of(R), t, f, m);
    for (size_t i = 0; i < f.size(); i++) frac[i] = f[i];
    for (size_t j = 0; j < m.size(); j++)
        for (size_t i = 0; i < m[j].size(); i++) modr[j * m[j].size() + i] = m[j][i].ConvertToInt();
}

// DCRTPoly::ScaleAndRound(paramsOutput, ...) (dcrtpoly-impl.h:1876-1960): batch over QR -> new batch over R
ref_polys* ref_polys_scale_and_round(const ref_polys* a, ref_ctx* Q, ref_ctx* R, uint64_t t) {
    std::vector<double> frac;
    std::vector<std::vector<NativeInteger>> modr;
    auto mr = moduli_of(R);
    make_scale_tables(moduli_of(Q), mr, t, frac, modr);
    auto mu = make_switch_tables(moduli_of(Q), mr).mu;
    auto* r = new ref_polys{R, {}};
    serial_omp([&] {
        for (auto& poly : a->v) r->v.push_back(poly.ScaleAndRound(R->params, modr, frac, mu));
    });
    return r;
}

// The RNS core of BFVrns EvalMult with multTech == HPS (src/pke/lib/scheme/bfvrns/bfvrns-leveledshe.cpp:193-207,
// 322-356), on the reference's own DCRTPoly: ExpandCRTBasis of the four input polynomials to QR (EVALUATION), the
// 2 x 2 tensor product (the plain product loop, :322-334), SetFormat(COEFFICIENT), ScaleAndRound to R, SwitchCRTBasis
// back to Q.  a0, a1, b0, b1: batches over Q in EVALUATION.  Returns out[0..2] (over Q, COEFFICIENT); 0 on success.
static int bfv_hps_mult_impl(const ref_polys* a0, const ref_polys* a1, const ref_polys* b0, const ref_polys* b1, ref_ctx* R,
                             ref_ctx* QR, uint64_t t, ref_polys** out, int outer_threads, double* seconds);

int ref_bfv_hps_mult(const ref_polys* a0, const ref_polys* a1, const ref_polys* b0, const ref_polys* b1, ref_ctx* R,
                     ref_ctx* QR, uint64_t t, ref_polys** out) {
    return bfv_hps_mult_impl(a0, a1, b0, b1, R, QR, t, out, 1, nullptr);
}
// Timed form for the CPU baseline of BASELINE configs[4]: outer_threads > 1 runs an omp-for over the batch (one thread per
// ciphertext pair, the reference's inner loops serial).  Only the hooked build may do that: in the plain build every
// NativeVector carries the fork's PimData member, whose bookkeeping is not thread-safe.  *seconds = the multiplication
// loop only (tables excluded, like the schemes' precomputation).
int ref_bfv_hps_mult_timed(const ref_polys* a0, const ref_polys* a1, const ref_polys* b0, const ref_polys* b1, ref_ctx* R,
                           ref_ctx* QR, uint64_t t, ref_polys** out, int outer_threads, double* seconds) {
#ifndef FHE_HOOKED
    outer_threads = 1;
#endif
    return bfv_hps_mult_impl(a0, a1, b0, b1, R, QR, t, out, outer_threads, seconds);
}

static int bfv_hps_mult_impl(const ref_polys* a0, const ref_polys* a1, const ref_polys* b0, const ref_polys* b1, ref_ctx* R,
                             ref_ctx* QR, uint64_t t, ref_polys** out, int outer_threads, double* seconds) {
    int rc = 0;
    std::string err;
    auto run = [&](auto body) {
        if (outer_threads > 1) body();
        else serial_omp(body);
    };
    run([&] {
        try {
            ref_ctx* Q = a0->ctx;
            auto mq = moduli_of(Q), mr = moduli_of(R);
            auto ex = make_exact_tables(mq, mr);          // Q -> R
            auto back = make_exact_tables(mr, mq);        // R -> Q
            std::vector<double> frac;
            std::vector<std::vector<NativeInteger>> modr;
            make_scale_tables(mq, mr, t, frac, modr);
            for (int k = 0; k < 3; k++) {
                out[k] = new ref_polys{Q, {}};
                out[k]->v.resize(a0->v.size());
            }
            auto t0 = std::chrono::steady_clock::now();
#pragma omp parallel for num_threads(outer_threads) schedule(dynamic) if (outer_threads > 1)
            for (long i = 0; i < (long)a0->v.size(); i++) {
                DCRTPoly cv1[2] = {a0->v[i], a1->v[i]}, cv2[2] = {b0->v[i], b1->v[i]};
                for (auto* cv : {cv1, cv2})
                    for (int k = 0; k < 2; k++)
                        cv[k].ExpandCRTBasis(QR->params, R->params, ex.sw.hatInv, ex.sw.hatInvPrecon, ex.hatModpT, ex.alphaQModp,
                                             ex.sw.mu, ex.qInv, Format::EVALUATION);
                DCRTPoly cvMult[3];
                bool first[3] = {true, true, true};
                for (int x = 0; x < 2; x++)
                    for (int y = 0; y < 2; y++) {
                        if (first[x + y]) { cvMult[x + y] = cv1[x] * cv2[y]; first[x + y] = false; }
                        else cvMult[x + y] += cv1[x] * cv2[y];
                    }
                for (int k = 0; k < 3; k++) {
                    cvMult[k].SetFormat(Format::COEFFICIENT);
                    cvMult[k] = cvMult[k].ScaleAndRound(R->params, modr, frac, ex.sw.mu);
                    cvMult[k] = cvMult[k].SwitchCRTBasis(Q->params, back.sw.hatInv, back.sw.hatInvPrecon, back.hatModpT,
                                                         back.alphaQModp, back.sw.mu, back.qInv);
                    out[k]->v[i] = std::move(cvMult[k]);
                }
            }
            if (seconds) *seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
        }
        catch (const std::exception& e) { err = e.what(); rc = 1; }
    });
    if (rc) g_ref_err = err;
    return rc;
}

// The RNS core of BFVrns EvalMult with multTech == HPSPOVERQ (bfvrns-leveledshe.cpp:208-235, 322-334, 357-366) on the
// reference's own DCRTPoly: cv1 ExpandCRTBasis Q -> QR, cv2 SetFormat(COEFFICIENT) + FastExpandCRTBasisPloverQ +
// SetFormat(EVALUATION), the 2 x 2 tensor product, SetFormat(COEFFICIENT), ScaleAndRound QR -> Q with the t*Q/(QR)
// tables (bfvrns-cryptoparameters.cpp:531-557; the same recipe as make_scale_tables with Q and R exchanged).
// sizeR must equal sizeQ (FastExpandCRTBasisPloverQ: sizeQl = sizePl).  out[0..2] over Q in COEFFICIENT.
static int bfv_hps_p_over_q_mult_impl(const ref_polys* a0, const ref_polys* a1, const ref_polys* b0, const ref_polys* b1, ref_ctx* R,
                                      ref_ctx* QR, uint64_t t, ref_polys** out, int outer_threads, double* seconds) {
    int rc = 0;
    std::string err;
    auto run = [&](auto body) {
        if (outer_threads > 1) body();
        else serial_omp(body);
    };
    run([&] {
        try {
            ref_ctx* Q = a0->ctx;
            auto mq = moduli_of(Q), mr = moduli_of(R);
            auto ex   = make_exact_tables(mq, mr);        // Q -> R
            auto back = make_exact_tables(mr, mq);        // R -> Q
            std::vector<NativeInteger> neg, negPrecon;
            std::vector<std::vector<NativeInteger>> qInvModr;
            make_pl_over_q_tables(mq, mr, neg, negPrecon, qInvModr);
            DCRTPoly::CRTBasisExtensionPrecomputations basisPQ(QR->params, R->params, Q->params, neg, negPrecon, qInvModr, ex.sw.mu,
                                                               back.sw.hatInv, back.sw.hatInvPrecon, back.hatModpT,
                                                               back.alphaQModp, back.sw.mu, back.qInv);
            std::vector<double> frac;
            std::vector<std::vector<NativeInteger>> modq;
            make_scale_tables(mr, mq, t, frac, modq);     // frac [sizeR], modq [sizeQ][sizeR+1]
            for (int k = 0; k < 3; k++) {
                out[k] = new ref_polys{Q, {}};
                out[k]->v.resize(a0->v.size());
            }
            auto t0 = std::chrono::steady_clock::now();
#pragma omp parallel for num_threads(outer_threads) schedule(dynamic) if (outer_threads > 1)
            for (long i = 0; i < (long)a0->v.size(); i++) {
                DCRTPoly cv1[2] = {a0->v[i], a1->v[i]}, cv2[2] = {b0->v[i], b1->v[i]};
                for (int k = 0; k < 2; k++) {
                    cv1[k].ExpandCRTBasis(QR->params, R->params, ex.sw.hatInv, ex.sw.hatInvPrecon, ex.hatModpT, ex.alphaQModp,
                                          ex.sw.mu, ex.qInv, Format::EVALUATION);
                    cv2[k].SetFormat(Format::COEFFICIENT);
                    cv2[k].FastExpandCRTBasisPloverQ(basisPQ);
                    cv2[k].SetFormat(Format::EVALUATION);
                }
                DCRTPoly cvMult[3];
                bool first[3] = {true, true, true};
                for (int x = 0; x < 2; x++)
                    for (int y = 0; y < 2; y++) {
                        if (first[x + y]) { cvMult[x + y] = cv1[x] * cv2[y]; first[x + y] = false; }
                        else cvMult[x + y] += cv1[x] * cv2[y];
                    }
                for (int k = 0; k < 3; k++) {
                    cvMult[k].SetFormat(Format::COEFFICIENT);
                    out[k]->v[i] = cvMult[k].ScaleAndRound(Q->params, modq, frac, back.sw.mu);
                }
            }
            if (seconds) *seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
        }
        catch (const std::exception& e) { err = e.what(); rc = 1; }
    });
    if (rc) g_ref_err = err;
    return rc;
}
int ref_bfv_hps_p_over_q_mult(const ref_polys* a0, const ref_polys* a1, const ref_polys* b0, const ref_polys* b1, ref_ctx* R,
                              ref_ctx* QR, uint64_t t, ref_polys** out) {
    return bfv_hps_p_over_q_mult_impl(a0, a1, b0, b1, R, QR, t, out, 1, nullptr);
}
// timed form, as ref_bfv_hps_mult_timed
int ref_bfv_hps_p_over_q_mult_timed(const ref_polys* a0, const ref_polys* a1, const ref_polys* b0, const ref_polys* b1, ref_ctx* R,
                                    ref_ctx* QR, uint64_t t, ref_polys** out, int outer_threads, double* seconds) {
#ifndef FHE_HOOKED
    outer_threads = 1;
#endif
    return bfv_hps_p_over_q_mult_impl(a0, a1, b0, b1, R, QR, t, out, outer_threads, seconds);
}

// mod-ops of the PimData surface executed by the reference's own NativeVector (in-place forms:
// mubintvecnat.cpp:111-195,370-487, mubintvecnat.h:560-566).  kind = enum mod_kernel.
// Returns 0, or 1 if the reference threw.
int ref_vec_modop(int kind, uint64_t* data, size_t n, uint64_t mod, uint64_t p, uint64_t q) {
    try {
        NativeVector v(n, NativeInteger(mod));
        for (size_t i = 0; i < n; i++) v[i] = NativeInteger(data[i]);
        switch (kind) {
            case 0: v.SwitchModulus(NativeInteger(p)); break;
            case 1: case 2: v.ModEq(NativeInteger(p)); break;
            case 3: case 4: v.ModByTwoEq(); break;
            case 5: case 6: v.ModExpEq(NativeInteger(p)); break;
            case 7: case 8: v.ModInverseEq(); break;
            case 9: case 10: v.MultiplyAndRoundEq(NativeInteger(p), NativeInteger(q)); break;
            case 11: case 12: v.DivideAndRoundEq(NativeInteger(p)); break;
            default: return 1;
        }
        for (size_t i = 0; i < n; i++) data[i] = v[i].ConvertToInt();
        return 0;
    }
    catch (...) {
        return 1;
    }
}

// The reference's own binary wire format of one DCRTPoly (Serial::Serialize(..., SerType::BINARY), utils/serial.h:99-103:
// a cereal PortableBinaryOutputArchive over DCRTPolyImpl::save, dcrtpoly.h:349-354).  Returns the byte count; copies
// min(count, cap) bytes to out.
size_t ref_polys_serialize(const ref_polys* p, size_t index, uint8_t* out, size_t cap) {
    try {
        std::stringstream ss;
        Serial::Serialize(p->v.at(index), ss, SerType::BINARY);
        std::string s = ss.str();
        if (out && cap) std::memcpy(out, s.data(), std::min(cap, s.size()));
        return s.size();
    }
    catch (const std::exception& e) {
        g_ref_err = e.what();
        return 0;
    }
}
// Same poly, but every tower owns a private copy of its params, so the stream carries the tower params of
// ILDCRTParams inline instead of by id (a shape the reference's reader accepts and other producers may write).
size_t ref_polys_serialize_detached(const ref_polys* p, size_t index, uint8_t* out, size_t cap) {
    try {
        DCRTPoly poly(p->v.at(index));
        for (uint32_t t = 0; t < p->ctx->T; t++) {
            const NativePoly& src = poly.GetElementAtIndex(t);
            auto own = std::make_shared<ILNativeParams>(*src.GetParams());
            NativePoly np(own, src.GetFormat(), false);
            np.SetValues(src.GetValues(), src.GetFormat());
            poly.SetElementAtIndex(t, std::move(np));
        }
        std::stringstream ss;
        Serial::Serialize(poly, ss, SerType::BINARY);
        std::string s = ss.str();
        if (out && cap) std::memcpy(out, s.data(), std::min(cap, s.size()));
        return s.size();
    }
    catch (const std::exception& e) {
        g_ref_err = e.what();
        return 0;
    }
}
// Serial::Deserialize of one DCRTPoly (utils/serial.h:111-115).  The handle keeps ctx c for (N, T) only: the poly carries
// the params it was serialised with.  NULL if the reference threw or the shape differs from c.
ref_polys* ref_polys_deserialize(ref_ctx* c, const uint8_t* bytes, size_t len) {
    try {
        std::stringstream ss(std::string(reinterpret_cast<const char*>(bytes), len));
        DCRTPoly poly;
        Serial::Deserialize(poly, ss, SerType::BINARY);
        if (poly.GetNumOfElements() != c->T || poly.GetRingDimension() != c->N) {
            g_ref_err = "deserialised shape differs from the context";
            return nullptr;
        }
        auto* p = new ref_polys{c, {}};
        p->v.push_back(std::move(poly));
        return p;
    }
    catch (const std::exception& e) {
        g_ref_err = e.what();
        return nullptr;
    }
}

// ---- BFV decryption: DCRTPoly::ScaleAndRound(t, ...) -> NativePoly (dcrtpoly-impl.h:1537-1816) ------------------------
// Tables built with the reference's BigInteger exactly as bfvrns-cryptoparameters.cpp:447-494 builds them.
// out [B][N]; the tables come back for comparison with the restated ones (bmodt / bfrac are zero when the split
// branch is not taken).
int ref_polys_scale_and_round_native(const ref_polys* a, uint64_t t, uint64_t* out, uint64_t* modt_out, uint64_t* bmodt_out,
                                     double* frac_out, double* bfrac_out) {
    int rc = 0;
    std::string err;
    serial_omp([&] {
        try {
            ref_ctx* Q   = a->ctx;
            auto moduliQ = moduli_of(Q);
            size_t sizeQ = moduliQ.size();
            BigInteger modulusQ(1);
            for (auto& qi : moduliQ) modulusQ *= BigInteger(qi);
            const BigInteger T(t);
            usint qMSB = moduliQ[0].GetMSB(), sizeQMSB = GetMSB64(sizeQ);
            std::vector<NativeInteger> modt(sizeQ), modtPrecon(sizeQ), bmodt, bmodtPrecon;
            std::vector<double> frac(sizeQ), bfrac;
            const bool split = !(qMSB + sizeQMSB < 52);
            if (split) { bmodt.resize(sizeQ); bmodtPrecon.resize(sizeQ); bfrac.resize(sizeQ); }
            usint qMSBHf = qMSB >> 1;
            for (size_t i = 0; i < sizeQ; i++) {
                BigInteger qi(moduliQ[i].ConvertToInt());
                BigInteger tQHatInvModqi     = ((modulusQ.DividedBy(qi)).ModInverse(qi) * T);
                BigInteger tQHatInvModqDivqi = tQHatInvModqi.DividedBy(qi);
                modt[i]                      = tQHatInvModqDivqi.Mod(T).ConvertToInt();
                modtPrecon[i]                = modt[i].PrepModMulConst(NativeInteger(t));
                int64_t numerator            = tQHatInvModqi.Mod(qi).ConvertToInt();
                int64_t denominator          = moduliQ[i].ConvertToInt();
                frac[i]                      = static_cast<double>(numerator) / static_cast<double>(denominator);
                if (split) {
                    tQHatInvModqi.LShiftEq(qMSBHf);
                    tQHatInvModqDivqi = tQHatInvModqi.DividedBy(qi);
                    bmodt[i]          = tQHatInvModqDivqi.Mod(T).ConvertToInt();
                    bmodtPrecon[i]    = bmodt[i].PrepModMulConst(NativeInteger(t));
                    numerator         = tQHatInvModqi.Mod(qi).ConvertToInt();
                    bfrac[i]          = static_cast<double>(numerator) / static_cast<double>(denominator);
                }
                modt_out[i] = modt[i].ConvertToInt(); frac_out[i] = frac[i];
                bmodt_out[i] = split ? bmodt[i].ConvertToInt() : 0; bfrac_out[i] = split ? bfrac[i] : 0.0;
            }
            size_t b = 0;
            for (auto& poly : a->v) {
                NativePoly r = poly.ScaleAndRound(NativeInteger(t), modt, modtPrecon, bmodt, bmodtPrecon, frac, bfrac);
                const auto& vals = r.GetValues();
                for (uint32_t j = 0; j < Q->N; j++) out[b * Q->N + j] = vals[j].ConvertToInt();
                b++;
            }
        }
        catch (const std::exception& e) { err = e.what(); rc = 1; }
    });
    if (rc) g_ref_err = err;
    return rc;
}

// ---- hybrid key switching on the reference's own DCRTPoly -----------------------------------------------------------
// The statements of KeySwitchHYBRID::EvalKeySwitchPrecomputeCore / EvalFastKeySwitchCoreExt / EvalFastKeySwitchCore
// (src/pke/lib/keyswitch/keyswitch-hybrid.cpp:330-480) — src/pke itself is not compiled here — with the tables the
// scheme parameters would hold (rns-cryptoparameters.cpp:236-340: the products run over the (truncated) digit basis,
// i.e. make_switch_tables of its moduli).  c over Ql (EVALUATION); QlP = Ql towers then P; the key polys evk_b / evk_a
// are given over the towers of QlP ([numPart][T][N], EVALUATION).  out[0], out[1] over Ql.
static std::shared_ptr<ILDCRTParams<BigInteger>> sub_params(const ref_ctx* c, const std::vector<uint32_t>& idx) {
    std::vector<NativeInteger> moduli, roots;
    for (uint32_t i : idx) {
        moduli.push_back(c->params->GetParams()[i]->GetModulus());
        roots.push_back(c->params->GetParams()[i]->GetRootOfUnity());
    }
    return std::make_shared<ILDCRTParams<BigInteger>>(2 * c->N, moduli, roots);
}
static std::vector<NativeInteger> moduli_of_params(const std::shared_ptr<ILDCRTParams<BigInteger>>& p) {
    std::vector<NativeInteger> m;
    for (auto& tp : p->GetParams()) m.push_back(tp->GetModulus());
    return m;
}

int ref_keyswitch_hybrid(const ref_polys* c, ref_ctx* QlP, uint32_t sizeQl, uint32_t alpha, const uint64_t* evk_b,
                         const uint64_t* evk_a, uint64_t t, ref_polys** out, uint64_t* digits_out) {
    int rc = 0;
    std::string err;
    serial_omp([&] {
        try {
            ref_ctx* Ql       = c->ctx;
            const uint32_t T = QlP->T, N = QlP->N, sizeP = T - sizeQl;
            const uint32_t numPartQl = (sizeQl + alpha - 1) / alpha;
            auto paramsQlP = QlP->params;
            std::vector<uint32_t> pidx;
            for (uint32_t i = sizeQl; i < T; i++) pidx.push_back(i);
            auto paramsP = sub_params(QlP, pidx);
            // key polys over QlP
            auto load_key = [&](const uint64_t* data, uint32_t j) {
                DCRTPoly k(paramsQlP, Format::EVALUATION, true);
                for (uint32_t i = 0; i < T; i++) {
                    auto tp = paramsQlP->GetParams()[i];
                    NativeVector vec(N, tp->GetModulus());
                    const uint64_t* src = data + ((size_t)j * T + i) * N;
                    for (uint32_t n = 0; n < N; n++) vec[n] = NativeInteger(src[n]);
                    NativePoly np(tp, Format::EVALUATION, false);
                    np.SetValues(std::move(vec), Format::EVALUATION);
                    k.SetElementAtIndex(i, std::move(np));
                }
                return k;
            };
            std::vector<DCRTPoly> bv, av;
            for (uint32_t j = 0; j < numPartQl; j++) { bv.push_back(load_key(evk_b, j)); av.push_back(load_key(evk_a, j)); }
            // ApproxModDown tables (rns-cryptoparameters.cpp:190-229)
            auto mq = moduli_of(Ql), mp = moduli_of_params(paramsP);
            auto tb = make_switch_tables(mp, mq);
            BigInteger modP(1);
            for (auto& pj : mp) modP *= BigInteger(pj);
            std::vector<NativeInteger> PInv(mq.size()), PInvPrecon(mq.size()), tInv(mp.size()), tInvPrecon(mp.size()), tModqPrecon(mq.size());
            for (size_t i = 0; i < mq.size(); i++) {
                PInv[i]       = modP.ModInverse(BigInteger(mq[i])).ConvertToInt();
                PInvPrecon[i] = PInv[i].PrepModMulConst(mq[i]);
                if (t) tModqPrecon[i] = NativeInteger(t).PrepModMulConst(mq[i]);
            }
            if (t)
                for (size_t j = 0; j < mp.size(); j++) {
                    tInv[j]       = NativeInteger(t).ModInverse(mp[j]);
                    tInvPrecon[j] = tInv[j].PrepModMulConst(mp[j]);
                }
            for (int k = 0; k < 2; k++) out[k] = new ref_polys{Ql, {}};
            size_t bi = 0;
            for (const DCRTPoly& cpoly : c->v) {
                // ---- EvalKeySwitchPrecomputeCore, :352-417 ----
                std::vector<DCRTPoly> partsCt(numPartQl), partsCtCompl(numPartQl), partsCtExt(numPartQl);
                std::vector<std::shared_ptr<ILDCRTParams<BigInteger>>> partParams(numPartQl), complParams(numPartQl);
                for (uint32_t part = 0; part < numPartQl; part++) {
                    uint32_t start = alpha * part, sizePartQl = std::min(alpha, sizeQl - start);
                    std::vector<uint32_t> idx, cidx;
                    for (uint32_t i = 0; i < sizePartQl; i++) idx.push_back(start + i);
                    for (uint32_t i = 0; i < T; i++)
                        if (i < start || i >= start + sizePartQl) cidx.push_back(i);
                    partParams[part]  = sub_params(QlP, idx);
                    complParams[part] = sub_params(QlP, cidx);
                    partsCt[part]     = DCRTPoly(partParams[part], Format::EVALUATION, true);
                    for (uint32_t i = 0, id = start; i < sizePartQl; i++, id++)
                        partsCt[part].SetElementAtIndex(i, cpoly.GetElementAtIndex(id));
                }
                for (uint32_t part = 0; part < numPartQl; part++) {
                    auto partCtClone = partsCt[part].Clone();
                    partCtClone.SetFormat(Format::COEFFICIENT);
                    uint32_t sizePartQl = partsCt[part].GetNumOfElements();
                    auto tbl = make_switch_tables(moduli_of_params(partParams[part]), moduli_of_params(complParams[part]));
                    partsCtCompl[part] = partCtClone.ApproxSwitchCRTBasis(partParams[part], complParams[part], tbl.hatInv,
                                                                          tbl.hatInvPrecon, tbl.hatModp, tbl.mu);
                    partsCtCompl[part].SetFormat(Format::EVALUATION);
                    partsCtExt[part] = DCRTPoly(paramsQlP, Format::EVALUATION, true);
                    uint32_t startPartIdx = alpha * part, endPartIdx = startPartIdx + sizePartQl;
                    for (uint32_t i = 0; i < startPartIdx; i++)
                        partsCtExt[part].SetElementAtIndex(i, partsCtCompl[part].GetElementAtIndex(i));
                    for (uint32_t i = startPartIdx, id = 0; i < endPartIdx; i++, id++)
                        partsCtExt[part].SetElementAtIndex(i, partsCt[part].GetElementAtIndex(id));
                    for (uint32_t i = endPartIdx; i < T; ++i)
                        partsCtExt[part].SetElementAtIndex(i, partsCtCompl[part].GetElementAtIndex(i - sizePartQl));
                    if (digits_out)
                        for (uint32_t i = 0; i < T; i++) {
                            const auto& vals = partsCtExt[part].GetElementAtIndex(i).GetValues();
                            uint64_t* dst    = digits_out + (((size_t)part * c->v.size() + bi) * T + i) * N;
                            for (uint32_t n = 0; n < N; n++) dst[n] = vals[n].ConvertToInt();
                        }
                }
                // ---- EvalFastKeySwitchCoreExt, :454-476 (the key towers are already those of QlP) ----
                DCRTPoly cTilda0(paramsQlP, Format::EVALUATION, true), cTilda1(paramsQlP, Format::EVALUATION, true);
                for (uint32_t j = 0; j < numPartQl; j++)
                    for (uint32_t i = 0; i < T; i++) {
                        const auto& cji = partsCtExt[j].GetElementAtIndex(i);
                        cTilda0.SetElementAtIndex(i, cTilda0.GetElementAtIndex(i) + cji * bv[j].GetElementAtIndex(i));
                        cTilda1.SetElementAtIndex(i, cTilda1.GetElementAtIndex(i) + cji * av[j].GetElementAtIndex(i));
                    }
                // ---- EvalFastKeySwitchCore, :430-440 ----
                out[0]->v.push_back(cTilda0.ApproxModDown(Ql->params, paramsP, PInv, PInvPrecon, tb.hatInv, tb.hatInvPrecon, tb.hatModp,
                                                          tb.mu, tInv, tInvPrecon, NativeInteger(t), tModqPrecon));
                out[1]->v.push_back(cTilda1.ApproxModDown(Ql->params, paramsP, PInv, PInvPrecon, tb.hatInv, tb.hatInvPrecon, tb.hatModp,
                                                          tb.mu, tInv, tInvPrecon, NativeInteger(t), tModqPrecon));
                bi++;
            }
        }
        catch (const std::exception& e) { err = e.what(); rc = 1; }
    });
    if (rc) g_ref_err = err;
    return rc;
}

int ref_max_threads(void) { return omp_get_max_threads(); }
// Thread count for the reference's own `omp parallel for` regions (the batch loops above pass theirs explicitly).
// OpenFHE caps its tower loops at the value omp_get_max_threads() had when the library was loaded
// (ParallelControls::machineThreads, utils/parallel.h:52), so OMP_NUM_THREADS must also be right at load time.
void ref_set_threads(int n) { if (n > 0) omp_set_num_threads(n); }

// ---- BEHZ: FastBaseConvqToBskMontgomery / FastRNSFloorq / FastBaseConvSK / ScaleAndRound(moduliQ, t, tgamma, ...) ----
// The tables are the ones CryptoParametersBFVRNS::PrecomputeCRTTables builds for multTech == BEHZ
// (src/pke/lib/scheme/bfvrns/bfvrns-cryptoparameters.cpp:662-885; the pke library is not part of this build), computed
// here with the reference's BigInteger from the same definitions; mtilde = 2^16, gamma = 2^26
// (src/pke/include/scheme/bfvrns/bfvrns-cryptoparameters.h).  QBsk = Q followed by B and msk.
struct ref_behz {
    ref_ctx* QBsk;
    std::shared_ptr<ILDCRTParams<BigInteger>> paramsQ;
    uint32_t nq, nb;
    NativeInteger t, tgamma, BInvModmsk, BInvModmskPrecon;
    uint64_t negQInvModmtilde;
    std::vector<NativeInteger> moduliQ, moduliBsk;
    std::vector<DoubleNativeInt> modqMu, modbskMu;
    std::vector<NativeInteger> mtildeQHatInvModq, mtildeQHatInvModqPrecon, QModbsk, QModbskPrecon, mtildeInvModbsk,
        mtildeInvModbskPrecon, tQHatInvModq, tQHatInvModqPrecon, tQInvModbsk, tQInvModbskPrecon, BHatInvModb, BHatInvModbPrecon,
        BHatModmsk, BModq, BModqPrecon, negInvqModtgamma, negInvqModtgammaPrecon, tgammaQHatInvModq, tgammaQHatInvModqPrecon;
    std::vector<std::vector<NativeInteger>> QHatModbsk, qInvModbsk, BHatModq;
    std::vector<uint64_t> QHatModmtilde;
};
static std::vector<DoubleNativeInt> barrett_mu_128(const std::vector<NativeInteger>& m) {
    const BigInteger B128("340282366920938463463374607431768211456"), T64("18446744073709551616");
    std::vector<DoubleNativeInt> mu(m.size());
    for (size_t j = 0; j < m.size(); j++) {
        BigInteger v = B128 / BigInteger(m[j]);
        uint64_t val[2] = {(v % T64).ConvertToInt(), v.RShift(64).ConvertToInt()};
        memcpy(&mu[j], val, sizeof(DoubleNativeInt));
    }
    return mu;
}
static void with_precon(const std::vector<NativeInteger>& v, const std::vector<NativeInteger>& m, std::vector<NativeInteger>& pre) {
    pre.resize(v.size());
    for (size_t i = 0; i < v.size(); i++) pre[i] = v[i].PrepModMulConst(m[i]);
}

ref_behz* ref_behz_create(ref_ctx* QBsk, uint32_t nq, uint64_t t) {
    auto* z = new ref_behz;
    z->QBsk = QBsk; z->nq = nq; z->nb = QBsk->T - nq; z->t = NativeInteger(t);
    const uint32_t nb = z->nb, nB = nb - 1;
    auto all = moduli_of(QBsk);
    z->moduliQ.assign(all.begin(), all.begin() + nq);
    z->moduliBsk.assign(all.begin() + nq, all.end());
    std::vector<NativeInteger> rootsQ;
    for (uint32_t i = 0; i < nq; i++) rootsQ.push_back(QBsk->params->GetParams()[i]->GetRootOfUnity());
    z->paramsQ = std::make_shared<ILDCRTParams<BigInteger>>(2 * QBsk->N, z->moduliQ, rootsQ);
    const auto& mq = z->moduliQ;
    const auto& mb = z->moduliBsk;
    const NativeInteger msk = mb[nB];
    const uint64_t mtilde = (uint64_t)1 << 16, gamma = (uint64_t)1 << 26;
    BigInteger Q(1), B(1);
    for (auto& qi : mq) Q *= BigInteger(qi);
    for (uint32_t i = 0; i < nB; i++) B *= BigInteger(mb[i]);
    z->modqMu   = barrett_mu_128(mq);
    z->modbskMu = barrett_mu_128(mb);
    const BigInteger bt(t), bmtilde(mtilde), bgamma(gamma);
    z->tgamma = NativeInteger(t * gamma);
    const BigInteger btgamma(z->tgamma);
    z->mtildeQHatInvModq.resize(nq); z->tQHatInvModq.resize(nq); z->QHatModbsk.resize(nq); z->QHatModmtilde.resize(nq);
    z->qInvModbsk.resize(nq); z->negInvqModtgamma.resize(nq); z->tgammaQHatInvModq.resize(nq); z->BModq.resize(nq);
    for (uint32_t i = 0; i < nq; i++) {
        const BigInteger qi(mq[i]);
        const BigInteger QHat = Q / qi;
        const BigInteger QHatInv = QHat.Mod(qi).ModInverse(qi);
        z->mtildeQHatInvModq[i] = (QHatInv * bmtilde).Mod(qi).ConvertToInt();
        z->tQHatInvModq[i]      = QHatInv.ModMul(bt, qi).ConvertToInt();
        z->tgammaQHatInvModq[i] = (((QHatInv * bgamma) % qi) * bt % qi).ConvertToInt();
        z->QHatModmtilde[i]     = QHat.Mod(bmtilde).ConvertToInt();
        z->negInvqModtgamma[i]  = (BigInteger(z->tgamma - NativeInteger(1)) * qi.ModInverse(btgamma)).Mod(btgamma).ConvertToInt();
        z->BModq[i]             = B.Mod(qi).ConvertToInt();
        z->QHatModbsk[i].resize(nb); z->qInvModbsk[i].resize(nb);
        for (uint32_t j = 0; j < nb; j++) {
            z->QHatModbsk[i][j] = QHat.Mod(BigInteger(mb[j])).ConvertToInt();
            z->qInvModbsk[i][j] = mq[i].ModInverse(mb[j]);
        }
    }
    z->negQInvModmtilde = (BigInteger(mtilde - 1) * Q.ModInverse(bmtilde)).Mod(bmtilde).ConvertToInt();
    z->QModbsk.resize(nb); z->mtildeInvModbsk.resize(nb); z->tQInvModbsk.resize(nb);
    for (uint32_t j = 0; j < nb; j++) {
        const BigInteger bj(mb[j]);
        z->QModbsk[j]         = Q.Mod(bj).ConvertToInt();
        z->mtildeInvModbsk[j] = (bmtilde % bj).ModInverse(bj).ConvertToInt();
        z->tQInvModbsk[j]     = Q.ModInverse(bj).ModMul(bt, bj).ConvertToInt();
    }
    z->BHatInvModb.resize(nB); z->BHatModq.resize(nB); z->BHatModmsk.resize(nB);
    for (uint32_t i = 0; i < nB; i++) {
        const BigInteger bi(mb[i]);
        const BigInteger BHat = B / bi;
        z->BHatInvModb[i] = BHat.Mod(bi).ModInverse(bi).ConvertToInt();
        z->BHatModmsk[i]  = BHat.Mod(BigInteger(msk)).ConvertToInt();
        z->BHatModq[i].resize(nq);
        for (uint32_t j = 0; j < nq; j++) z->BHatModq[i][j] = BHat.Mod(BigInteger(mq[j])).ConvertToInt();
    }
    z->BInvModmsk       = B.ModInverse(BigInteger(msk)).ConvertToInt();
    z->BInvModmskPrecon = z->BInvModmsk.PrepModMulConst(msk);
    with_precon(z->mtildeQHatInvModq, mq, z->mtildeQHatInvModqPrecon);
    with_precon(z->tQHatInvModq, mq, z->tQHatInvModqPrecon);
    with_precon(z->tgammaQHatInvModq, mq, z->tgammaQHatInvModqPrecon);
    with_precon(z->BModq, mq, z->BModqPrecon);
    with_precon(z->QModbsk, mb, z->QModbskPrecon);
    with_precon(z->mtildeInvModbsk, mb, z->mtildeInvModbskPrecon);
    with_precon(z->tQInvModbsk, mb, z->tQInvModbskPrecon);
    with_precon(z->BHatInvModb, mb, z->BHatInvModbPrecon);
    z->negInvqModtgammaPrecon.resize(nq);
    for (uint32_t i = 0; i < nq; i++) z->negInvqModtgammaPrecon[i] = z->negInvqModtgamma[i].PrepModMulConst(z->tgamma);
    return z;
}
void ref_behz_destroy(ref_behz* z) { delete z; }
// table `which` in the numbering of oracle/fhe_oracle.h (orc_behz_table); returns the length in words
size_t ref_behz_table(const ref_behz* z, int which, uint64_t* out) {
    std::vector<uint64_t> v;
    auto flat = [&](const std::vector<NativeInteger>& a) { for (auto& x : a) v.push_back(x.ConvertToInt()); };
    auto flat2 = [&](const std::vector<std::vector<NativeInteger>>& a) { for (auto& r : a) flat(r); };
    switch (which) {
        case 0: flat(z->mtildeQHatInvModq); break;
        case 1: flat2(z->QHatModbsk); break;
        case 2: v = z->QHatModmtilde; break;
        case 3: flat(z->QModbsk); break;
        case 4: v.push_back(z->negQInvModmtilde); break;
        case 5: flat(z->mtildeInvModbsk); break;
        case 6: flat(z->tQHatInvModq); break;
        case 7: flat2(z->qInvModbsk); break;
        case 8: flat(z->tQInvModbsk); break;
        case 9: flat(z->BHatInvModb); break;
        case 10: flat2(z->BHatModq); break;
        case 11: flat(z->BHatModmsk); break;
        case 12: v.push_back(z->BInvModmsk.ConvertToInt()); break;
        case 13: flat(z->BModq); break;
        case 14: v.push_back(z->tgamma.ConvertToInt()); break;
        case 15: flat(z->negInvqModtgamma); break;
        case 16: flat(z->tgammaQHatInvModq); break;
        default: return 0;
    }
    if (out) memcpy(out, v.data(), v.size() * 8);
    return v.size();
}

static int behz_run(const std::function<void()>& f) {
    int rc = 0;
    std::string err;
    serial_omp([&] {
        try { f(); }
        catch (const std::exception& e) { err = e.what(); rc = 1; }
    });
    if (rc) g_ref_err = err;
    return rc;
}
// DCRTPoly::FastBaseConvqToBskMontgomery (dcrtpoly-impl.h:2069-2209), in place: the batch moves from Q to QBsk (EVALUATION)
int ref_polys_behz_q_to_bsk(ref_polys* a, ref_behz* z) {
    return behz_run([&] {
        for (auto& poly : a->v)
            poly.FastBaseConvqToBskMontgomery(z->QBsk->params, z->moduliQ, z->moduliBsk, z->modbskMu, z->mtildeQHatInvModq,
                                              z->mtildeQHatInvModqPrecon, z->QHatModbsk, z->QHatModmtilde, z->QModbsk, z->QModbskPrecon,
                                              z->negQInvModmtilde, z->mtildeInvModbsk, z->mtildeInvModbskPrecon);
        a->ctx = z->QBsk;
    });
}
// DCRTPoly::FastRNSFloorq (dcrtpoly-impl.h:2212-2271), in place over QBsk
int ref_polys_behz_floor(ref_polys* a, ref_behz* z) {
    return behz_run([&] {
        for (auto& poly : a->v)
            poly.FastRNSFloorq(z->t, z->moduliQ, z->moduliBsk, z->modbskMu, z->tQHatInvModq, z->tQHatInvModqPrecon, z->QHatModbsk,
                               z->qInvModbsk, z->tQInvModbsk, z->tQInvModbskPrecon);
    });
}
// DCRTPoly::FastBaseConvSK (dcrtpoly-impl.h:2309-2420), in place: the batch moves from QBsk to Q (`Q` = a context over the
// first nq towers, used by ref_polys_store afterwards)
int ref_polys_behz_conv_sk(ref_polys* a, ref_behz* z, ref_ctx* Q) {
    return behz_run([&] {
        for (auto& poly : a->v)
            poly.FastBaseConvSK(z->paramsQ, z->modqMu, z->moduliBsk, z->modbskMu, z->BHatInvModb, z->BHatInvModbPrecon, z->BHatModmsk,
                                z->BInvModmsk, z->BInvModmskPrecon, z->BHatModq, z->BModq, z->BModqPrecon);
        a->ctx = Q;
    });
}
// DCRTPoly::ScaleAndRound(moduliQ, t, tgamma, ...) -> NativePoly (dcrtpoly-impl.h:2005-2050): out [B][N]
int ref_polys_behz_scale_and_round(const ref_polys* a, ref_behz* z, uint64_t* out) {
    return behz_run([&] {
        const uint32_t N = a->ctx->N;
        size_t b = 0;
        for (auto& poly : a->v) {
            NativePoly r = poly.ScaleAndRound(z->moduliQ, z->t, z->tgamma, z->tgammaQHatInvModq, z->tgammaQHatInvModqPrecon,
                                              z->negInvqModtgamma, z->negInvqModtgammaPrecon);
            const auto& vals = r.GetValues();
            for (uint32_t j = 0; j < N; j++) out[b * N + j] = vals[j].ConvertToInt();
            b++;
        }
    });
}

// DCRTPoly::CRTDecompose(0) (dcrtpoly-impl.h:266-288): digit `digit` of every polynomial of the batch, as a new batch
ref_polys* ref_polys_crt_decompose(const ref_polys* a, uint32_t digit) {
    auto* r = new ref_polys{a->ctx, {}};
    serial_omp([&] {
        for (auto& poly : a->v) {
            auto digits = poly.CRTDecompose(0);
            r->v.push_back(digits.at(digit));
        }
    });
    return r;
}
// KeySwitchBV::EvalFastKeySwitchCore (keyswitch-bv.cpp:314-338) on DCRTPolys of this driver: c over Q (EVALUATION),
// keys kb / ka = T polynomials over Q each (batch handles of T polys); returns ct0, ct1 (batch of two)
ref_polys* ref_keyswitch_bv(const ref_polys* c, const ref_polys* kb, const ref_polys* ka) {
    auto* r = new ref_polys{c->ctx, {}};
    serial_omp([&] {
        auto digits = c->v.at(0).CRTDecompose(0);
        std::vector<DCRTPoly> bv(kb->v), av(ka->v);
        DCRTPoly ct1 = (av[0] *= digits[0]);
        DCRTPoly ct0 = (bv[0] *= digits[0]);
        for (usint i = 1; i < digits.size(); ++i) {
            ct0 += (bv[i] *= digits[i]);
            ct1 += (av[i] *= digits[i]);
        }
        r->v.push_back(ct0);
        r->v.push_back(ct1);
    });
    return r;
}

}  // extern "C"
