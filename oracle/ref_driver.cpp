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

#include <chrono>
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

extern "C" {

const char* ref_last_error(void) { return g_ref_err.c_str(); }

// Live intercept (only in the hooked build, oracle/Makefile `hooked`): attach / detach the backend handle
// the way src/pke/examples/simple-integers.cpp:120-130 sketches, DCRTPolyImpl::SetPim (dcrtpoly.h:410-417).
// Returns 0, 1 on failure (message in ref_last_error), -1 when this build has no hook.
int ref_polys_set_pim(ref_polys* p, int on) {
#ifdef FHE_HOOKED
    try {
        std::shared_ptr<PimManager> h;
        if (on) h = std::shared_ptr<PimManager>(PimManager::getPim(4), [](PimManager*) {});
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
        else r->v.push_back(a->v[i].Times(b->v[i]));
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

int ref_max_threads(void) { return omp_get_max_threads(); }

}  // extern "C"
