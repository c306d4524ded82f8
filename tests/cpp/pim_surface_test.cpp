// pim_surface_test.cpp — C++ parity test of the kept intercept surface (include/pim/*.h) on a B200.
// Calls PimData / PimManager / PimDCRTPoly exactly like code written against the UPMEM fork would
// (src/pke/examples/simple-integers.cpp:120-130 shows the intended use) and checks every result
// against the CPU oracle (oracle/fhe_oracle.h, test infrastructure) — bit exact.
#include "pim/PimData.h"
#include "pim/PimDCRTPoly.h"

#include "fhe_oracle.h"

#include <cstdio>
#include <cstdlib>
#include <random>
#include <vector>

static int failures = 0;
#define EXPECT(cond, ...)                                                  \
    do {                                                                   \
        if (!(cond)) { failures++; printf("FAIL %s:%d ", __FILE__, __LINE__); printf(__VA_ARGS__); printf("\n"); } \
    } while (0)

static std::vector<uint64_t> rnd(size_t n, uint64_t q, uint64_t seed) {
    std::vector<uint64_t> v(n);
    orc_fill_uniform(v.data(), n, q, seed, 0);
    return v;
}

static void test_pimdata_kat() {
    // UnitTestDCRTElements.cpp:285-406 values through the PimData operators, q = 8353
    PimData a{2, 4, 3, 2}, b{2, 1, 2, 0};
    const uint64_t q = 8353;
    PimData s = a.ModAdd(b, q), d = a.ModSub(b, q), m = a.ModMul(b, q);
    s.sync(); d.sync(); m.sync();
    EXPECT((s.host() == std::vector<uint64_t>{4, 5, 5, 2}), "ModAdd");
    EXPECT((d.host() == std::vector<uint64_t>{0, 3, 1, 2}), "ModSub");
    EXPECT((m.host() == std::vector<uint64_t>{4, 4, 6, 0}), "ModMul");
    PimData c = a;                 // shallow, shared block: must not double free (reference bug)
    c.ModAddEq(b, q);
    a.sync();
    EXPECT((a.host() == std::vector<uint64_t>{4, 5, 5, 2}), "shared block after ModAddEq");
    PimData w = a + b;             // default modulus 2^64-1: plain wrapping add
    w.sync();
    EXPECT((w.host() == std::vector<uint64_t>{6, 6, 7, 2}), "operator+ default modulus");
    PimData e;                     // default-constructed handle is destructible
}

static void test_pimdata_random(uint64_t q, size_t n) {
    std::vector<uint64_t> ha = rnd(n, q, 11), hb = rnd(n, q, 12), exp(n);
    PimData a(ha), b(hb);
    const uint64_t scalar = 0x123456789abcdefull;
    struct { int op; void (*vec)(uint64_t*, const uint64_t*, const uint64_t*, size_t, uint64_t);
             void (*sc)(uint64_t*, const uint64_t*, uint64_t, size_t, uint64_t); } ops[3] = {
        {0, orc_vec_modadd, orc_vec_modadd_scalar}, {1, orc_vec_modsub, orc_vec_modsub_scalar},
        {2, orc_vec_modmul, orc_vec_modmul_scalar}};
    for (auto& o : ops) {
        PimData r = o.op == 0 ? a.ModAdd(b, q) : o.op == 1 ? a.ModSub(b, q) : a.ModMul(b, q);
        r.sync();
        o.vec(exp.data(), ha.data(), hb.data(), n, q);
        EXPECT(r.host() == exp, "vector op %d q=%llu n=%zu", o.op, (unsigned long long)q, n);
        PimData rs = o.op == 0 ? a.ModAdd(scalar, q) : o.op == 1 ? a.ModSub(scalar, q) : a.ModMul(scalar, q);
        rs.sync();
        o.sc(exp.data(), ha.data(), scalar, n, q);
        EXPECT(rs.host() == exp, "scalar op %d q=%llu n=%zu", o.op, (unsigned long long)q, n);
        PimData t(ha);             // in-place forms
        if (o.op == 0) t.ModAddEq(b, q); else if (o.op == 1) t.ModSubEq(b, q); else t.ModMulEq(b, q);
        t.sync();
        o.vec(exp.data(), ha.data(), hb.data(), n, q);
        EXPECT(t.host() == exp, "vector_eq op %d", o.op);
    }
}

static void test_pimdata_modops(uint64_t mod, size_t n) {
    // pim/PimData.h:174-233 of the reference: Mod / switchmod / ModByTwo / ModExp / ModInverse /
    // MultiplyAndRound / DivideAndRound, through the mod_ops descriptor and ELEM_MOD_OPS
    std::vector<uint64_t> hx = rnd(n, mod, 21), exp(n);
    for (auto& v : hx) if (v == 0) v = 1;
    hx[0] = mod - 1; hx[1] = mod >> 1; hx[2] = (mod >> 1) + 1;
    auto expect = [&](int kind, uint64_t p, uint64_t q) {
        int rc = orc_vec_modop(kind, exp.data(), hx.data(), n, mod, p, q);
        EXPECT(rc == 0, "oracle modop %d", kind);
    };
    PimData x(hx);
    { PimData r = x.Mod(mod, mod / 3 + 2); r.sync(); expect(MOD_NORMAL, mod / 3 + 2, 0); EXPECT(r.host() == exp, "Mod"); }
    { PimData r = x.ModByTwo(mod); r.sync(); expect(MODBYTWO, 0, 0); EXPECT(r.host() == exp, "ModByTwo"); }
    { PimData r = x.ModExp(mod, 65537); r.sync(); expect(MODEXP, 65537, 0); EXPECT(r.host() == exp, "ModExp"); }
    { PimData r = x.ModInverse(mod); r.sync(); expect(MODINVERSE, 0, 0); EXPECT(r.host() == exp, "ModInverse"); }
    { PimData r = x.MultiplyAndRound(mod, 12345, mod / 5 + 1); r.sync(); expect(MULTIPLYANDROUND, 12345, mod / 5 + 1); EXPECT(r.host() == exp, "MultiplyAndRound"); }
    { PimData r = x.DivideAndRound(mod, 7); r.sync(); expect(DIVIDEANDROUND, 7, 0); EXPECT(r.host() == exp, "DivideAndRound"); }
    { PimData t(hx); t.switchmod(mod, mod + 1000); t.sync(); expect(SWITCHMODULUS, mod + 1000, 0); EXPECT(t.host() == exp, "switchmod"); }
    { PimData t(hx); t.ModEq(mod, 2); t.sync(); expect(MODEQ, 2, 0); EXPECT(t.host() == exp, "ModEq(2)"); }
    { PimData t(hx); t.ModExpEq(mod, mod + 3); t.sync(); expect(MODEXPEQ, mod + 3, 0); EXPECT(t.host() == exp, "ModExpEq"); }
    { PimData t(hx); t.DivideAndRoundEq(mod, mod / 9 + 1); t.sync(); expect(DIVIDEANDROUNDEQ, mod / 9 + 1, 0); EXPECT(t.host() == exp, "DivideAndRoundEq"); }
    std::vector<uint64_t> hz = hx; hz[3] = 0;
    PimData z(hz);
    bool threw = false;
    try { z.ModInverseEq(mod); } catch (const std::runtime_error&) { threw = true; }
    EXPECT(threw, "ModInverse of 0 must throw (math_error in the reference)");
}

static void test_dcrtpoly(uint32_t N, uint32_t T, size_t B) {
    PimManager* pim = PimManager::getPim(4);
    auto ring = PimRing::Generate(pim, 60, N, T);
    std::vector<uint64_t> q(T), psi(T);
    orc_gen_dcrt_params(60, N, T, q.data(), psi.data());
    EXPECT(q == ring->GetModuli() && psi == ring->GetRoots(), "parameter recipe N=%u", N);
    orc_ctx* oc = orc_ctx_create(N, T, q.data(), psi.data());
    size_t words = B * T * (size_t)N;
    std::vector<uint64_t> ha(words), hb(words), exp(words);
    for (size_t b = 0; b < B; b++)
        for (uint32_t t = 0; t < T; t++) {
            orc_fill_uniform(&ha[(b * T + t) * N], N, q[t], 5, (b * T + t) * (size_t)N);
            orc_fill_uniform(&hb[(b * T + t) * N], N, q[t], 6, (b * T + t) * (size_t)N);
        }
    PimDCRTPoly a(ring, B, PIM_EVALUATION, ha.data()), b(ring, B, PIM_EVALUATION, hb.data());
    for (int op = 0; op < 3; op++) {
        PimDCRTPoly c = op == 0 ? a + b : op == 1 ? a - b : a * b;
        orc_dcrt_elementwise(oc, op, exp.data(), ha.data(), hb.data(), B, 0);
        EXPECT(c.GetValues() == exp, "DCRTPoly op %d N=%u", op, N);
        PimDCRTPoly d(a);
        if (op == 0) d += b; else if (op == 1) d -= b; else d *= b;
        EXPECT(d.GetValues() == exp, "DCRTPoly in-place op %d N=%u", op, N);
    }
    PimDCRTPoly x(ring, B, PIM_COEFFICIENT, ha.data());
    bool threw = false;
    try { x.Times(x); } catch (const pim_not_implemented_error&) { threw = true; }
    EXPECT(threw, "Times outside EVALUATION must throw");
    threw = false;
    try { x.Plus(a); } catch (const pim_not_implemented_error&) { threw = true; }
    EXPECT(threw, "Plus with a format mismatch must throw");
    x.SwitchFormat();
    exp = ha;
    orc_dcrt_ntt(oc, 0, exp.data(), B, 0);
    EXPECT(x.GetFormat() == PIM_EVALUATION && x.GetValues() == exp, "forward NTT N=%u", N);
    x.SwitchFormat();
    EXPECT(x.GetFormat() == PIM_COEFFICIENT && x.GetValues() == ha, "round trip N=%u", N);
    PimDCRTPoly y(ring, B, PIM_EVALUATION, hb.data());
    y.SetFormat(PIM_COEFFICIENT);
    exp = hb;
    orc_dcrt_ntt(oc, 1, exp.data(), B, 0);
    EXPECT(y.GetValues() == exp, "inverse NTT N=%u", N);
    orc_ctx_destroy(oc);
}

// SURVEY §8(f) ranks 3 and 4 through the C++ surface
static void test_dcrtpoly_ext(uint32_t N, uint32_t sq, uint32_t sp, size_t B) {
    PimManager* pim = PimManager::getPim(4);
    uint32_t T = sq + sp;
    auto ring = PimRing::Generate(pim, 60, N, T);
    auto ringQ = PimRing::View(ring, 0, sq), ringP = PimRing::View(ring, sq, sp);
    const auto &q = ring->GetModuli(), &psi = ring->GetRoots();
    orc_ctx *oQP = orc_ctx_create(N, T, q.data(), psi.data()), *oQ = orc_ctx_create(N, sq, q.data(), psi.data()),
            *oP = orc_ctx_create(N, sp, q.data() + sq, psi.data() + sq), *oR = orc_ctx_create(N, T - 1, q.data(), psi.data()),
            *oL = orc_ctx_create(N, 1, q.data() + T - 1, psi.data() + T - 1);
    std::vector<uint64_t> hx(B * sq * (size_t)N), hy(B * T * (size_t)N);
    for (size_t u = 0; u < B * sq; u++) orc_fill_uniform(&hx[u * N], N, q[u % sq], 7, u * (size_t)N);
    for (size_t u = 0; u < B * T; u++) orc_fill_uniform(&hy[u * N], N, q[u % T], 8, u * (size_t)N);
    PimDCRTPoly x(ringQ, B, PIM_EVALUATION, hx.data()), y(ring, B, PIM_EVALUATION, hy.data());

    std::vector<uint64_t> exp(hy.size());
    orc_dcrt_automorphism(oQP, exp.data(), hy.data(), B, 5, 0);
    EXPECT(y.AutomorphismTransform(5).GetValues() == exp, "AutomorphismTransform N=%u", N);
    bool threw = false;
    try { y.AutomorphismTransform(6); } catch (const pim_math_error&) { threw = true; }
    EXPECT(threw, "even automorphism index must throw");

    // wire format round trip (bytes against the reference are pinned in tests/test_gpu_serial.py)
    std::string blob = y.Serialize();
    size_t used = 0;
    PimDCRTPoly back = PimDCRTPoly::Deserialize(ring, blob.data(), blob.size(), B, &used);
    EXPECT(used == blob.size() && back.GetFormat() == PIM_EVALUATION && back.GetValues() == hy, "Serialize/Deserialize N=%u", N);
    threw = false;
    try { PimDCRTPoly::Deserialize(ringQ, blob.data(), blob.size(), B); } catch (const std::exception&) { threw = true; }
    EXPECT(threw, "a stream written for another ring must be rejected");

    PimBasis qToP(ringQ, ringP), pToQ(ringP, ringQ);
    orc_approx_mod_up(oQ, oP, exp.data(), hx.data(), B, 0);
    PimDCRTPoly up = x.ApproxModUp(qToP, ring);
    EXPECT(up.GetFormat() == PIM_EVALUATION && up.GetValues() == exp, "ApproxModUp N=%u", N);
    for (uint64_t t : {0ull, 65537ull}) {
        std::vector<uint64_t> e2(B * sq * (size_t)N);
        orc_approx_mod_down(oQ, oP, e2.data(), hy.data(), B, t);
        EXPECT(y.ApproxModDown(pToQ, t).GetValues() == e2, "ApproxModDown t=%llu N=%u", (unsigned long long)t, N);
        std::vector<uint64_t> e3(B * (T - 1) * (size_t)N);
        orc_rescale(oQP, oR, oL, e3.data(), hy.data(), B, 0, t);
        PimRescale plan(ring);
        PimDCRTPoly r = t ? y.ModReduce(plan, t) : y.DropLastElementAndScale(plan);
        EXPECT(r.GetParams()->GetNumTowers() == T - 1 && r.GetValues() == e3, "rescale t=%llu N=%u", (unsigned long long)t, N);
    }
    PimDCRTPoly xc(x);
    xc.SetFormat(PIM_COEFFICIENT);
    std::vector<uint64_t> hi(sq), hm((size_t)sq * sp), e4(B * sp * (size_t)N), coef(hx);
    orc_dcrt_ntt(oQ, 1, coef.data(), B, 0);
    orc_switch_tables(q.data(), sq, q.data() + sq, sp, hi.data(), hm.data());
    orc_approx_switch(e4.data(), coef.data(), B, N, q.data(), sq, q.data() + sq, sp, hi.data(), hm.data());
    EXPECT(xc.ApproxSwitchCRTBasis(qToP).GetValues() == e4, "ApproxSwitchCRTBasis N=%u", N);
    // BFV (HPS) set
    std::vector<double> qinv(sq);
    std::vector<uint64_t> aq((size_t)(sq + 1) * sp), e5(B * sp * (size_t)N);
    orc_exact_tables(q.data(), sq, q.data() + sq, sp, qinv.data(), aq.data());
    orc_switch_crt_basis(e5.data(), coef.data(), B, N, q.data(), sq, q.data() + sq, sp, hi.data(), hm.data(), qinv.data(), aq.data());
    EXPECT(xc.SwitchCRTBasis(qToP).GetValues() == e5, "SwitchCRTBasis N=%u", N);
    orc_expand_crt_basis(oQ, oP, exp.data(), hx.data(), B, 0, 1);
    PimDCRTPoly ex = x.ExpandCRTBasis(qToP, ring, PIM_COEFFICIENT);
    EXPECT(ex.GetFormat() == PIM_COEFFICIENT && ex.GetValues() == exp, "ExpandCRTBasis N=%u", N);
    std::vector<double> frac(sq);
    std::vector<uint64_t> modo((size_t)sp * (sq + 1)), e6(B * sp * (size_t)N);
    for (uint32_t i = 0; i < sq; i++) frac[i] = (double)(i + 1) / (sq + 2.5);
    for (size_t k = 0; k < modo.size(); k++) modo[k] = orc_uniform_at(99, k, q[sq + k / (sq + 1)]);
    orc_scale_and_round(e6.data(), hy.data(), B, N, sq, q.data() + sq, sp, 1, frac.data(), modo.data());
    PimScaleRound sr(ring, sp, true, frac, modo);
    EXPECT(y.ScaleAndRound(sr).GetValues() == e6, "ScaleAndRound N=%u", N);
    // BFV HPSPOVERQ tails: ExpandCRTBasisQlHat (Q as the first towers of QP) and ScaleAndRoundPOverQ (QP minus its last tower)
    std::vector<uint64_t> cst(T), e7(B * T * (size_t)N), e8(B * (T - 1) * (size_t)N);
    for (uint32_t i = 0; i < T; i++) cst[i] = orc_uniform_at(77, i, q[i]);
    orc_expand_ql_hat(e7.data(), hx.data(), B, N, q.data(), T, sq, cst.data());
    PimTowerScale hat(ring, cst);
    PimDCRTPoly eh = x.ExpandCRTBasisQlHat(hat);
    EXPECT(eh.GetParams()->GetNumTowers() == T && eh.GetValues() == e7, "ExpandCRTBasisQlHat N=%u", N);
    orc_scale_round_p_over_q(e8.data(), hy.data(), B, N, q.data(), T - 1, cst.data());
    PimTowerScale pov(PimRing::View(ring, 0, T - 1), std::vector<uint64_t>(cst.begin(), cst.end() - 1));
    EXPECT(y.ScaleAndRoundPOverQ(pov).GetValues() == e8, "ScaleAndRoundPOverQ N=%u", N);
    threw = false;
    try { x.ScaleAndRoundPOverQ(pov); } catch (const std::exception&) { threw = true; }
    EXPECT(threw || sq == T, "ScaleAndRoundPOverQ must reject an input without the extra tower");
    for (orc_ctx* c : {oQP, oQ, oP, oR, oL}) orc_ctx_destroy(c);
}

int main() {
    try {
        test_pimdata_kat();
        test_pimdata_random(73, 1000);
        test_pimdata_random(1152921504606748673ull, 16384);
        test_pimdata_random(8353, 3);
        test_pimdata_modops(8353, 1000);
        test_pimdata_modops(1152921504606748673ull, 4099);
        test_dcrtpoly(1 << 14, 2, 2);      // BASELINE configs[0]
        test_dcrtpoly(1 << 16, 3, 2);
        test_dcrtpoly(1 << 11, 4, 3);
        test_dcrtpoly_ext(1 << 12, 3, 2, 2);
        test_dcrtpoly_ext(1 << 14, 5, 3, 1);
    } catch (const std::exception& e) {
        printf("FAIL exception: %s\n", e.what());
        failures++;
    }
    PimManager::releasePim();
    printf(failures ? "pim_surface_test: %d FAILURE(S)\n" : "pim_surface_test: all ok\n", failures);
    return failures ? 1 : 0;
}
