// pke_driver.cpp — the reference's OWN scheme layer (src/pke: BFV / BGV / CKKS) run on top of the hooked DCRTPoly
// (TEST INFRASTRUCTURE ONLY).  Compiled by `make -C oracle pke` with the reference's sources where they lie and
// -DFIXED_SEED (math/distributiongenerator.h:51), so the PRNG stream is the same in every run on one thread.
//
// Every scenario prints one line per ciphertext / plaintext: a name and a digest of every tower word.  Run once with
// FHE_B200_ATTACH=1 (every DCRTPoly the schemes create carries the backend handle: their +, *, SwitchFormat, automorphisms,
// basis conversions, rescaling and the BEHZ set run on the GPU, pim/PimHook.h) and once without (the reference's CPU
// statements): the two outputs must be identical, which tests/test_pke_under_hook.py checks.  Nothing here restates
// arithmetic.
#include "openfhe.h"

#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <string>
#include <vector>

#include "fhe_b200.h"
#include <omp.h>

using namespace lbcrypto;

static uint64_t digest_poly(const DCRTPoly& poly, uint64_t h) {
    for (const auto& tower : poly.GetAllElements()) {        // reads the host towers: a device copy is downloaded first
        const auto& v = tower.GetValues();
        for (usint i = 0; i < v.GetLength(); i++) {
            h ^= v[i].ConvertToInt();
            h *= 0x100000001B3ull;                           // FNV-1a over 64-bit words
        }
        h ^= tower.GetModulus().ConvertToInt();
        h *= 0x100000001B3ull;
    }
    h ^= (uint64_t)poly.GetFormat() + 17;
    return h * 0x100000001B3ull;
}
static void show(const std::string& name, ConstCiphertext<DCRTPoly> ct) {
    uint64_t h = 0xcbf29ce484222325ull;
    for (const auto& e : ct->GetElements()) h = digest_poly(e, h);
    std::printf("%s ct elements=%zu towers=%zu digest=%016llx\n", name.c_str(), ct->GetElements().size(),
                (size_t)ct->GetElements()[0].GetNumOfElements(), (unsigned long long)h);
}
static void show(const std::string& name, const std::vector<int64_t>& v) {
    std::printf("%s pt", name.c_str());
    for (auto x : v) std::printf(" %lld", (long long)x);
    std::printf("\n");
}
static void show(const std::string& name, const std::vector<double>& v) {
    std::printf("%s pt", name.c_str());
    for (auto x : v) std::printf(" %.6f", x);
    std::printf("\n");
}

struct Clock {
    std::chrono::steady_clock::time_point t0 = std::chrono::steady_clock::now();
    double ms() const { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count(); }
};

static const std::vector<int64_t> V1 = {1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 12};
static const std::vector<int64_t> V2 = {3, 2, 1, 4, 5, 6, 7, 8, 9, 10, 11, 12};

// BFVrns: KeyGen, Encrypt, EvalAdd, EvalMult + relinearisation (twice: depth 2), EvalRotate, Decrypt
static void run_bfv(const char* label, MultiplicationTechnique tech, KeySwitchTechnique ks, uint32_t ringDim) {
    Clock clk;
    CCParams<CryptoContextBFVRNS> P;
    P.SetPlaintextModulus(65537);
    P.SetMultiplicativeDepth(2);
    P.SetMultiplicationTechnique(tech);
    P.SetKeySwitchTechnique(ks);
    P.SetSecurityLevel(HEStd_NotSet);
    P.SetRingDim(ringDim);
    CryptoContext<DCRTPoly> cc = GenCryptoContext(P);
    cc->Enable(PKE);
    cc->Enable(KEYSWITCH);
    cc->Enable(LEVELEDSHE);
    auto kp = cc->KeyGen();
    cc->EvalMultKeyGen(kp.secretKey);
    cc->EvalRotateKeyGen(kp.secretKey, {1, -2});
    auto c1 = cc->Encrypt(kp.publicKey, cc->MakePackedPlaintext(V1));
    auto c2 = cc->Encrypt(kp.publicKey, cc->MakePackedPlaintext(V2));
    const std::string L(label);
    show(L + ".c1", c1);
    auto cadd = cc->EvalAdd(c1, c2);
    show(L + ".add", cadd);
    auto csub = cc->EvalSub(c1, c2);
    show(L + ".sub", csub);
    auto cmul = cc->EvalMult(c1, c2);
    show(L + ".mul", cmul);
    auto cmul2 = cc->EvalMult(cmul, cadd);
    show(L + ".mul2", cmul2);
    auto crot = cc->EvalRotate(cmul, 1);
    show(L + ".rot1", crot);
    auto crot2 = cc->EvalRotate(cadd, -2);
    show(L + ".rot-2", crot2);
    std::vector<std::pair<std::string, Ciphertext<DCRTPoly>>> results{
        {"add", cadd}, {"sub", csub}, {"mul", cmul}, {"mul2", cmul2}, {"rot1", crot}, {"rot-2", crot2}};
    if (tech != HPSPOVERQLEVELED && ks == HYBRID) {     // (the reference's own hoisted rotation throws for HPSPOVERQLEVELED on the CPU path too)
        auto pre = cc->EvalFastRotationPrecompute(cmul);
        auto cfr1 = cc->EvalFastRotation(cmul, 1, 2 * ringDim, pre);
        show(L + ".fastrot1", cfr1);
        results.push_back({"fastrot1", cfr1});
    }
    for (auto& item : results) {
        Plaintext r;
        cc->Decrypt(kp.secretKey, item.second, &r);
        r->SetLength(V1.size());
        show(L + ".dec." + item.first, r->GetPackedValue());
    }
    std::fprintf(stderr, "%s: %.1f ms, launches so far %ld\n", label, clk.ms(), fhe_launch_count());
}

// BGVrns: EvalMult, ModReduce (explicit with FIXEDMANUAL), EvalRotate
static void run_bgv(const char* label, ScalingTechnique st, uint32_t ringDim) {
    Clock clk;
    CCParams<CryptoContextBGVRNS> P;
    P.SetPlaintextModulus(65537);
    P.SetMultiplicativeDepth(2);
    P.SetScalingTechnique(st);
    P.SetSecurityLevel(HEStd_NotSet);
    P.SetRingDim(ringDim);
    CryptoContext<DCRTPoly> cc = GenCryptoContext(P);
    cc->Enable(PKE);
    cc->Enable(KEYSWITCH);
    cc->Enable(LEVELEDSHE);
    auto kp = cc->KeyGen();
    cc->EvalMultKeyGen(kp.secretKey);
    cc->EvalRotateKeyGen(kp.secretKey, {1});
    auto c1 = cc->Encrypt(kp.publicKey, cc->MakePackedPlaintext(V1));
    auto c2 = cc->Encrypt(kp.publicKey, cc->MakePackedPlaintext(V2));
    const std::string L(label);
    auto cadd = cc->EvalAdd(c1, c2);
    show(L + ".add", cadd);
    auto cmul = cc->EvalMult(c1, c2);
    show(L + ".mul", cmul);
    if (st == FIXEDMANUAL) {
        cmul = cc->ModReduce(cmul);
        show(L + ".modreduce", cmul);
    }
    auto cmul2 = cc->EvalMult(cmul, cmul);
    show(L + ".mul2", cmul2);
    auto crot = cc->EvalRotate(cmul, 1);
    show(L + ".rot1", crot);
    for (auto& item : std::vector<std::pair<std::string, Ciphertext<DCRTPoly>>>{{"add", cadd}, {"mul", cmul}, {"mul2", cmul2}, {"rot1", crot}}) {
        Plaintext r;
        cc->Decrypt(kp.secretKey, item.second, &r);
        r->SetLength(V1.size());
        show(L + ".dec." + item.first, r->GetPackedValue());
    }
    std::fprintf(stderr, "%s: %.1f ms, launches so far %ld\n", label, clk.ms(), fhe_launch_count());
}

// CKKSrns: EvalMult, Rescale, EvalAdd, plaintext multiplication, EvalRotate (hybrid key switching: ApproxModUp / ApproxModDown)
static void run_ckks(const char* label, ScalingTechnique st, KeySwitchTechnique ks, uint32_t ringDim) {
    Clock clk;
    CCParams<CryptoContextCKKSRNS> P;
    P.SetMultiplicativeDepth(3);
    P.SetScalingModSize(50);
    P.SetBatchSize(8);
    P.SetScalingTechnique(st);
    P.SetKeySwitchTechnique(ks);
    if (ks == BV) P.SetDigitSize(10);
    P.SetSecurityLevel(HEStd_NotSet);
    P.SetRingDim(ringDim);
    CryptoContext<DCRTPoly> cc = GenCryptoContext(P);
    cc->Enable(PKE);
    cc->Enable(KEYSWITCH);
    cc->Enable(LEVELEDSHE);
    auto kp = cc->KeyGen();
    cc->EvalMultKeyGen(kp.secretKey);
    cc->EvalRotateKeyGen(kp.secretKey, {1, -2});
    const std::vector<double> x1 = {0.25, 0.5, 0.75, 1.0, 2.0, 3.0, 4.0, 5.0}, x2 = {5.0, 4.0, 3.0, 2.0, 1.0, 0.75, 0.5, 0.25};
    auto p1 = cc->MakeCKKSPackedPlaintext(x1), p2 = cc->MakeCKKSPackedPlaintext(x2);
    auto c1 = cc->Encrypt(kp.publicKey, p1), c2 = cc->Encrypt(kp.publicKey, p2);
    const std::string L(label);
    auto cadd = cc->EvalAdd(c1, c2);
    show(L + ".add", cadd);
    auto cmul = cc->EvalMult(c1, c2);
    show(L + ".mul", cmul);
    if (st == FIXEDMANUAL) {
        cmul = cc->Rescale(cmul);
        show(L + ".rescale", cmul);
    }
    auto cpt = cc->EvalMult(cmul, p1);
    if (st == FIXEDMANUAL) cpt = cc->Rescale(cpt);
    show(L + ".mulpt", cpt);
    auto csq = cc->EvalMult(cmul, cmul);
    if (st == FIXEDMANUAL) csq = cc->Rescale(csq);
    show(L + ".square", csq);
    auto crot = cc->EvalRotate(cmul, 1);
    show(L + ".rot1", crot);
    auto crot2 = cc->EvalRotate(c1, -2);
    show(L + ".rot-2", crot2);
    auto pre = cc->EvalFastRotationPrecompute(cmul);                       // hoisted rotations: the digits once, two rotations
    auto cfr1 = cc->EvalFastRotation(cmul, 1, 2 * ringDim, pre);
    auto cfr2 = cc->EvalFastRotation(cmul, -2, 2 * ringDim, pre);
    show(L + ".fastrot1", cfr1);
    show(L + ".fastrot-2", cfr2);
    for (auto& item : std::vector<std::pair<std::string, Ciphertext<DCRTPoly>>>{
             {"add", cadd}, {"mul", cmul}, {"mulpt", cpt}, {"square", csq}, {"rot1", crot}, {"rot-2", crot2}, {"fastrot1", cfr1}, {"fastrot-2", cfr2}}) {
        Plaintext r;
        cc->Decrypt(kp.secretKey, item.second, &r);
        r->SetLength(8);
        show(L + ".dec." + item.first, r->GetRealPackedValue());
    }
    std::fprintf(stderr, "%s: %.1f ms, launches so far %ld\n", label, clk.ms(), fhe_launch_count());
}

// Throughput of the schemes' own operations over `count` independent ciphertexts, the callers' OpenMP loop over
// ciphertexts outside (the reference's inner loops over towers then run serially, utils/parallel.h): usage
//   pke-driver-hooked bench <bfv-hps|bfv-behz|bfv-hpspoverq|ckks> <ringDim> <depth> <threads> <count>
// Prints one JSON line per operation; run with and without FHE_B200_ATTACH=1.
static void device_wait() {
    const char* att = std::getenv("FHE_B200_ATTACH");
    if (att && att[0] == '1') pimhook::default_handle()->sync();
}
template <class F>
static double time_loop(int threads, int count, int reps, F f) {
    double best = 1e30;
    for (int r = 0; r < reps; r++) {
        Clock c;
#pragma omp parallel for num_threads(threads) schedule(dynamic, 1)
        for (int i = 0; i < count; i++) f(i);
        device_wait();                                   // the attached path is asynchronous: the work counts when it has finished
        best = std::min(best, c.ms());
    }
    return best;
}
static int run_bench(const std::string& scheme, uint32_t n, uint32_t depth, int threads, int count) {
    CryptoContext<DCRTPoly> cc;
    const bool ckks = scheme == "ckks";
    if (ckks) {
        CCParams<CryptoContextCKKSRNS> P;
        P.SetMultiplicativeDepth(depth);
        P.SetScalingModSize(50);
        P.SetScalingTechnique(FIXEDMANUAL);
        P.SetSecurityLevel(HEStd_NotSet);
        P.SetRingDim(n);
        cc = GenCryptoContext(P);
    } else {
        CCParams<CryptoContextBFVRNS> P;
        P.SetPlaintextModulus(65537);
        P.SetMultiplicativeDepth(depth);
        P.SetMultiplicationTechnique(scheme == "bfv-behz" ? BEHZ : scheme == "bfv-hpspoverq" ? HPSPOVERQ : HPS);
        P.SetKeySwitchTechnique(scheme == "bfv-bv" ? BV : HYBRID);      // bfv-bv: HPS with the scheme's default key switching
        P.SetSecurityLevel(HEStd_NotSet);
        P.SetRingDim(n);
        cc = GenCryptoContext(P);
    }
    cc->Enable(PKE);
    cc->Enable(KEYSWITCH);
    cc->Enable(LEVELEDSHE);
    auto kp = cc->KeyGen();
    cc->EvalMultKeyGen(kp.secretKey);
    cc->EvalRotateKeyGen(kp.secretKey, {1});
    std::vector<Ciphertext<DCRTPoly>> a(count), b(count), out(count);
    for (int i = 0; i < count; i++) {
        Plaintext pa, pb;
        if (ckks) {
            pa = cc->MakeCKKSPackedPlaintext(std::vector<double>{0.5 + i, 1.5, 2.5});
            pb = cc->MakeCKKSPackedPlaintext(std::vector<double>{1.0, 0.25 * i, 3.0});
        } else {
            pa = cc->MakePackedPlaintext(std::vector<int64_t>{1 + i, 2, 3});
            pb = cc->MakePackedPlaintext(std::vector<int64_t>{4, 5 + i, 6});
        }
        a[i] = cc->Encrypt(kp.publicKey, pa);
        b[i] = cc->Encrypt(kp.publicKey, pb);
    }
    const size_t towers = a[0]->GetElements()[0].GetNumOfElements();
    auto report = [&](const char* op, double ms) {
        std::printf("{\"scheme\": \"%s\", \"op\": \"%s\", \"ring_dim\": %u, \"towers\": %zu, \"threads\": %d, \"count\": %d, \"ms\": %.3f, "
                    "\"ops_per_s\": %.2f, \"attached\": %s, \"launches\": %ld}\n", scheme.c_str(), op, n, towers, threads, count, ms,
                    count / (ms * 1e-3), std::getenv("FHE_B200_ATTACH") && std::getenv("FHE_B200_ATTACH")[0] == '1' ? "true" : "false",
                    fhe_launch_count());
        std::fflush(stdout);
    };
    time_loop(threads, count, 1, [&](int i) { out[i] = cc->EvalMult(a[i], b[i]); });          // warm-up: tables, plans, uploads
    report("EvalAdd", time_loop(threads, count, 3, [&](int i) { out[i] = cc->EvalAdd(a[i], b[i]); }));
    report("EvalMult+relin", time_loop(threads, count, 3, [&](int i) { out[i] = cc->EvalMult(a[i], b[i]); }));
    if (ckks)
        report("EvalMult+relin+Rescale", time_loop(threads, count, 3, [&](int i) { out[i] = cc->Rescale(cc->EvalMult(a[i], b[i])); }));
    report("EvalRotate", time_loop(threads, count, 3, [&](int i) { out[i] = cc->EvalRotate(a[i], 1); }));
    if (scheme != "bfv-bv") {
        cc->EvalRotateKeyGen(kp.secretKey, {2, 3, 4});
        report("4 hoisted rotations (EvalFastRotationPrecompute + 4 x EvalFastRotation)", time_loop(threads, count, 3, [&](int i) {
                   auto pre = cc->EvalFastRotationPrecompute(a[i]);
                   for (int r = 1; r <= 4; r++) out[i] = cc->EvalFastRotation(a[i], r, 2 * n, pre);
               }));
    }
    uint64_t h = 0;                                      // read one result back: the digest must not depend on the path
    for (const auto& e : out[count - 1]->GetElements()) h = digest_poly(e, h);
    std::fprintf(stderr, "last result digest %016llx\n", (unsigned long long)h);
    return 0;
}

// CKKS bootstrapping with the parameters of the reference's src/pke/examples/simple-ckks-bootstrapping.cpp at a chosen ring
// dimension and slot count: times EvalBootstrap alone.  usage: pke-driver-hooked bootstrap <ringDim> <slots> <reps>
static int run_bootstrap(uint32_t n, uint32_t slots, int reps) {
    CCParams<CryptoContextCKKSRNS> P;
    SecretKeyDist skd = UNIFORM_TERNARY;
    P.SetSecretKeyDist(skd);
    P.SetSecurityLevel(HEStd_NotSet);
    P.SetRingDim(n);
    P.SetScalingModSize(59);
    P.SetScalingTechnique(FLEXIBLEAUTO);
    P.SetFirstModSize(60);
    std::vector<uint32_t> levelBudget = {4, 4};
    usint depth = 10 + FHECKKSRNS::GetBootstrapDepth(levelBudget, skd);
    P.SetMultiplicativeDepth(depth);
    CryptoContext<DCRTPoly> cc = GenCryptoContext(P);
    cc->Enable(PKE);
    cc->Enable(KEYSWITCH);
    cc->Enable(LEVELEDSHE);
    cc->Enable(ADVANCEDSHE);
    cc->Enable(FHE);
    Clock setup;
    cc->EvalBootstrapSetup(levelBudget, {0, 0}, slots);
    auto kp = cc->KeyGen();
    cc->EvalMultKeyGen(kp.secretKey);
    cc->EvalBootstrapKeyGen(kp.secretKey, slots);
    const double setup_ms = setup.ms();
    std::vector<double> x = {0.25, 0.5, 0.75, 1.0, 2.0, 3.0, 4.0, 5.0};
    Plaintext pt = cc->MakeCKKSPackedPlaintext(x, 1, depth - 1, nullptr, slots);
    pt->SetLength(x.size());
    auto ct = cc->Encrypt(kp.publicKey, pt);
    Ciphertext<DCRTPoly> after;
    double best = 1e30;
    for (int r = 0; r < reps; r++) {
        const long up0 = pimhook::resident_uploads(), down0 = pimhook::resident_downloads(), l0 = fhe_launch_count();
        Clock c;
        after = cc->EvalBootstrap(ct);
        std::fprintf(stderr, "bootstrap rep %d: %ld polynomials uploaded, %ld downloaded, %ld launches\n", r,
                     pimhook::resident_uploads() - up0, pimhook::resident_downloads() - down0, fhe_launch_count() - l0);
        uint64_t h = 0;                                  // reading the result ends the measurement (the device work is asynchronous)
        for (const auto& e : after->GetElements()) h = digest_poly(e, h);
        best = std::min(best, c.ms());
        std::fprintf(stderr, "bootstrap rep %d: digest %016llx\n", r, (unsigned long long)h);
    }
    Plaintext res;
    cc->Decrypt(kp.secretKey, after, &res);
    res->SetLength(x.size());
    show("bootstrap.dec", res->GetRealPackedValue());
    const char* att = std::getenv("FHE_B200_ATTACH");
    std::printf("{\"scheme\": \"ckks\", \"op\": \"EvalBootstrap\", \"ring_dim\": %u, \"slots\": %u, \"towers\": %zu, \"threads\": %d, \"ms\": %.1f, "
                "\"setup_keygen_ms\": %.0f, \"levels_after\": %u, \"attached\": %s, \"launches\": %ld}\n", n, slots,
                (size_t)cc->GetCryptoParameters()->GetElementParams()->GetParams().size(), omp_get_max_threads(), best, setup_ms,
                (unsigned)(depth - after->GetLevel()), att && att[0] == '1' ? "true" : "false", fhe_launch_count());
    return 0;
}

int main(int argc, char** argv) {
    if (argc > 1 && std::string(argv[1]) == "bootstrap") {
        if (argc < 5) { std::fprintf(stderr, "usage: %s bootstrap <ringDim> <slots> <reps>\n", argv[0]); return 2; }
        return run_bootstrap((uint32_t)std::atoi(argv[2]), (uint32_t)std::atoi(argv[3]), std::atoi(argv[4]));
    }
    if (argc > 1 && std::string(argv[1]) == "bench") {
        if (argc < 7) { std::fprintf(stderr, "usage: %s bench <scheme> <ringDim> <depth> <threads> <count>\n", argv[0]); return 2; }
        return run_bench(argv[2], (uint32_t)std::atoi(argv[3]), (uint32_t)std::atoi(argv[4]), std::atoi(argv[5]), std::atoi(argv[6]));
    }
    const uint32_t n = argc > 1 ? (uint32_t)std::atoi(argv[1]) : 4096;
    const std::string only = argc > 2 ? argv[2] : "";
    auto want = [&](const char* s) { return only.empty() || only == s; };
    if (want("bfv-hps")) run_bfv("bfv-hps", HPS, HYBRID, n);
    if (want("bfv-hpspoverq")) run_bfv("bfv-hpspoverq", HPSPOVERQ, HYBRID, n);
    if (want("bfv-hpspoverqleveled")) run_bfv("bfv-hpspoverqleveled", HPSPOVERQLEVELED, HYBRID, n);
    if (want("bfv-behz")) run_bfv("bfv-behz", BEHZ, BV, n);
    if (want("bgv-manual")) run_bgv("bgv-manual", FIXEDMANUAL, n);
    if (want("bgv-auto")) run_bgv("bgv-auto", FLEXIBLEAUTO, n);
    if (want("ckks-manual")) run_ckks("ckks-manual", FIXEDMANUAL, HYBRID, n);
    if (want("ckks-auto")) run_ckks("ckks-auto", FLEXIBLEAUTO, HYBRID, n);
    if (want("ckks-bv")) run_ckks("ckks-bv", FIXEDMANUAL, BV, n);
    std::fprintf(stderr, "total kernel launches %ld\n", fhe_launch_count());
    return 0;
}
