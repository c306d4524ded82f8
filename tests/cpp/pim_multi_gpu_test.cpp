// pim_multi_gpu_test.cpp — the kept surface (include/pim/*.h) over SEVERAL B200s of one node, driven from C++.
// Usage: pim_multi_gpu_test [gpus]   (default 2).  A C++ caller asks PimManager for k devices ("gpus=k"), exactly as
// the reference's callers ask for `nr_dpus` DPUs (src/core/include/pim/PimManager.h:23-29); vectors (PimData) and
// batches of DCRTPolys (PimDCRTPoly) are then split over the devices the way PimManager::copy_to_pim scatters over
// DPUs (src/core/pim/host/PimManager.cpp:15-24).  Every result is checked against the CPU oracle (test
// infrastructure) bit for bit; the gather runs over NCCL inside libfhe_b200.so (no torch).
#include "pim/PimData.h"
#include "pim/PimDCRTPoly.h"

#include "fhe_oracle.h"

#include <cstdio>
#include <cstdlib>
#include <vector>

static int failures = 0;
#define EXPECT(cond, ...)                                                  \
    do {                                                                   \
        if (!(cond)) { failures++; printf("FAIL %s:%d ", __FILE__, __LINE__); printf(__VA_ARGS__); printf("\n"); } \
    } while (0)

int main(int argc, char** argv) {
    const int want = argc > 1 ? atoi(argv[1]) : 2;
    PimManager* pim = nullptr;
    try {
        pim = PimManager::getPim(8, "gpus=" + std::to_string(want) + ",heap_mb=64");
    } catch (const std::exception& e) {
        printf("cannot take %d devices: %s\n", want, e.what());
        return 2;
    }
    const uint32_t k = pim->getNumDpus();
    EXPECT((int)k == want, "getNumDpus() = %u, asked for %d", k, want);

    // ---- allocate: one (device index, byte offset) pair per device, equal offsets (PimManager.cpp:56-85) ----
    {
        auto a = pim->allocate(1000 * k), b = pim->allocate(16);
        EXPECT(a.size() == k && b.size() == k, "allocate returned %zu / %zu pairs for %u devices", a.size(), b.size(), k);
        for (uint32_t i = 0; i < a.size(); i++) {
            EXPECT(a[i].first == i, "pair %u names device %zu", i, a[i].first);
            EXPECT(a[i].second == a[0].second && b[i].second == b[0].second, "offsets differ between devices");
        }
        EXPECT(b[0].second >= a[0].second + 1000 * 8, "second block overlaps the first");
        pim->deallocate(a);
        pim->deallocate(b);
    }

    // ---- PimData: vectors scattered over the devices, the reference's descriptor route on all of them at once ----
    {
        const uint64_t q = 1152921504606748673ull;
        const size_t n = 4096 * k;
        std::vector<uint64_t> ha(n), hb(n), exp(n);
        orc_fill_uniform(ha.data(), n, q, 31, 0);
        orc_fill_uniform(hb.data(), n, q, 32, 0);
        PimData a(ha), b(hb);
        PimData s = a.ModAdd(b, q), m = a.ModMul(b, q);
        s.sync(); m.sync();
        orc_vec_modadd(exp.data(), ha.data(), hb.data(), n, q);
        EXPECT(s.host() == exp, "PimData ModAdd over %u devices", k);
        orc_vec_modmul(exp.data(), ha.data(), hb.data(), n, q);
        EXPECT(m.host() == exp, "PimData ModMul over %u devices", k);
        a.ModSubEq(b, q);
        a.sync();
        orc_vec_modsub(exp.data(), ha.data(), hb.data(), n, q);
        EXPECT(a.host() == exp, "PimData ModSubEq over %u devices", k);
        bool threw = false;
        try { std::vector<uint64_t> odd(k + 1, 1); PimData o(odd); } catch (const std::exception&) { threw = true; }
        EXPECT(k == 1 || threw, "a length that does not divide over the devices must be refused");
    }

    // ---- PimDCRTPoly: the hot path on a batch split over the devices ----
    for (size_t batch : {(size_t)k * 4, (size_t)k * 3 + 1, (size_t)1}) {
        const uint32_t N = 1u << 14, T = 3;
        auto ring = PimRing::Generate(pim, 60, N, T);
        orc_ctx* oc = orc_ctx_create(N, T, ring->GetModuli().data(), ring->GetRoots().data());
        const size_t words = batch * T * N;
        std::vector<uint64_t> ha(words), hb(words), exp(words);
        for (size_t u = 0; u < batch * T; u++) {
            orc_fill_uniform(&ha[u * N], N, ring->GetModuli()[u % T], 41, u * N);
            orc_fill_uniform(&hb[u * N], N, ring->GetModuli()[u % T], 42, u * N);
        }
        PimDCRTPoly a(ring, batch, PIM_EVALUATION, ha.data()), b(ring, batch, PIM_EVALUATION, hb.data());
        size_t covered = 0;
        for (uint32_t i = 0; i < k; i++) { size_t f, c; a.GetBlock(i, &f, &c); EXPECT(f == covered, "blocks not contiguous"); covered += c; }
        EXPECT(covered == batch, "blocks cover %zu of %zu polynomials", covered, batch);

        PimDCRTPoly c = a + b;
        orc_dcrt_elementwise(oc, 0, exp.data(), ha.data(), hb.data(), batch, 0);
        EXPECT(c.GetValues() == exp, "Plus, batch %zu over %u devices", batch, k);
        PimDCRTPoly d = a * b;
        orc_dcrt_elementwise(oc, 2, exp.data(), ha.data(), hb.data(), batch, 0);
        EXPECT(d.GetValues() == exp, "Times, batch %zu", batch);
        d.SwitchFormat();                                      // EVALUATION -> COEFFICIENT on every device
        orc_dcrt_ntt(oc, 1, exp.data(), batch, 0);
        EXPECT(d.GetFormat() == PIM_COEFFICIENT && d.GetValues() == exp, "SwitchFormat (inverse), batch %zu", batch);
        d.SwitchFormat();
        orc_dcrt_ntt(oc, 0, exp.data(), batch, 0);
        EXPECT(d.GetValues() == exp, "SwitchFormat (forward), batch %zu", batch);
        PimDCRTPoly e = d;                                     // copy: HBM to HBM on every device
        e -= b;
        orc_dcrt_elementwise(oc, 1, exp.data(), exp.data(), hb.data(), batch, 0);
        EXPECT(e.GetValues() == exp, "copy and -=, batch %zu", batch);

        // the optional gather: every device ends up with the whole result (NCCL over NVLink, library streams)
        std::vector<fhe_dptr> all = e.AllGather();
        pim->sync();
        for (uint32_t i = 0; i < k; i++) {
            std::vector<uint64_t> got(words);
            PimManager::check(fhe_d2h(pim->context(i), got.data(), all[i], words * 8, 0), "d2h");
            PimManager::check(fhe_sync(pim->context(i), 0), "sync");
            EXPECT(got == exp, "AllGather: device %u holds the whole batch %zu", i, batch);
            fhe_free(pim->context(i), all[i]);
        }
        if (k > 1) {
            bool threw = false;
            try { (void)e.AutomorphismTransform(5); } catch (const pim_not_implemented_error&) { threw = true; }
            EXPECT(threw, "methods beside the path say that they need one device");
        }
        orc_ctx_destroy(oc);
    }

    PimManager::releasePim();
    if (failures) { printf("%d failure(s)\n", failures); return 1; }
    printf("all ok on %u device(s)\n", k);
    return 0;
}
