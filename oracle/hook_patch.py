#!/usr/bin/env python3
"""Prepares the header tree for the LIVE-INTERCEPT build of the reference (TEST INFRASTRUCTURE).

    python3 oracle/hook_patch.py /root/reference /tmp/fhe_b200_hook_include

Copies the reference's src/core/include to a temporary directory OUTSIDE this repository (reference
sources are never copied into the repo) and applies the few-line patch INTEGRATION.md §4 describes:

  1. drop the fork's pim/*.h, so that `#include "pim/PimManager.h"` / "pim/PimData.h" resolve to this
     repository's include/pim/*.h (the B200 implementation of the same surface);
  2. remove the `PimData pim_m_data = {1,2,3,4};` member every NativeVector carries (mubintvecnat.h:128),
     which would allocate device memory per vector;
  3. insert the intercept calls (pim/PimHook.h) at the top of DCRTPolyImpl::operator+= / -= / *= and
     SwitchFormat (dcrtpoly-impl.h:410-416, 438-444, 2518-2524; dcrtpoly.h:142-148);
  4. resident form: rename DCRTPolyImpl's tower member to m_vectors_ (dcrtpoly.h:421), add a device copy handle
     next to it, and route every other use of `m_vectors` in dcrtpoly.h / dcrtpoly-impl.h through the accessor
     PimHostVectors() (a #define scoped to those two headers), which refreshes the host towers from the device
     copy when that is newer.  Hooked operators then leave their result in HBM.

Only the inserted lines are ours; anchors are matched on the reference's function signatures.
"""
import os
import shutil
import sys

ref, out = sys.argv[1], sys.argv[2]
src = os.path.join(ref, "src", "core", "include")
if os.path.isdir(out):
    shutil.rmtree(out)
shutil.copytree(src, out)
shutil.rmtree(os.path.join(out, "pim"))


def patch(rel, edits):
    path = os.path.join(out, rel)
    text = open(path).read()
    for anchor, insert, mode in edits:
        if text.count(anchor) != 1:
            raise SystemExit(f"hook_patch: anchor not found exactly once in {rel}: {anchor!r}")
        if mode == "after":
            text = text.replace(anchor, anchor + insert)
        elif mode == "delete_line":
            start = text.rfind("\n", 0, text.index(anchor)) + 1
            end = text.index("\n", text.index(anchor)) + 1
            text = text[:start] + text[end:]
    open(path, "w").write(text)


patch("math/hal/intnat/mubintvecnat.h", [("PimData pim_m_data", "", "delete_line")])

# ---- resident form: the member is renamed, every other use of `m_vectors` goes through an accessor ----------
ACCESSORS = """
    /* B200 intercept, resident form (pim/PimHook.h): host view of the towers.  const: refresh from the device copy
     * if that is newer.  non-const: refresh, then drop the device copy (the caller may write). */
    const std::vector<PolyType>& PimHostVectors() const {
        if (pimres_) {       /* one predictable branch: the reference reads m_vectors[i][j] in inner loops */
            auto r = std::atomic_load(&pimres_);      /* may be entered from the threads of an OpenMP loop over towers at once */
            if (r) pimhook::sync_host(m_vectors_, r, m_format);
        }
        return m_vectors_;
    }
    std::vector<PolyType>& PimHostVectors() {
        if (pimres_) {
            auto r = std::atomic_load(&pimres_);
            if (r) {
                pimhook::sync_host(m_vectors_, r, m_format);      /* serialised by the copy's own mutex; the second thread finds it done */
                std::atomic_store(&pimres_, std::shared_ptr<pimhook::Resident>());
            }
        }
        return m_vectors_;
    }
    /* copy construction / assignment: an attached source is copied HBM to HBM (uploaded first if it never was) */
    void PimCopyFrom(const DCRTPolyImpl& e) {
        if (&e == this) return;
        std::shared_ptr<pimhook::Resident> r;
        if (e.pim && pimhook::copy_resident(e.pim.get(), m_vectors_, r, e.m_vectors_, e.pimres_)) {
            pimres_ = r;
            return;
        }
        m_vectors_ = e.PimHostVectors();
        pimres_.reset();
    }
    mutable std::vector<PolyType> m_vectors_;
    mutable std::shared_ptr<pimhook::Resident> pimres_;

public:
    /* the tower count needs no host values: GetNumOfElements() must not force a download */
    size_t PimNumTowers() const { return m_vectors_.size(); }

protected:
"""
HOOK_EQ = ("\n    if (pim && pimhook::resident_eq(pim.get(), m_vectors_, pimres_, rhs.m_vectors_, rhs.pimres_, %s))"
           "\n        return *this;")


HOOK_BIN = ("        if (pim) {\n"
            "            DCRTPolyType dev(m_params, m_format, false);\n"
            "            if (pimhook::resident_binary(pim.get(), m_vectors_, pimres_, rhs.m_vectors_, rhs.pimres_, dev.m_vectors_,\n"
            "                                         dev.pimres_, %s)) {\n"
            "                dev.pim = pim;\n"
            "                return dev;\n"
            "            }\n"
            "        }\n")


def rewrite(rel, fn):
    path = os.path.join(out, rel)
    text = fn(open(path).read())
    open(path, "w").write(text)


def once(text, old, new, rel):
    if text.count(old) != 1:
        raise SystemExit(f"hook_patch: anchor not found exactly once in {rel}: {old!r}")
    return text.replace(old, new)


def patch_class(text):
    rel = "dcrtpoly.h"
    text = once(text, '#include "utils/parallel.h"\n', '#include "utils/parallel.h"\n#include "pim/PimHook.h"\n', rel)
    text = once(text, "    std::vector<PolyType> m_vectors;\n", ACCESSORS, rel)
    # copies of a resident poly copy the device buffer HBM to HBM instead of downloading it; the handle travels along
    text = once(text, "m_vectors{e.m_vectors} {}",
                "m_vectors_{}, pim{e.pim} { PimCopyFrom(e); }", rel)
    text = once(text, "        m_format  = rhs.m_format;\n        m_vectors = rhs.m_vectors;\n        return *this;\n",
                "        m_format  = rhs.m_format;\n        PimCopyFrom(rhs);\n"
                "        if (rhs.pim) pim = rhs.pim;\n        return *this;\n", rel)
    # ciphertext level: KeySwitchHYBRID::KeySwitchCore as one device composition (pim/PimHook.h resident_keyswitch_hybrid);
    # a static member, because it reaches into the key's polynomials
    text = once(text, "    std::shared_ptr<PimManager> GetPim() {\n        return pim;\n    }\n",
                "    std::shared_ptr<PimManager> GetPim() {\n        return pim;\n    }\n\n"
                "    static bool PimKeySwitchHybrid(const DCRTPolyImpl& c, const std::vector<DCRTPolyImpl>& bv,\n"
                "                                   const std::vector<DCRTPolyImpl>& av, size_t sizeP, size_t alpha, uint64_t t,\n"
                "                                   const void* keyId, DCRTPolyImpl& ct0, DCRTPolyImpl& ct1) {\n"
                "        if (!c.pim || c.m_format != Format::EVALUATION || bv.empty() || bv.size() != av.size()) return false;\n"
                "        for (size_t j = 0; j < bv.size(); j++)\n"
                "            if (bv[j].m_format != Format::EVALUATION || av[j].m_format != Format::EVALUATION) return false;\n"
                "        using Towers = std::vector<PolyType>;\n"
                "        auto keyAt = [&](int which, size_t j) {\n"
                "            const DCRTPolyImpl& k = which ? av[j] : bv[j];\n"
                "            return pimhook::KeyRef<Towers>{k.m_vectors_, k.pimres_};\n"
                "        };\n"
                "        std::shared_ptr<pimhook::Resident> r0, r1;\n"
                "        if (!pimhook::resident_keyswitch_hybrid(c.pim.get(), c.m_vectors_, c.pimres_, bv.size(), keyAt, sizeP, alpha, t,\n"
                "                                                keyId, r0, r1))\n"
                "            return false;\n"
                "        ct0         = DCRTPolyImpl(c.m_params, Format::EVALUATION, false);\n"
                "        ct1         = DCRTPolyImpl(c.m_params, Format::EVALUATION, false);\n"
                "        ct0.pimres_ = r0;\n"
                "        ct1.pimres_ = r1;\n"
                "        ct0.pim = ct1.pim = c.pim;\n"
                "        return true;\n"
                "    }\n"
                "    /* cv0 += ks0, cv1 += ks1 of the hybrid key switch of c: the relinearisation step of the schemes' EvalMult\n"
                "     * (base-leveledshe.cpp:210-213 and the six places like it) as one device composition */\n"
                "    static bool PimKeySwitchHybridAdd(const DCRTPolyImpl& c, const std::vector<DCRTPolyImpl>& bv,\n"
                "                                      const std::vector<DCRTPolyImpl>& av, size_t sizeP, size_t alpha, uint64_t t,\n"
                "                                      const void* keyId, DCRTPolyImpl& cv0, DCRTPolyImpl& cv1) {\n"
                "        if (!c.pim || c.m_format != Format::EVALUATION || cv0.m_format != Format::EVALUATION ||\n"
                "            cv1.m_format != Format::EVALUATION || bv.empty() || bv.size() != av.size())\n"
                "            return false;\n"
                "        for (size_t j = 0; j < bv.size(); j++)\n"
                "            if (bv[j].m_format != Format::EVALUATION || av[j].m_format != Format::EVALUATION) return false;\n"
                "        using Towers = std::vector<PolyType>;\n"
                "        auto keyAt = [&](int which, size_t j) {\n"
                "            const DCRTPolyImpl& k = which ? av[j] : bv[j];\n"
                "            return pimhook::KeyRef<Towers>{k.m_vectors_, k.pimres_};\n"
                "        };\n"
                "        std::shared_ptr<pimhook::Resident> r0, r1;\n"
                "        return pimhook::resident_keyswitch_hybrid(c.pim.get(), c.m_vectors_, c.pimres_, bv.size(), keyAt, sizeP, alpha, t,\n"
                "                                                  keyId, r0, r1, &cv0.m_vectors_, &cv0.pimres_, &cv1.m_vectors_, &cv1.pimres_);\n"
                "    }\n"
                "    /* LeveledSHEBase::EvalMultCore, 2 x 2 (base-leveledshe.cpp:667-672): the three products of two ciphertexts' elements\n"
                "     * as one launch on the device copies */\n"
                "    static bool PimTensorProduct(const DCRTPolyImpl& a0, const DCRTPolyImpl& a1, const DCRTPolyImpl& b0, const DCRTPolyImpl& b1,\n"
                "                                 std::vector<DCRTPolyImpl>& out) {\n"
                "        if (!a0.pim || a0.m_format != Format::EVALUATION || a1.m_format != Format::EVALUATION ||\n"
                "            b0.m_format != Format::EVALUATION || b1.m_format != Format::EVALUATION || out.size() != 3)\n"
                "            return false;\n"
                "        std::shared_ptr<pimhook::Resident> r[3];\n"
                "        if (!pimhook::resident_tensor_product(a0.pim.get(), a0.m_vectors_, a0.pimres_, a1.m_vectors_, a1.pimres_, b0.m_vectors_,\n"
                "                                              b0.pimres_, b1.m_vectors_, b1.pimres_, r))\n"
                "            return false;\n"
                "        for (int i = 0; i < 3; i++) {\n"
                "            out[i]         = DCRTPolyImpl(a0.m_params, Format::EVALUATION, false);\n"
                "            out[i].pimres_ = r[i];\n"
                "            out[i].pim     = a0.pim;\n"
                "        }\n"
                "        return true;\n"
                "    }\n"
                "    /* KeySwitchExt / EvalFastRotationExt (keyswitch-hybrid.cpp:231-256, ckksrns-leveledshe.cpp:402-460): c over Ql times\n"
                "     * per-tower constants, extended by zero towers to paramsExt (Ql u P) - instead of TimesNoCheck + SetElementAtIndex */\n"
                "    static bool PimExtendTimes(const DCRTPolyImpl& c, const std::vector<NativeInteger>& consts,\n"
                "                               const std::shared_ptr<Params>& paramsExt, DCRTPolyImpl& out) {\n"
                "        if (!c.pim || c.m_format != Format::EVALUATION || c.m_vectors_.empty() || consts.size() < c.m_vectors_.size() ||\n"
                "            paramsExt->GetParams().size() <= c.m_vectors_.size())\n"
                "            return false;\n"
                "        DCRTPolyImpl dev(paramsExt, Format::EVALUATION, false);\n"
                "        if (!pimhook::resident_tower_scale(c.pim.get(), 0, c.m_vectors_, c.pimres_, dev.m_vectors_, dev.pimres_, consts)) return false;\n"
                "        dev.pim = c.pim;\n"
                "        out     = std::move(dev);\n"
                "        return true;\n"
                "    }\n"
                "    /* EvalKeySwitchPrecomputeCore (keyswitch-hybrid.cpp:330-420): the digits of c over Ql u P, resident */\n"
                "    static bool PimKeySwitchDigits(const DCRTPolyImpl& c, const std::shared_ptr<Params>& paramsP,\n"
                "                                   const std::shared_ptr<Params>& paramsQlP, size_t alpha, size_t maxDigits,\n"
                "                                   std::vector<DCRTPolyImpl>& out) {\n"
                "        if (!c.pim || c.m_format != Format::EVALUATION || c.m_vectors_.empty()) return false;\n"
                "        std::vector<uint64_t> pq, ppsi;\n"
                "        for (const auto& tp : paramsP->GetParams()) {\n"
                "            pq.push_back(tp->GetModulus().ConvertToInt());\n"
                "            ppsi.push_back(tp->GetRootOfUnity().ConvertToInt());\n"
                "        }\n"
                "        std::vector<std::shared_ptr<pimhook::Resident>> dig;\n"
                "        if (!pimhook::resident_keyswitch_digits(c.pim.get(), c.m_vectors_, c.pimres_, pq, ppsi, alpha, maxDigits, dig)) return false;\n"
                "        out.clear();\n"
                "        out.reserve(dig.size());\n"
                "        for (auto& r : dig) {\n"
                "            DCRTPolyImpl d(paramsQlP, Format::EVALUATION, false);\n"
                "            d.pimres_ = r;\n"
                "            d.pim     = c.pim;\n"
                "            out.push_back(std::move(d));\n"
                "        }\n"
                "        return true;\n"
                "    }\n"
                "    /* EvalFastKeySwitchCoreExt (keyswitch-hybrid.cpp:442-480): sum_j digits[j] * key[j], both halves, over Ql u P */\n"
                "    /* paramsQl given: EvalFastKeySwitchCore (:415-437), the inner product and ApproxModDown of both halves as one call */\n"
                "    static bool PimKeySwitchInner(const std::vector<DCRTPolyImpl>& digits, const std::vector<DCRTPolyImpl>& bv,\n"
                "                                  const std::vector<DCRTPolyImpl>& av, size_t sizeP, size_t alpha, const void* keyId,\n"
                "                                  DCRTPolyImpl& c0, DCRTPolyImpl& c1, const std::shared_ptr<Params>& paramsQl = nullptr,\n"
                "                                  uint64_t t = 0) {\n"
                "        if (digits.empty() || !digits[0].pim || bv.empty() || bv.size() != av.size()) return false;\n"
                "        for (const auto& d : digits)\n"
                "            if (d.m_format != Format::EVALUATION) return false;\n"
                "        for (size_t j = 0; j < bv.size(); j++)\n"
                "            if (bv[j].m_format != Format::EVALUATION || av[j].m_format != Format::EVALUATION) return false;\n"
                "        using Towers = std::vector<PolyType>;\n"
                "        auto digitAt = [&](size_t j) { return pimhook::KeyRef<Towers>{digits[j].m_vectors_, digits[j].pimres_}; };\n"
                "        auto keyAt = [&](int which, size_t j) {\n"
                "            const DCRTPolyImpl& k = which ? av[j] : bv[j];\n"
                "            return pimhook::KeyRef<Towers>{k.m_vectors_, k.pimres_};\n"
                "        };\n"
                "        std::shared_ptr<pimhook::Resident> r0, r1;\n"
                "        if (paramsQl && paramsQl->GetParams().size() + sizeP != digits[0].m_vectors_.size()) return false;\n"
                "        if (!pimhook::resident_keyswitch_inner<Towers>(digits[0].pim.get(), digits.size(), digitAt, bv.size(), keyAt, sizeP,\n"
                "                                                       alpha, keyId, r0, r1, paramsQl != nullptr, t))\n"
                "            return false;\n"
                "        c0         = DCRTPolyImpl(paramsQl ? paramsQl : digits[0].m_params, Format::EVALUATION, false);\n"
                "        c1         = DCRTPolyImpl(paramsQl ? paramsQl : digits[0].m_params, Format::EVALUATION, false);\n"
                "        c0.pimres_ = r0;\n"
                "        c1.pimres_ = r1;\n"
                "        c0.pim = c1.pim = digits[0].pim;\n"
                "        return true;\n"
                "    }\n", rel)
    # FHE_B200_ATTACH=1: every DCRTPoly is born attached (pim/PimHook.h default_handle)
    text = once(text, "    std::shared_ptr<PimManager> pim;\n", "    std::shared_ptr<PimManager> pim = pimhook::default_handle();\n", rel)
    text = once(text, "m_vectors{std::move(e.m_vectors)} {}",
                "m_vectors_{std::move(e.m_vectors_)}, pimres_{std::move(e.pimres_)}, pim{std::move(e.pim)} {}", rel)
    # move assignment: take the device copy along instead of forcing a download
    text = once(text, "        m_vectors = std::move(rhs.m_vectors);\n",
                "        m_vectors_ = std::move(rhs.m_vectors_);\n        pimres_    = std::move(rhs.pimres_);\n"
                "        pim        = std::move(rhs.pim);\n", rel)
    text = once(text, "DCRTPolyType& operator*=(const DCRTPolyType& rhs) override {",
                "DCRTPolyType& operator*=(const DCRTPolyType& rhs) override {" + HOOK_EQ % "FHE_OP_MUL", rel)
    # out-of-place Plus / Times: after the reference's own checks, compute on the device copies; the result carries
    # the handle so that a chain of operators stays in HBM
    for name, opc in (("Plus", "FHE_OP_ADD"), ("Times", "FHE_OP_MUL")):
        start = text.index("DCRTPolyType %s(const DCRTPolyType& rhs) const override {" % name)
        at = text.index("        DCRTPolyType tmp(m_params, m_format);\n", start)
        # the parameter checks read only sizes and moduli: let them look at the member itself, not force a download
        checks = text[start:at].replace("m_vectors", "m_vectors_")
        text = text[:start] + checks + HOOK_BIN % opc + text[at:]
    # everything else in this header reads the towers through the accessor
    # the tower COUNT needs no host values (and must not drop the device copy): read it from the member directly
    text = text.replace("m_vectors.size()", "m_vectors_.size()")
    head, sep, tail = text.partition("namespace lbcrypto {")
    return head + "#define m_vectors PimHostVectors()\n" + sep + tail + "\n#undef m_vectors\n"


def patch_impl(text):
    rel = "dcrtpoly-impl.h"
    text = once(text, "m_vectors(params->GetParams().size(), rhs) {", "m_vectors_(params->GetParams().size(), rhs) {", rel)
    text = once(text, "m_vectors(towers) {", "m_vectors_(towers) {", rel)
    text = once(text, "DCRTPolyImpl<VecType>& DCRTPolyImpl<VecType>::operator+=(const DCRTPolyImpl& rhs) {",
                "DCRTPolyImpl<VecType>& DCRTPolyImpl<VecType>::operator+=(const DCRTPolyImpl& rhs) {" + HOOK_EQ % "FHE_OP_ADD", rel)
    text = once(text, "DCRTPolyImpl<VecType>& DCRTPolyImpl<VecType>::operator-=(const DCRTPolyImpl& rhs) {",
                "DCRTPolyImpl<VecType>& DCRTPolyImpl<VecType>::operator-=(const DCRTPolyImpl& rhs) {" + HOOK_EQ % "FHE_OP_SUB", rel)
    text = once(text, "    if (m_vectors.size() != rhs.m_vectors.size())\n        OPENFHE_THROW(math_error, \"tower size mismatch; cannot subtract\");\n",
                "    if (m_vectors_.size() != rhs.m_vectors_.size())\n        OPENFHE_THROW(math_error, \"tower size mismatch; cannot subtract\");\n", rel)
    text = once(text, "        OPENFHE_THROW(math_error, \"tower size mismatch; cannot subtract\");\n",
                "        OPENFHE_THROW(math_error, \"tower size mismatch; cannot subtract\");\n"
                + (HOOK_BIN % "FHE_OP_SUB").replace("DCRTPolyType dev", "DCRTPolyImpl<VecType> dev").replace("\n    ", "\n"), rel)
    text = once(text, "void DCRTPolyImpl<VecType>::SwitchFormat() {",
                "void DCRTPolyImpl<VecType>::SwitchFormat() {"
                "\n    if (pim) {\n"
                "        const bool toEval = (m_format == Format::COEFFICIENT);\n"
                "        if (pimhook::resident_switch_format(pim.get(), m_vectors_, pimres_, toEval)) {\n"
                "            m_format = toEval ? Format::EVALUATION : Format::COEFFICIENT;\n"
                "            return;\n"
                "        }\n"
                "    }", rel)
    # ---- the neighbouring methods: same pattern, device forms of csrc/automorph.cu and
    # csrc/basis.cu.  Each hook leaves the reference's own checks / exceptions to the CPU path (odd index, formats).
    auto_hook = ("\n    if (pim && (i & 1)%s) {\n"
                 "        DCRTPolyImpl<VecType> dev(m_params, m_format, false);\n"
                 "        if (pimhook::resident_automorphism(pim.get(), m_vectors_, pimres_, dev.m_vectors_, dev.pimres_, i,\n"
                 "                                           m_format == Format::EVALUATION ? 0 : 1)) {\n"
                 "            dev.pim = pim;\n"
                 "            return dev;\n"
                 "        }\n"
                 "    }")
    sig = "DCRTPolyImpl<VecType> DCRTPolyImpl<VecType>::AutomorphismTransform(uint32_t i) const {"
    text = once(text, sig, sig + auto_hook % "", rel)
    sig = "DCRTPolyImpl<VecType> DCRTPolyImpl<VecType>::AutomorphismTransform(uint32_t i, const std::vector<uint32_t>& vec) const {"
    text = once(text, sig, sig + auto_hook % " && m_format == Format::EVALUATION", rel)
    # DropLastElementAndScale, EVALUATION input: shrink the host tower vector and the parameters exactly as
    # DropLastElement (dcrtpoly-impl.h:719-728) does, without touching the (stale) host values
    sig = "                                                    const std::vector<NativeInteger>& qlInvModqPrecon) {\n    // TODO: move this inside loop if omp enabled\n    auto lastPoly(m_vectors.back());"
    text = once(text, sig,
                "                                                    const std::vector<NativeInteger>& qlInvModqPrecon) {\n"
                "    if (pim && m_format == Format::EVALUATION && m_vectors_.size() > 1) {\n"
                "        auto full = pimhook::ensure_device(pim.get(), m_vectors_, pimres_);\n"
                "        m_vectors_.resize(m_vectors_.size() - 1);\n"
                "        DCRTPolyImpl::Params* newP = new DCRTPolyImpl::Params(*m_params);\n"
                "        newP->PopLastParam();\n"
                "        m_params.reset(newP);\n"
                "        pimres_.reset();\n"
                "        pimhook::resident_rescale(pim.get(), full, m_vectors_, pimres_);\n"
                "        return;\n"
                "    }\n"
                "    // TODO: move this inside loop if omp enabled\n    auto lastPoly(m_vectors.back());", rel)
    # ModReduce (in place, either format): same shrink as DropLastElementAndScale
    sig = ("                                      const std::vector<NativeInteger>& qlInvModqPrecon) {\n"
           "    // TODO: move this inside loop if omp enabled\n    DCRTPolyImpl::PolyType delta(m_vectors.back());")
    text = once(text, sig,
                "                                      const std::vector<NativeInteger>& qlInvModqPrecon) {\n"
                "    if (pim && m_vectors_.size() > 1) {\n"
                "        auto full = pimhook::ensure_device(pim.get(), m_vectors_, pimres_);\n"
                "        m_vectors_.resize(m_vectors_.size() - 1);\n"
                "        DCRTPolyImpl::Params* newP = new DCRTPolyImpl::Params(*m_params);\n"
                "        newP->PopLastParam();\n"
                "        m_params.reset(newP);\n"
                "        pimres_.reset();\n"
                "        pimhook::resident_mod_reduce(pim.get(), full, m_vectors_, pimres_, t.ConvertToInt(),\n"
                "                                     m_format == Format::EVALUATION ? 0 : 1);\n"
                "        return;\n"
                "    }\n"
                "    // TODO: move this inside loop if omp enabled\n    DCRTPolyImpl::PolyType delta(m_vectors.back());", rel)
    # ApproxModUp (in place: Q -> QP, result in EVALUATION format)
    sig = ("                                        const std::vector<DoubleNativeInt>& modpBarrettMu) {\n"
           "    std::vector<DCRTPolyImpl::PolyType> polyInNTT;\n")
    text = once(text, sig,
                "                                        const std::vector<DoubleNativeInt>& modpBarrettMu) {\n"
                "    if (pim && !m_vectors_.empty() && m_vectors_.size() == paramsQ->GetParams().size() &&\n"
                "        paramsQP->GetParams().size() == m_vectors_.size() + paramsP->GetParams().size()) {\n"
                "        std::vector<DCRTPolyImpl::PolyType> shell;\n"
                "        for (const auto& tp : paramsQP->GetParams()) shell.emplace_back(tp, Format::EVALUATION, false);\n"
                "        std::shared_ptr<pimhook::Resident> up;\n"
                "        if (pimhook::resident_mod_up(pim.get(), m_vectors_, pimres_, shell, up, m_format == Format::EVALUATION ? 0 : 1, QHatInvModq)) {\n"
                "            m_vectors_ = std::move(shell);\n"
                "            pimres_    = up;\n"
                "            m_format   = Format::EVALUATION;\n"
                "            m_params   = paramsQP;\n"
                "            return;\n"
                "        }\n"
                "    }\n"
                "    std::vector<DCRTPolyImpl::PolyType> polyInNTT;\n", rel)
    # ApproxSwitchCRTBasis (const, returns a polynomial over paramsP).  The hooks of the const methods sit BEFORE the
    # reference's zero-initialised `ans` so that a device result costs no host allocation.
    sig = ("    const std::vector<std::vector<NativeInteger>>& QHatModp, const std::vector<DoubleNativeInt>& modpBarrettMu) const {\n"
           "#if defined(HAVE_INT128) && NATIVEINT == 64\n    DCRTPolyImpl<VecType> ans(paramsP, m_format, true);\n")
    text = once(text, sig,
                "    const std::vector<std::vector<NativeInteger>>& QHatModp, const std::vector<DoubleNativeInt>& modpBarrettMu) const {\n"
                "    if (pim && m_vectors_.size() <= paramsQ->GetParams().size()) {\n"
                "        DCRTPolyImpl<VecType> dev(paramsP, m_format, false);\n"
                "        if (pimhook::resident_approx_switch(pim.get(), m_vectors_, pimres_, m_vectors_.size(), dev.m_vectors_, dev.pimres_, QHatInvModq)) {\n"
                "            dev.pim = pim;\n"
                "            return dev;\n"
                "        }\n"
                "    }\n"
                "#if defined(HAVE_INT128) && NATIVEINT == 64\n    DCRTPolyImpl<VecType> ans(paramsP, m_format, true);\n", rel)
    # BFV (HPS) set: SwitchCRTBasis (const), ExpandCRTBasis (in place), ScaleAndRound(paramsOutput, ...) (const)
    sig = ("                                                            const std::vector<double>& qInv) const {\n"
           "#if defined(HAVE_INT128) && NATIVEINT == 64\n    DCRTPolyImpl<VecType> ans(paramsP, m_format, true);\n")
    text = once(text, sig,
                "                                                            const std::vector<double>& qInv) const {\n"
                "    if (pim) {\n"
                "        DCRTPolyImpl<VecType> dev(paramsP, m_format, false);\n"
                "        if (pimhook::resident_switch_exact(pim.get(), m_vectors_, pimres_, dev.m_vectors_, dev.pimres_, QHatInvModq)) {\n"
                "            dev.pim = pim;\n"
                "            return dev;\n"
                "        }\n"
                "    }\n"
                "#if defined(HAVE_INT128) && NATIVEINT == 64\n    DCRTPolyImpl<VecType> ans(paramsP, m_format, true);\n", rel)
    start = text.index("void DCRTPolyImpl<VecType>::ExpandCRTBasis(\n")
    sig = ("    const std::vector<DoubleNativeInt>& modpBarrettMu, const std::vector<double>& qInv, Format resultFormat) {\n"
           "    std::vector<DCRTPolyImpl::PolyType> polyInNTT;\n")
    at = text.index(sig, start)
    text = text[:at] + text[at:].replace(sig,
                "    const std::vector<DoubleNativeInt>& modpBarrettMu, const std::vector<double>& qInv, Format resultFormat) {\n"
                "    if (pim && !m_vectors_.empty() && paramsQP->GetParams().size() == m_vectors_.size() + paramsP->GetParams().size()) {\n"
                "        std::vector<DCRTPolyImpl::PolyType> shell;\n"
                "        for (const auto& tp : paramsQP->GetParams()) shell.emplace_back(tp, resultFormat, false);\n"
                "        std::shared_ptr<pimhook::Resident> up;\n"
                "        if (pimhook::resident_expand(pim.get(), m_vectors_, pimres_, shell, up, m_format == Format::EVALUATION ? 0 : 1,\n"
                "                                     resultFormat == Format::EVALUATION ? 0 : 1, QHatInvModq)) {\n"
                "            m_vectors_ = std::move(shell);\n"
                "            pimres_    = up;\n"
                "            m_format   = resultFormat;\n"
                "            m_params   = paramsQP;\n"
                "            return;\n"
                "        }\n"
                "    }\n"
                "    std::vector<DCRTPolyImpl::PolyType> polyInNTT;\n", 1)
    sig = ("    DCRTPolyImpl<VecType> ans(paramsOutput, m_format, true);\n"
           "    usint ringDim      = m_params->GetRingDimension();\n")
    text = once(text, sig,
                "    if (pim) {\n"
                "        bool outLast = !(paramsOutput->GetParams()[0]->GetModulus() == m_params->GetParams()[0]->GetModulus());\n"
                "        DCRTPolyImpl<VecType> dev(paramsOutput, m_format, false);\n"
                "        if (pimhook::resident_scale_round(pim.get(), m_vectors_, pimres_, dev.m_vectors_, dev.pimres_, outLast,\n"
                "                                          tOSHatInvModsDivsFrac, tOSHatInvModsDivsModo)) {\n"
                "            dev.pim = pim;\n"
                "            return dev;\n"
                "        }\n"
                "    }\n"
                "    DCRTPolyImpl<VecType> ans(paramsOutput, m_format, true);\n"
                "    usint ringDim      = m_params->GetRingDimension();\n", rel)
    # FastExpandCRTBasisPloverQ (in place, Q -> QlPl, format unchanged; BFV HPSPOVERQ / HPSPOVERQLEVELED)
    sig = ("void DCRTPolyImpl<VecType>::FastExpandCRTBasisPloverQ(const Precomputations& precomputed) {\n"
           "    usint ringDim = m_params->GetRingDimension();\n")
    text = once(text, sig,
                "void DCRTPolyImpl<VecType>::FastExpandCRTBasisPloverQ(const Precomputations& precomputed) {\n"
                "    if (pim && !m_vectors_.empty()) {\n"
                "        std::vector<DCRTPolyImpl::PolyType> shell;\n"
                "        for (const auto& tp : precomputed.paramsQlPl->GetParams()) shell.emplace_back(tp, m_format, false);\n"
                "        std::shared_ptr<pimhook::Resident> up;\n"
                "        if (pimhook::resident_fast_expand_pl_over_q(pim.get(), m_vectors_, pimres_, shell, up, precomputed.mPlQHatInvModq,\n"
                "                                                    precomputed.qInvModp, precomputed.PlHatInvModp)) {\n"
                "            m_vectors_ = std::move(shell);\n"
                "            pimres_    = up;\n"
                "            m_params   = precomputed.paramsQlPl;\n"
                "            return;\n"
                "        }\n"
                "    }\n"
                "    usint ringDim = m_params->GetRingDimension();\n", rel)
    # ExpandCRTBasisReverseOrder (in place, Q -> PQ)
    start = text.index("void DCRTPolyImpl<VecType>::ExpandCRTBasisReverseOrder(\n")
    sig = ("    const std::vector<DoubleNativeInt>& modpBarrettMu, const std::vector<double>& qInv, Format resultFormat) {\n"
           "    std::vector<DCRTPolyImpl::PolyType> polyInNTT;\n")
    at = text.index(sig, start)
    text = text[:at] + text[at:].replace(sig,
                "    const std::vector<DoubleNativeInt>& modpBarrettMu, const std::vector<double>& qInv, Format resultFormat) {\n"
                "    if (pim && !m_vectors_.empty() && paramsQP->GetParams().size() == m_vectors_.size() + paramsP->GetParams().size()) {\n"
                "        std::vector<DCRTPolyImpl::PolyType> shell;\n"
                "        for (const auto& tp : paramsQP->GetParams()) shell.emplace_back(tp, resultFormat, false);\n"
                "        std::shared_ptr<pimhook::Resident> up;\n"
                "        if (pimhook::resident_expand(pim.get(), m_vectors_, pimres_, shell, up, m_format == Format::EVALUATION ? 0 : 1,\n"
                "                                     resultFormat == Format::EVALUATION ? 0 : 1, QHatInvModq, true)) {\n"
                "            m_vectors_ = std::move(shell);\n"
                "            pimres_    = up;\n"
                "            m_format   = resultFormat;\n"
                "            m_params   = paramsQP;\n"
                "            return;\n"
                "        }\n"
                "    }\n"
                "    std::vector<DCRTPolyImpl::PolyType> polyInNTT;\n", 1)
    # ExpandCRTBasisQlHat (in place, Ql -> Q) and ScaleAndRoundPOverQ (in place, Q plus one tower -> Q)
    sig = ("                                                const std::vector<NativeInteger>& QlHatModqPrecon, const usint sizeQ) {\n"
           "    size_t sizeQl(m_vectors.size());\n")
    text = once(text, sig,
                "                                                const std::vector<NativeInteger>& QlHatModqPrecon, const usint sizeQ) {\n"
                "    if (pim && !m_vectors_.empty() && paramsQ->GetParams().size() == sizeQ) {\n"
                "        std::vector<DCRTPolyImpl::PolyType> shell;\n"
                "        for (const auto& tp : paramsQ->GetParams()) shell.emplace_back(tp, m_format, false);\n"
                "        std::shared_ptr<pimhook::Resident> up;\n"
                "        if (pimhook::resident_tower_scale(pim.get(), 0, m_vectors_, pimres_, shell, up, QlHatModq)) {\n"
                "            m_vectors_ = std::move(shell);\n"
                "            pimres_    = up;\n"
                "            m_params   = paramsQ;\n"
                "            return;\n"
                "        }\n"
                "    }\n"
                "    size_t sizeQl(m_vectors.size());\n", rel)
    sig = ("                                                const std::vector<NativeInteger>& pInvModq) {\n"
           "    const usint sizeQ   = m_vectors.size() - 1;\n")
    text = once(text, sig,
                "                                                const std::vector<NativeInteger>& pInvModq) {\n"
                "    if (pim && m_vectors_.size() == paramsQ->GetParams().size() + 1) {\n"
                "        std::vector<DCRTPolyImpl::PolyType> shell;\n"
                "        for (const auto& tp : paramsQ->GetParams()) shell.emplace_back(tp, m_format, false);\n"
                "        std::shared_ptr<pimhook::Resident> up;\n"
                "        if (pimhook::resident_tower_scale(pim.get(), 1, m_vectors_, pimres_, shell, up, pInvModq)) {\n"
                "            m_vectors_ = std::move(shell);\n"
                "            pimres_    = up;\n"
                "            m_params   = paramsQ;\n"
                "            return;\n"
                "        }\n"
                "    }\n"
                "    const usint sizeQ   = m_vectors.size() - 1;\n", rel)
    # Times(const std::vector<NativeInteger>&) (const): one constant per tower
    sig = ("DCRTPolyImpl<VecType> DCRTPolyImpl<VecType>::Times(const std::vector<NativeInteger>& rhs) const {\n"
           "    if (m_vectors.size() != rhs.size())\n"
           "        OPENFHE_THROW(math_error, \"tower size mismatch; cannot multiply\");\n")
    text = once(text, sig,
                "DCRTPolyImpl<VecType> DCRTPolyImpl<VecType>::Times(const std::vector<NativeInteger>& rhs) const {\n"
                "    if (m_vectors_.size() != rhs.size())\n"
                "        OPENFHE_THROW(math_error, \"tower size mismatch; cannot multiply\");\n"
                "    if (pim && !m_vectors_.empty()) {\n"
                "        DCRTPolyImpl<VecType> dev(m_params, m_format, false);\n"
                "        if (pimhook::resident_tower_scale(pim.get(), 2, m_vectors_, pimres_, dev.m_vectors_, dev.pimres_, rhs)) {\n"
                "            dev.pim = pim;\n"
                "            return dev;\n"
                "        }\n"
                "    }\n", rel)
    # ApproxModDown (const, EVALUATION input over QP, returns a polynomial over Q)
    sig = ("    usint sizeQP = m_vectors.size();\n    usint sizeP  = paramsP->GetParams().size();\n    usint sizeQ  = sizeQP - sizeP;\n")
    text = once(text, sig,
                "    if (pim && m_format == Format::EVALUATION && m_vectors_.size() > paramsP->GetParams().size()) {\n"
                "        usint devQ = m_vectors_.size() - paramsP->GetParams().size();\n"
                "        DCRTPolyImpl<VecType> dev(paramsQ, Format::EVALUATION, false);\n"
                "        if (paramsQ->GetParams().size() > devQ)\n"
                "            dev.DropLastElements(paramsQ->GetParams().size() - devQ);\n"
                "        if (pimhook::resident_mod_down(pim.get(), m_vectors_, pimres_, devQ, t.ConvertToInt(), dev.m_vectors_, dev.pimres_, PHatInvModp)) {\n"
                "            dev.pim = pim;\n"
                "            return dev;\n"
                "        }\n"
                "    }\n" + sig, rel)
    # Times(std::vector<Integer>) (dcrtpoly-impl.h:608-615: CKKS scaling-factor adjustments): the per-tower constants as
    # NativeIntegers, then the same device path as Times(std::vector<NativeInteger>)
    sig = ("DCRTPolyImpl<VecType> DCRTPolyImpl<VecType>::Times(const std::vector<Integer>& crtElement) const {\n"
           "    DCRTPolyImpl<VecType> tmp(m_params, m_format);\n")
    text = once(text, sig,
                "DCRTPolyImpl<VecType> DCRTPolyImpl<VecType>::Times(const std::vector<Integer>& crtElement) const {\n"
                "    if (pim && !m_vectors_.empty() && crtElement.size() >= m_vectors_.size()) {\n"
                "        std::vector<NativeInteger> nat(m_vectors_.size());\n"
                "        for (size_t i = 0; i < nat.size(); ++i) nat[i] = NativeInteger(crtElement[i]);\n"
                "        DCRTPolyImpl<VecType> dev(m_params, m_format, false);\n"
                "        if (pimhook::resident_tower_scale(pim.get(), 2, m_vectors_, pimres_, dev.m_vectors_, dev.pimres_, nat)) {\n"
                "            dev.pim = pim;\n"
                "            return dev;\n"
                "        }\n"
                "    }\n"
                "    DCRTPolyImpl<VecType> tmp(m_params, m_format);\n", rel)
    # Plus / Minus(std::vector<Integer>) (dcrtpoly-impl.h:556-563, 577-584: CKKS EvalAdd / EvalSub with a real constant)
    for name, opc, cond in (("Plus", "FHE_OP_ADD", " && m_format == Format::EVALUATION"), ("Minus", "FHE_OP_SUB", "")):
        sig = ("DCRTPolyImpl<VecType> DCRTPolyImpl<VecType>::%s(const std::vector<Integer>& crtElement) const {\n"
               "    DCRTPolyImpl<VecType> tmp(m_params, m_format);\n" % name)
        text = once(text, sig,
                    "DCRTPolyImpl<VecType> DCRTPolyImpl<VecType>::%s(const std::vector<Integer>& crtElement) const {\n"
                    "    if (pim && !m_vectors_.empty() && crtElement.size() >= m_vectors_.size()%s) {\n"
                    "        std::vector<NativeInteger> nat(m_vectors_.size());\n"
                    "        for (size_t i = 0; i < nat.size(); ++i) nat[i] = NativeInteger(crtElement[i]);\n"
                    "        DCRTPolyImpl<VecType> dev(m_params, m_format, false);\n"
                    "        if (pimhook::resident_tower_scalars(pim.get(), %s, m_vectors_, pimres_, dev.m_vectors_, dev.pimres_, nat)) {\n"
                    "            dev.pim = pim;\n"
                    "            return dev;\n"
                    "        }\n"
                    "    }\n"
                    "    DCRTPolyImpl<VecType> tmp(m_params, m_format);\n" % (name, cond, opc), rel)
    # DropLastElement / DropLastElements on a resident polynomial: the device copy shrinks to a prefix view, nothing moves
    sig = ("        OPENFHE_THROW(config_error, \"Removing last element of DCRTPoly object renders it invalid!\");\n"
           "    m_vectors.resize(m_vectors.size() - 1);\n")
    text = once(text, sig,
                "        OPENFHE_THROW(config_error, \"Removing last element of DCRTPoly object renders it invalid!\");\n"
                "    if (pimres_) {\n"
                "        m_vectors_.resize(m_vectors_.size() - 1);\n"
                "        DCRTPolyImpl::Params* shorter = new DCRTPolyImpl::Params(*m_params);\n"
                "        shorter->PopLastParam();\n"
                "        m_params.reset(shorter);\n"
                "        pimres_ = pimhook::resident_prefix(pimres_->ring->manager(), pimres_, m_vectors_);\n"
                "        return;\n"
                "    }\n"
                "    m_vectors.resize(m_vectors.size() - 1);\n", rel)
    sig = ("                      \"perform the modulus reduction\");\n    }\n"
           "    m_vectors.resize(m_vectors.size() - i);\n")
    text = once(text, sig,
                "                      \"perform the modulus reduction\");\n    }\n"
                "    if (pimres_) {\n"
                "        m_vectors_.resize(m_vectors_.size() - i);\n"
                "        DCRTPolyImpl::Params* shorter = new DCRTPolyImpl::Params(*m_params);\n"
                "        for (size_t j = 0; j < i; j++)\n"
                "            shorter->PopLastParam();\n"
                "        m_params.reset(shorter);\n"
                "        pimres_ = pimhook::resident_prefix(pimres_->ring->manager(), pimres_, m_vectors_);\n"
                "        return;\n"
                "    }\n"
                "    m_vectors.resize(m_vectors.size() - i);\n", rel)
    # CRTDecompose(0) (const, returns one polynomial per tower: the digits of KeySwitchBV)
    sig = ("std::vector<DCRTPolyImpl<VecType>> DCRTPolyImpl<VecType>::CRTDecompose(uint32_t baseBits) const {\n"
           "    DCRTPolyImpl<VecType> cp(*this);\n")
    text = once(text, sig,
                "std::vector<DCRTPolyImpl<VecType>> DCRTPolyImpl<VecType>::CRTDecompose(uint32_t baseBits) const {\n"
                "    if (pim && baseBits == 0 && !m_vectors_.empty()) {\n"
                "        std::vector<std::shared_ptr<pimhook::Resident>> dig;\n"
                "        if (pimhook::resident_crt_decompose(pim.get(), m_vectors_, pimres_, m_format == Format::EVALUATION ? 0 : 1, dig)) {\n"
                "            std::vector<DCRTPolyType> result;\n"
                "            result.reserve(dig.size());\n"
                "            for (auto& r : dig) {\n"
                "                DCRTPolyImpl<VecType> d(m_params, Format::EVALUATION, false);\n"
                "                d.pimres_ = r;\n"
                "                d.pim     = pim;\n"
                "                result.push_back(std::move(d));\n"
                "            }\n"
                "            return result;\n"
                "        }\n"
                "    }\n"
                "    DCRTPolyImpl<VecType> cp(*this);\n", rel)
    # BEHZ: FastBaseConvqToBskMontgomery (in place, Q -> QBsk, EVALUATION), FastRNSFloorq (in place over QBsk),
    # FastBaseConvSK (in place, QBsk -> Q)
    sig = ("    const std::vector<NativeInteger>& mtildeInvModbskPrecon) {\n"
           "    // Input: dcrtpoly in basis Q\n")
    text = once(text, sig,
                "    const std::vector<NativeInteger>& mtildeInvModbskPrecon) {\n"
                "    if (pim && !m_vectors_.empty() && m_vectors_.size() == moduliQ.size() &&\n"
                "        paramsQBsk->GetParams().size() == moduliQ.size() + moduliBsk.size()) {\n"
                "        std::vector<DCRTPolyImpl::PolyType> shell;\n"
                "        for (const auto& tp : paramsQBsk->GetParams()) shell.emplace_back(tp, Format::EVALUATION, false);\n"
                "        std::shared_ptr<pimhook::Resident> up;\n"
                "        if (pimhook::resident_behz_q_to_bsk(pim.get(), m_vectors_, pimres_, shell, up,\n"
                "                                            m_format == Format::EVALUATION ? 0 : 1, mtildeQHatInvModq)) {\n"
                "            m_vectors_ = std::move(shell);\n"
                "            pimres_    = up;\n"
                "            m_format   = Format::EVALUATION;\n"
                "            m_params   = paramsQBsk;\n"
                "            return;\n"
                "        }\n"
                "    }\n"
                "    // Input: dcrtpoly in basis Q\n", rel)
    sig = ("    const std::vector<NativeInteger>& tQInvModbskPrecon) {\n"
           "    // Input: poly in basis {q U Bsk}\n")
    text = once(text, sig,
                "    const std::vector<NativeInteger>& tQInvModbskPrecon) {\n"
                "    if (pim && m_vectors_.size() == moduliQ.size() + moduliBsk.size() &&\n"
                "        pimhook::resident_behz_floor(pim.get(), m_vectors_, pimres_, moduliQ.size(), t.ConvertToInt(), tQHatInvModq))\n"
                "        return;\n"
                "    // Input: poly in basis {q U Bsk}\n", rel)
    sig = ("    const std::vector<NativeInteger>& BModq, const std::vector<NativeInteger>& BModqPrecon) {\n"
           "    // Input: poly in basis Bsk\n")
    text = once(text, sig,
                "    const std::vector<NativeInteger>& BModq, const std::vector<NativeInteger>& BModqPrecon) {\n"
                "    if (pim && m_vectors_.size() == paramsQ->GetParams().size() + moduliBsk.size()) {\n"
                "        const size_t devQ = paramsQ->GetParams().size();\n"
                "        std::shared_ptr<pimhook::Resident> down;\n"
                "        if (pimhook::resident_behz_conv_sk(pim.get(), m_vectors_, pimres_, devQ, down, BHatInvModb)) {\n"
                "            m_vectors_.resize(devQ);\n"
                "            pimres_  = down;\n"
                "            m_params = paramsQ;\n"
                "            return;\n"
                "        }\n"
                "    }\n"
                "    // Input: poly in basis Bsk\n", rel)
    # the tower COUNT needs no host values (and must not drop the device copy): read it from the member directly
    text = text.replace("m_vectors.size()", "m_vectors_.size()")
    head, sep, tail = text.partition("namespace lbcrypto {")
    return head + "#define m_vectors PimHostVectors()\n" + sep + tail + "\n#undef m_vectors\n"


def patch_interface(text):
    return once(text, "        return this->GetDerived().GetAllElements().size();\n",
                "        return this->GetDerived().PimNumTowers();\n", "dcrtpoly-interface.h")


rewrite("lattice/hal/dcrtpoly-interface.h", patch_interface)
rewrite("lattice/hal/default/dcrtpoly.h", patch_class)
rewrite("lattice/hal/default/dcrtpoly-impl.h", patch_impl)

# ---- scheme layer (only the `pke` build uses it): a patched COPY of keyswitch-hybrid.cpp next to the header tree ----
pke_src = os.path.join(ref, "src", "pke", "lib", "keyswitch", "keyswitch-hybrid.cpp")
if os.path.exists(pke_src):
    text = open(pke_src).read()
    sig = ("                                                                      const EvalKey<DCRTPoly> evalKey) const {\n"
           "    return EvalFastKeySwitchCore(EvalKeySwitchPrecomputeCore(a, evalKey->GetCryptoParameters()), evalKey,\n")
    text = once(text, sig,
                "                                                                      const EvalKey<DCRTPoly> evalKey) const {\n"
                "    {   // B200 intercept: the whole key switch as one device composition when `a` carries the handle\n"
                "        const auto cp = std::dynamic_pointer_cast<CryptoParametersRNS>(evalKey->GetCryptoParameters());\n"
                "        PlaintextModulus t = (cp->GetNoiseScale() == 1) ? 0 : cp->GetPlaintextModulus();\n"
                "        auto res = std::make_shared<std::vector<DCRTPoly>>(2);\n"
                "        if (DCRTPoly::PimKeySwitchHybrid(a, evalKey->GetBVector(), evalKey->GetAVector(),\n"
                "                                         cp->GetParamsP()->GetParams().size(), cp->GetNumPerPartQ(), t, evalKey.get(),\n"
                "                                         (*res)[0], (*res)[1]))\n"
                "            return res;\n"
                "    }\n"
                "    return EvalFastKeySwitchCore(EvalKeySwitchPrecomputeCore(a, evalKey->GetCryptoParameters()), evalKey,\n",
                "keyswitch-hybrid.cpp")
    # the two halves, called separately by the hoisted rotations (EvalFastRotationPrecompute / EvalFastRotation)
    sig = ("    if (numPartQl > cryptoParams->GetNumberOfQPartitions())\n"
           "        numPartQl = cryptoParams->GetNumberOfQPartitions();\n\n"
           "    std::vector<DCRTPoly> partsCt(numPartQl);\n")
    text = once(text, sig,
                "    if (numPartQl > cryptoParams->GetNumberOfQPartitions())\n"
                "        numPartQl = cryptoParams->GetNumberOfQPartitions();\n\n"
                "    {   // B200 intercept: the digits as one device composition, resident\n"
                "        std::vector<DCRTPoly> dev;\n"
                "        if (DCRTPoly::PimKeySwitchDigits(c, paramsP, paramsQlP, alpha, cryptoParams->GetNumberOfQPartitions(), dev))\n"
                "            return std::make_shared<std::vector<DCRTPoly>>(std::move(dev));\n"
                "    }\n"
                "    std::vector<DCRTPoly> partsCt(numPartQl);\n", "keyswitch-hybrid.cpp")
    sig = ("    size_t sizeQ   = cryptoParams->GetElementParams()->GetParams().size();\n\n"
           "    DCRTPoly cTilda0(paramsQlP, Format::EVALUATION, true);\n")
    text = once(text, sig,
                "    size_t sizeQ   = cryptoParams->GetElementParams()->GetParams().size();\n\n"
                "    {   // B200 intercept: the inner product with the key on the device copies\n"
                "        auto res = std::make_shared<std::vector<DCRTPoly>>(2);\n"
                "        if (DCRTPoly::PimKeySwitchInner(*digits, bv, av, paramsP->GetParams().size(), cryptoParams->GetNumPerPartQ(),\n"
                "                                        evalKey.get(), (*res)[0], (*res)[1]))\n"
                "            return res;\n"
                "    }\n"
                "    DCRTPoly cTilda0(paramsQlP, Format::EVALUATION, true);\n", "keyswitch-hybrid.cpp")
    text = once(text,
                "    std::shared_ptr<std::vector<DCRTPoly>> cTilda = EvalFastKeySwitchCoreExt(digits, evalKey, paramsQl);\n\n"
                "    PlaintextModulus t = (cryptoParams->GetNoiseScale() == 1) ? 0 : cryptoParams->GetPlaintextModulus();\n",
                "    {   // B200 intercept: the inner product with the key and ApproxModDown of both halves as one device call\n"
                "        PlaintextModulus t0 = (cryptoParams->GetNoiseScale() == 1) ? 0 : cryptoParams->GetPlaintextModulus();\n"
                "        auto res = std::make_shared<std::vector<DCRTPoly>>(2);\n"
                "        if (DCRTPoly::PimKeySwitchInner(*digits, evalKey->GetBVector(), evalKey->GetAVector(),\n"
                "                                        cryptoParams->GetParamsP()->GetParams().size(), cryptoParams->GetNumPerPartQ(),\n"
                "                                        evalKey.get(), (*res)[0], (*res)[1], paramsQl, t0))\n"
                "            return res;\n"
                "    }\n"
                "    std::shared_ptr<std::vector<DCRTPoly>> cTilda = EvalFastKeySwitchCoreExt(digits, evalKey, paramsQl);\n\n"
                "    PlaintextModulus t = (cryptoParams->GetNoiseScale() == 1) ? 0 : cryptoParams->GetPlaintextModulus();\n",
                "keyswitch-hybrid.cpp")
    sig = ("        if ((addFirst) || (k > 0)) {\n"
           "            auto cMult = cv[k].TimesNoCheck(cryptoParams->GetPModq());\n")
    text = once(text, sig,
                "        if (((addFirst) || (k > 0)) && DCRTPoly::PimExtendTimes(cv[k], cryptoParams->GetPModq(), paramsQlP, resultElements[k]))\n"
                "            continue;      // B200 intercept: c * [P]_q extended by zero towers, on the device copy\n"
                "        if ((addFirst) || (k > 0)) {\n"
                "            auto cMult = cv[k].TimesNoCheck(cryptoParams->GetPModq());\n", "keyswitch-hybrid.cpp")
    os.makedirs(os.path.join(out, "_pke_src", "keyswitch"), exist_ok=True)
    open(os.path.join(out, "_pke_src", "keyswitch", "keyswitch-hybrid.cpp"), "w").write(text)
    # EvalFastRotationExt (ckksrns-leveledshe.cpp:402-460): the addFirst term built the same way
    rel2 = os.path.join("scheme", "ckksrns", "ckksrns-leveledshe.cpp")
    text = open(os.path.join(ref, "src", "pke", "lib", rel2)).read()
    sig = ("        DCRTPoly psiC0       = DCRTPoly(paramsQlP, Format::EVALUATION, true);\n"
           "        auto cMult           = ciphertext->GetElements()[0].TimesNoCheck(cryptoParams->GetPModq());\n"
           "        for (usint i = 0; i < sizeQl; i++) {\n"
           "            psiC0.SetElementAtIndex(i, cMult.GetElementAtIndex(i));\n"
           "        }\n")
    text = once(text, sig,
                "        DCRTPoly psiC0;\n"
                "        if (!DCRTPoly::PimExtendTimes(ciphertext->GetElements()[0], cryptoParams->GetPModq(), paramsQlP, psiC0)) {      // B200 intercept\n"
                "            psiC0      = DCRTPoly(paramsQlP, Format::EVALUATION, true);\n"
                "            auto cMult = ciphertext->GetElements()[0].TimesNoCheck(cryptoParams->GetPModq());\n"
                "            for (usint i = 0; i < sizeQl; i++) {\n"
                "                psiC0.SetElementAtIndex(i, cMult.GetElementAtIndex(i));\n"
                "            }\n"
                "        }\n", "ckksrns-leveledshe.cpp")
    # reading a tower's modulus must not fetch the tower: X.GetElementAtIndex(i).GetModulus() -> the same value from the
    # parameters (ckksrns-leveledshe.cpp:167,216,310,334; bfvrns-leveledshe.cpp:240,462,772,813,858)
    import re
    mod_rx = re.compile(r"(\b\w+(?:\[\d+\])?)\.GetElementAtIndex\((\w+)\)\.GetModulus\(\)")
    text, nsub = mod_rx.subn(r"\1.GetParams()->GetParams()[\2]->GetModulus()", text)
    if nsub != 4:
        raise SystemExit(f"hook_patch: expected 4 modulus lookups in ckksrns-leveledshe.cpp, found {nsub}")
    os.makedirs(os.path.join(out, "_pke_src", "scheme", "ckksrns"), exist_ok=True)
    open(os.path.join(out, "_pke_src", rel2), "w").write(text)
    rel3 = os.path.join("scheme", "bfvrns", "bfvrns-leveledshe.cpp")
    text, nsub = mod_rx.subn(r"\1.GetParams()->GetParams()[\2]->GetModulus()", open(os.path.join(ref, "src", "pke", "lib", rel3)).read())
    if nsub != 5:
        raise SystemExit(f"hook_patch: expected 5 modulus lookups in bfvrns-leveledshe.cpp, found {nsub}")
    # LeveledSHEBFVRNS::EvalMult's products over the extended basis (bfvrns-leveledshe.cpp:324-336), 2 x 2: one launch
    text = once(text,
                "#else\n"
                "    std::vector<bool> isFirstAdd(cvMultSize, true);\n"
                "    for (size_t i = 0; i < cv1Size; i++) {\n",
                "#else\n"
                "    std::vector<bool> isFirstAdd(cvMultSize, true);\n"
                "    if (cv1Size == 2 && cv2Size == 2 && DCRTPoly::PimTensorProduct(cv1[0], cv1[1], cv2[0], cv2[1], cvMult))\n"
                "        cv1Size = 0;      // B200 intercept: done, skip the loops\n"
                "    for (size_t i = 0; i < cv1Size; i++) {\n", "bfvrns-leveledshe.cpp")
    os.makedirs(os.path.join(out, "_pke_src", "scheme", "bfvrns"), exist_ok=True)
    open(os.path.join(out, "_pke_src", rel3), "w").write(text)
    # the scheme's bilinear step as one launch, and sums of two ciphertexts without the copy that precedes the in-place form
    # (base-leveledshe.cpp:591-598, 624-629, 657-672; bfvrns-leveledshe.cpp:324-336)
    rel4 = os.path.join("schemebase", "base-leveledshe.cpp")
    text = open(os.path.join(ref, "src", "pke", "lib", rel4)).read()
    text = once(text,
                "    std::vector<Element> cv1        = ciphertext1->GetElements();\n"
                "    const std::vector<Element>& cv2 = ciphertext2->GetElements();\n\n"
                "    size_t cResultSize = cv1.size() + cv2.size() - 1;\n",
                "    // B200 intercept: the copy of ciphertext1's elements (the CPU path multiplies two of them in place) only when needed\n"
                "    const std::vector<Element>& cv1r = ciphertext1->GetElements();\n"
                "    const std::vector<Element>& cv2  = ciphertext2->GetElements();\n"
                "    {\n"
                "        std::vector<Element> dev(3);\n"
                "        if (cv1r.size() == 2 && cv2.size() == 2 && DCRTPoly::PimTensorProduct(cv1r[0], cv1r[1], cv2[0], cv2[1], dev)) {\n"
                "            result->SetElements(std::move(dev));\n"
                "            result->SetNoiseScaleDeg(ciphertext1->GetNoiseScaleDeg() + ciphertext2->GetNoiseScaleDeg());\n"
                "            result->SetScalingFactor(ciphertext1->GetScalingFactor() * ciphertext2->GetScalingFactor());\n"
                "            const auto plainMod0 = ciphertext1->GetCryptoParameters()->GetPlaintextModulus();\n"
                "            result->SetScalingFactorInt(\n"
                "                ciphertext1->GetScalingFactorInt().ModMul(ciphertext2->GetScalingFactorInt(), plainMod0));\n"
                "            return result;\n"
                "        }\n"
                "    }\n"
                "    std::vector<Element> cv1 = cv1r;\n\n"
                "    size_t cResultSize = cv1.size() + cv2.size() - 1;\n", "base-leveledshe.cpp")
    for name, op in (("Add", "+"), ("Sub", "-")):
        text = once(text,
                    "    auto result = ciphertext1->Clone();\n"
                    f"    Eval{name}CoreInPlace(result, ciphertext2);\n"
                    "    return result;\n",
                    "    {   // B200 intercept: out-of-place sums (one launch each on the device copies) instead of a copy and an in-place sum\n"
                    "        const std::vector<Element>& cv1 = ciphertext1->GetElements();\n"
                    "        const std::vector<Element>& cv2 = ciphertext2->GetElements();\n"
                    "        if (ciphertext1->GetPim() == nullptr && !cv1.empty() && cv1.size() == cv2.size()) {\n"
                    "            auto sum = ciphertext1->CloneZero();\n"
                    "            std::vector<Element> cv;\n"
                    "            cv.reserve(cv1.size());\n"
                    f"            for (size_t i = 0; i < cv1.size(); i++) cv.push_back(cv1[i] {op} cv2[i]);\n"
                    "            sum->SetElements(std::move(cv));\n"
                    "            return sum;\n"
                    "        }\n"
                    "    }\n"
                    "    auto result = ciphertext1->Clone();\n"
                    f"    Eval{name}CoreInPlace(result, ciphertext2);\n"
                    "    return result;\n", "base-leveledshe.cpp")
    # relinearisation: cv[0] += (*ab)[0]; cv[1] += (*ab)[1] after KeySwitchCore, seven places (base-leveledshe.cpp:210-213, ...)
    site = ("    std::shared_ptr<std::vector<Element>> ab = algo->KeySwitchCore(cv[2], evalKey);\n\n"
            "    cv[0] += (*ab)[0];\n"
            "    cv[1] += (*ab)[1];\n")
    if text.count(site) != 7:
        raise SystemExit(f"hook_patch: expected 7 relinearisation sites in base-leveledshe.cpp, found {text.count(site)}")
    text = text.replace(site,
                        "    if (!PimRelinearizeAdd(cv, evalKey)) {      // B200 intercept: key switch and both sums as one device composition\n"
                        "        std::shared_ptr<std::vector<Element>> ab = algo->KeySwitchCore(cv[2], evalKey);\n"
                        "        cv[0] += (*ab)[0];\n"
                        "        cv[1] += (*ab)[1];\n"
                        "    }\n")
    if text.count("namespace lbcrypto {\n") != 2:
        raise SystemExit("hook_patch: base-leveledshe.cpp no longer opens namespace lbcrypto twice")
    text = text.replace("namespace lbcrypto {\n",
                "#include \"schemerns/rns-cryptoparameters.h\"\n\n"
                "namespace lbcrypto {\n\n"
                "// B200 intercept: cv[0] += ks0, cv[1] += ks1 for the hybrid key switch of cv[2] (KeySwitchHYBRID::KeySwitchCore)\n"
                "template <class Element>\n"
                "static bool PimRelinearizeAdd(std::vector<Element>& cv, const EvalKey<Element>& evalKey) {\n"
                "    const auto cp = std::dynamic_pointer_cast<CryptoParametersRNS>(evalKey->GetCryptoParameters());\n"
                "    if (!cp || cp->GetKeySwitchTechnique() != HYBRID || cv.size() < 3) return false;\n"
                "    PlaintextModulus t = (cp->GetNoiseScale() == 1) ? 0 : cp->GetPlaintextModulus();\n"
                "    return DCRTPoly::PimKeySwitchHybridAdd(cv[2], evalKey->GetBVector(), evalKey->GetAVector(),\n"
                "                                           cp->GetParamsP()->GetParams().size(), cp->GetNumPerPartQ(), t, evalKey.get(),\n"
                "                                           cv[0], cv[1]);\n"
                "}\n", 1)
    os.makedirs(os.path.join(out, "_pke_src", "schemebase"), exist_ok=True)
    open(os.path.join(out, "_pke_src", rel4), "w").write(text)
print("hook_patch: patched header tree in", out)
