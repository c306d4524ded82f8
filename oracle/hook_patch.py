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
     SwitchFormat (dcrtpoly-impl.h:410-416, 438-444, 2518-2524; dcrtpoly.h:142-148).

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

HOOK_EQ = "\n    if (pim && pimhook::dcrt_eq(pim.get(), m_vectors, rhs.m_vectors, %s))\n        return *this;"
patch("lattice/hal/default/dcrtpoly.h", [
    ('#include "utils/parallel.h"\n', '#include "pim/PimHook.h"\n', "after"),
    ("DCRTPolyType& operator*=(const DCRTPolyType& rhs) override {", HOOK_EQ % "FHE_OP_MUL", "after"),
])
patch("lattice/hal/default/dcrtpoly-impl.h", [
    ("DCRTPolyImpl<VecType>& DCRTPolyImpl<VecType>::operator+=(const DCRTPolyImpl& rhs) {", HOOK_EQ % "FHE_OP_ADD", "after"),
    ("DCRTPolyImpl<VecType>& DCRTPolyImpl<VecType>::operator-=(const DCRTPolyImpl& rhs) {", HOOK_EQ % "FHE_OP_SUB", "after"),
    ("void DCRTPolyImpl<VecType>::SwitchFormat() {",
     "\n    if (pim) {\n"
     "        const bool toEval = (m_format == Format::COEFFICIENT);\n"
     "        const Format nf   = toEval ? Format::EVALUATION : Format::COEFFICIENT;\n"
     "        if (pimhook::dcrt_switch_format(pim.get(), m_vectors, toEval, nf)) {\n"
     "            m_format = nf;\n"
     "            return;\n"
     "        }\n"
     "    }", "after"),
])
print("hook_patch: patched header tree in", out)
