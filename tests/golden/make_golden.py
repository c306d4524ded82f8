"""Generates tests/golden/ref_vectors.json from the REFERENCE itself.

Runs oracle/_ref/libopenfhe_ref.so — the reference's own sources compiled by oracle/Makefile —
on seeded inputs and records (a) the parameters its FirstPrime/PreviousPrime/RootOfUnity produce,
(b) SHA-256 digests of the outputs of DCRTPoly += / -= / *= and SwitchFormat (both directions),
(c) full output vectors for the small cases.  Inputs are reproducible from (seed, q, index) with the
counter-based generator of oracle/fhe_oracle.c, so only digests have to be stored for large N.

Only runs where /root/reference was available to build oracle/_ref (the build container):
    make -C oracle ref && python tests/golden/make_golden.py
"""
import ctypes
import hashlib
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from tests import oracle_py  # noqa: E402

vp, sz, ci, u32, u64 = ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int, ctypes.c_uint32, ctypes.c_uint64

CASES = [  # nbits, N, towers, batch, seed, store_full
    (14, 4, 3, 1, 20260001, True),
    (28, 8, 2, 2, 20260002, True),
    (60, 256, 2, 2, 20260003, True),
    (60, 1 << 10, 2, 1, 20260004, False),
    (50, 1 << 12, 3, 2, 20260005, False),
    (60, 1 << 14, 2, 2, 20260006, False),
    (60, 1 << 14, 8, 2, 20260007, False),
    (60, 1 << 15, 4, 1, 20260008, False),
    (60, 1 << 16, 3, 1, 20260009, False),
]


def load_ref():
    R = ctypes.CDLL(oracle_py.REF_SO)
    R.ref_gen_params.argtypes = [u32, u32, u32, vp, vp]
    R.ref_ctx_create.restype = vp
    R.ref_ctx_create.argtypes = [u32, u32, vp, vp]
    R.ref_ctx_destroy.argtypes = [vp]
    R.ref_polys_load.restype = vp
    R.ref_polys_load.argtypes = [vp, vp, sz, ci]
    R.ref_polys_store.argtypes = [vp, vp]
    R.ref_polys_free.argtypes = [vp]
    R.ref_polys_op.restype = ctypes.c_double
    R.ref_polys_op.argtypes = [vp, vp, ci, ci]
    R.ref_polys_switch_format.restype = ctypes.c_double
    R.ref_polys_switch_format.argtypes = [vp, ci]
    return R


def digest(a):
    return hashlib.sha256(np.ascontiguousarray(a, np.uint64).tobytes()).hexdigest()


def main():
    R = load_ref()
    O = oracle_py.load()
    p = lambda a: a.ctypes.data_as(vp)
    out = []
    for nbits, n, towers, batch, seed, full in CASES:
        q = np.zeros(towers, np.uint64)
        psi = np.zeros(towers, np.uint64)
        R.ref_gen_params(nbits, n, towers, p(q), p(psi))
        a = O.fill_ring(batch, n, q, seed)
        b = O.fill_ring(batch, n, q, seed + 1000)
        ctx = R.ref_ctx_create(n, towers, p(q), p(psi))
        rec = {"nbits": nbits, "N": n, "towers": towers, "batch": batch, "seed": seed,
               "q": [int(v) for v in q], "psi": [int(v) for v in psi],
               "sha_a": digest(a), "sha_b": digest(b)}
        res = np.empty_like(a)
        for op, name in enumerate(["add", "sub", "mul"]):
            ha, hb = R.ref_polys_load(ctx, p(a), batch, 0), R.ref_polys_load(ctx, p(b), batch, 0)
            R.ref_polys_op(ha, hb, op, 1)
            R.ref_polys_store(ha, p(res))
            rec["sha_" + name] = digest(res)
            if full:
                rec[name] = res.reshape(-1).tolist()
            R.ref_polys_free(ha); R.ref_polys_free(hb)
        ha = R.ref_polys_load(ctx, p(a), batch, 1)          # COEFFICIENT -> forward
        R.ref_polys_switch_format(ha, 1)
        R.ref_polys_store(ha, p(res))
        rec["sha_fwd"] = digest(res)
        if full:
            rec["fwd"] = res.reshape(-1).tolist()
        R.ref_polys_free(ha)
        hb = R.ref_polys_load(ctx, p(b), batch, 0)          # EVALUATION -> inverse
        R.ref_polys_switch_format(hb, 1)
        R.ref_polys_store(hb, p(res))
        rec["sha_inv"] = digest(res)
        if full:
            rec["inv"] = res.reshape(-1).tolist()
        R.ref_polys_free(hb)
        R.ref_ctx_destroy(ctx)
        out.append(rec)
        print("case", nbits, n, towers, batch, "done")
    path = os.path.join(ROOT, "tests", "golden", "ref_vectors.json")
    with open(path, "w") as f:
        json.dump({"generator": "tests/golden/make_golden.py over oracle/_ref/libopenfhe_ref.so "
                                "(reference commit e70da11 compiled by oracle/Makefile)",
                   "cases": out}, f, indent=1)
    print("wrote", path)


if __name__ == "__main__":
    main()
