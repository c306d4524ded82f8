"""Generates tests/golden/ext_vectors.json from the REFERENCE itself (oracle/_ref/libopenfhe_ref.so): the
rows of SURVEY §8(f) beyond add / mul / NTT.  Section "automorphism": SHA-256 of
DCRTPoly::AutomorphismTransform(k) (dcrtpoly-impl.h:348-369) in both formats and of PrecomputeAutoMap's
table (nbtheory2.cpp:274-286), on inputs reproducible from (seed, q, index) with the oracle's generator.

    make -C oracle ref && python tests/golden/make_golden_ext.py
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
AUTO_CASES = [  # nbits, N, towers, batch, seed, ks
    (20, 8, 2, 1, 20261001, [3, 5, 15]),
    (60, 1 << 10, 2, 2, 20261002, [5, 25, 2047]),
    (60, 1 << 14, 3, 1, 20261003, [5, 3125, 32767]),
    (60, 1 << 16, 2, 1, 20261004, [5, 65537]),
]


def digest(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def main():
    R = ctypes.CDLL(oracle_py.REF_SO)
    R.ref_gen_params.argtypes = [u32, u32, u32, vp, vp]
    R.ref_ctx_create.restype = vp
    R.ref_ctx_create.argtypes = [u32, u32, vp, vp]
    R.ref_ctx_destroy.argtypes = [vp]
    R.ref_polys_load.restype = vp
    R.ref_polys_load.argtypes = [vp, vp, sz, ci]
    R.ref_polys_store.argtypes = [vp, vp]
    R.ref_polys_free.argtypes = [vp]
    R.ref_polys_automorphism.restype = vp
    R.ref_polys_automorphism.argtypes = [vp, u32, ci]
    R.ref_precompute_automap.argtypes = [u32, u32, vp]
    O = oracle_py.load()
    p = lambda a: a.ctypes.data_as(vp)
    auto = []
    for nbits, n, towers, batch, seed, ks in AUTO_CASES:
        q, psi = np.zeros(towers, np.uint64), np.zeros(towers, np.uint64)
        R.ref_gen_params(nbits, n, towers, p(q), p(psi))
        a = O.fill_ring(batch, n, q, seed)
        a[0, :, :2] = 0
        ctx = R.ref_ctx_create(n, towers, p(q), p(psi))
        rec = {"nbits": nbits, "N": n, "towers": towers, "batch": batch, "seed": seed, "q": [int(v) for v in q],
               "psi": [int(v) for v in psi], "sha_a": digest(a), "k": {}}
        res = np.empty_like(a)
        for k in ks:
            m = np.empty(n, np.uint32)
            R.ref_precompute_automap(n, k, p(m))
            e = {"sha_map": digest(m)}
            for fmt, name in ((0, "eval"), (1, "coef")):
                ha = R.ref_polys_load(ctx, p(a), batch, fmt)
                hr = R.ref_polys_automorphism(ha, k, 0)
                R.ref_polys_store(hr, p(res))
                e["sha_" + name] = digest(res)
                if n <= 8:
                    e[name] = res.reshape(-1).tolist()
                R.ref_polys_free(hr); R.ref_polys_free(ha)
            rec["k"][str(k)] = e
        R.ref_ctx_destroy(ctx)
        auto.append(rec)
        print("automorphism case", n, towers, "done")
    path = os.path.join(ROOT, "tests", "golden", "ext_vectors.json")
    doc = {}
    if os.path.exists(path):
        with open(path) as f:
            doc = json.load(f)
    doc["generator"] = ("tests/golden/make_golden_ext.py over oracle/_ref/libopenfhe_ref.so "
                        "(reference commit e70da11 compiled by oracle/Makefile)")
    doc["automorphism"] = auto
    with open(path, "w") as f:
        json.dump(doc, f, indent=1)
    print("wrote", path)


if __name__ == "__main__":
    main()
