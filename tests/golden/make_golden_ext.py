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


BASIS_CASES = [  # nbits, N, sizeQ, sizeP, batch, seed
    (30, 16, 2, 1, 2, 20262001),
    (60, 1 << 10, 4, 2, 2, 20262002),
    (60, 1 << 14, 5, 3, 1, 20262003),
]


def basis_section(R, O, p):
    """Section "basis": digests of ApproxSwitchCRTBasis, ApproxModUp, ApproxModDown (t = 0 and 65537),
    DropLastElementAndScale and ModReduce (dcrtpoly-impl.h:746-812, 1034-1175) run by the reference, tables built
    with its BigInteger as the schemes do (oracle/ref_driver.cpp)."""
    R.ref_polys_approx_switch.restype = vp
    R.ref_polys_approx_switch.argtypes = [vp, vp]
    R.ref_polys_approx_mod_up.argtypes = [vp, vp, vp]
    R.ref_polys_approx_mod_down.restype = vp
    R.ref_polys_approx_mod_down.argtypes = [vp, vp, vp, u64]
    R.ref_polys_rescale.argtypes = [vp, vp, u64]
    out = []
    for nbits, n, sq, sp, batch, seed in BASIS_CASES:
        q, psi = np.zeros(sq + sp, np.uint64), np.zeros(sq + sp, np.uint64)
        R.ref_gen_params(nbits, n, sq + sp, p(q), p(psi))
        mk = lambda a, b: R.ref_ctx_create(n, b - a, p(q[a:b].copy()), p(psi[a:b].copy()))
        cQ, cP, cQP, cR = mk(0, sq), mk(sq, sq + sp), mk(0, sq + sp), mk(0, sq + sp - 1)
        x = O.fill_ring(batch, n, q[:sq].copy(), seed)
        y = O.fill_ring(batch, n, q, seed + 1)
        rec = {"nbits": nbits, "N": n, "sizeQ": sq, "sizeP": sp, "batch": batch, "seed": seed,
               "q": [int(v) for v in q], "psi": [int(v) for v in psi], "sha_x": digest(x), "sha_y": digest(y)}
        hx = R.ref_polys_load(cQ, p(x), batch, 1)
        hr = R.ref_polys_approx_switch(hx, cP)
        res = np.empty((batch, sp, n), np.uint64)
        R.ref_polys_store(hr, p(res))
        rec["sha_switch"] = digest(res)
        R.ref_polys_free(hr); R.ref_polys_free(hx)
        for fmt in (0, 1):
            hx = R.ref_polys_load(cQ, p(x), batch, fmt)
            assert R.ref_polys_approx_mod_up(hx, cP, cQP) == 0
            res = np.empty((batch, sq + sp, n), np.uint64)
            R.ref_polys_store(hx, p(res))
            rec[f"sha_modup_f{fmt}"] = digest(res)
            R.ref_polys_free(hx)
        for t in (0, 65537):
            hy = R.ref_polys_load(cQP, p(y), batch, 0)
            hr = R.ref_polys_approx_mod_down(hy, cQ, cP, t)
            res = np.empty((batch, sq, n), np.uint64)
            R.ref_polys_store(hr, p(res))
            rec[f"sha_moddown_t{t}"] = digest(res)
            R.ref_polys_free(hr); R.ref_polys_free(hy)
            for fmt in (0, 1):
                hy = R.ref_polys_load(cQP, p(y), batch, fmt)
                assert R.ref_polys_rescale(hy, cR, t) == 0
                res = np.empty((batch, sq + sp - 1, n), np.uint64)
                R.ref_polys_store(hy, p(res))
                rec[f"sha_rescale_t{t}_f{fmt}"] = digest(res)
                R.ref_polys_free(hy)
        for c in (cQ, cP, cQP, cR):
            R.ref_ctx_destroy(c)
        out.append(rec)
        print("basis case", n, sq, sp, "done")
    return out


BFV_CASES = [  # nbits, N, sizeQ, sizeP (= R), batch, seed, t
    (30, 16, 2, 3, 2, 20263001, 65537),
    (60, 1 << 10, 4, 5, 2, 20263002, 65537),
    (60, 1 << 13, 6, 7, 1, 20263003, 786433),
    (60, 16, 80, 81, 1, 20263004, 65537),          # nu >= 2^64: the 128-bit alpha path of ScaleAndRound
]


def bfv_section(R, O, p):
    """Section "bfv": SwitchCRTBasis, ExpandCRTBasis (all format pairs) and ScaleAndRound (QR -> R) run by the reference
    (dcrtpoly-impl.h:1178-1358, 1876-1960) with tables built as bfvrns-cryptoparameters.cpp:262-350 does; the
    ScaleAndRound tables are stored so that the GPU box can feed the same ones."""
    R.ref_polys_switch_crt_basis.restype = vp
    R.ref_polys_switch_crt_basis.argtypes = [vp, vp]
    R.ref_polys_expand_crt_basis.argtypes = [vp, vp, vp, ci]
    R.ref_scale_tables_get.argtypes = [vp, vp, u64, vp, vp]
    R.ref_polys_scale_and_round.restype = vp
    R.ref_polys_scale_and_round.argtypes = [vp, vp, vp, u64]
    R.ref_bfv_hps_mult.argtypes = [vp, vp, vp, vp, vp, vp, u64, ctypes.POINTER(vp)]
    out = []
    for nbits, n, sq, sp, batch, seed, t in BFV_CASES:
        q, psi = np.zeros(sq + sp, np.uint64), np.zeros(sq + sp, np.uint64)
        R.ref_gen_params(nbits, n, sq + sp, p(q), p(psi))
        mk = lambda a, b: R.ref_ctx_create(n, b - a, p(q[a:b].copy()), p(psi[a:b].copy()))
        cQ, cP, cQP = mk(0, sq), mk(sq, sq + sp), mk(0, sq + sp)
        x = O.fill_ring(batch, n, q[:sq].copy(), seed)
        y = O.fill_ring(batch, n, q, seed + 1)
        frac, modr = np.empty(sq, np.float64), np.empty((sp, sq + 1), np.uint64)
        R.ref_scale_tables_get(cQ, cP, t, p(frac), p(modr))
        rec = {"nbits": nbits, "N": n, "sizeQ": sq, "sizeP": sp, "batch": batch, "seed": seed, "t": t,
               "q": [int(v) for v in q], "psi": [int(v) for v in psi], "sha_x": digest(x), "sha_y": digest(y),
               "frac_hex": [float(v).hex() for v in frac], "modr": [int(v) for v in modr.reshape(-1)]}
        hx = R.ref_polys_load(cQ, p(x), batch, 1)
        hr = R.ref_polys_switch_crt_basis(hx, cP)
        res = np.empty((batch, sp, n), np.uint64)
        R.ref_polys_store(hr, p(res))
        rec["sha_switch_exact"] = digest(res)
        R.ref_polys_free(hr); R.ref_polys_free(hx)
        for fmt in (0, 1):
            for rf in (0, 1):
                hx = R.ref_polys_load(cQ, p(x), batch, fmt)
                assert R.ref_polys_expand_crt_basis(hx, cP, cQP, rf) == 0
                res = np.empty((batch, sq + sp, n), np.uint64)
                R.ref_polys_store(hx, p(res))
                rec[f"sha_expand_f{fmt}_r{rf}"] = digest(res)
                R.ref_polys_free(hx)
        hy = R.ref_polys_load(cQP, p(y), batch, 1)
        hr = R.ref_polys_scale_and_round(hy, cQ, cP, t)
        res = np.empty((batch, sp, n), np.uint64)
        R.ref_polys_store(hr, p(res))
        rec["sha_scale_round"] = digest(res)
        R.ref_polys_free(hr); R.ref_polys_free(hy)
        if sq <= 8:      # the composed EvalMult core (bfvrns-leveledshe.cpp:193-207, 322-356) on four seeded inputs
            ins = [O.fill_ring(batch, n, q[:sq].copy(), seed + 10 + k) for k in range(4)]
            hs = [R.ref_polys_load(cQ, p(v), batch, 0) for v in ins]
            outs = (vp * 3)()
            assert R.ref_bfv_hps_mult(hs[0], hs[1], hs[2], hs[3], cP, cQP, t, outs) == 0
            rec["sha_mult_in"] = [digest(v) for v in ins]
            rec["sha_mult_out"] = []
            for k in range(3):
                res = np.empty((batch, sq, n), np.uint64)
                R.ref_polys_store(outs[k], p(res))
                rec["sha_mult_out"].append(digest(res))
                R.ref_polys_free(outs[k])
            for h in hs:
                R.ref_polys_free(h)
        for c in (cQ, cP, cQP):
            R.ref_ctx_destroy(c)
        out.append(rec)
        print("bfv case", n, sq, sp, "done")
    return out


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
    basis = basis_section(R, O, p)
    bfv = bfv_section(R, O, p)
    path = os.path.join(ROOT, "tests", "golden", "ext_vectors.json")
    doc = {}
    if os.path.exists(path):
        with open(path) as f:
            doc = json.load(f)
    doc["generator"] = ("tests/golden/make_golden_ext.py over oracle/_ref/libopenfhe_ref.so "
                        "(reference commit e70da11 compiled by oracle/Makefile)")
    doc["automorphism"] = auto
    doc["basis"] = basis
    doc["bfv"] = bfv
    with open(path, "w") as f:
        json.dump(doc, f, indent=1)
    print("wrote", path)


if __name__ == "__main__":
    main()
