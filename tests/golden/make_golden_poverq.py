"""Adds section "bfv_p_over_q" to tests/golden/ext_vectors.json from the REFERENCE itself (oracle/_ref/libopenfhe_ref.so):
SHA-256 of DCRTPoly::ExpandCRTBasisReverseOrder, FastExpandCRTBasisPloverQ (+ its two tables), ExpandCRTBasisQlHat and
ScaleAndRoundPOverQ (dcrtpoly-impl.h:1361-1468, 1514-1534, 2052-2066) on inputs reproducible from (seed, q, index).

    make -C oracle ref && python tests/golden/make_golden_poverq.py
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
CASES = [  # nbits, N, sizeQ, sizeQl (= sizePl), batch, seed
    (30, 16, 2, 2, 2, 20263001),
    (60, 1 << 10, 4, 3, 2, 20263002),
    (60, 1 << 13, 6, 6, 1, 20263003),
]


def digest(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def constants(O, q, seed):
    """one reduced constant per tower, reproducible from the seed"""
    return O.fill_ring(1, 1, np.ascontiguousarray(q, np.uint64), seed).reshape(-1).copy()


MULT_CASES = [  # nbits, N, sizeQ (= sizeR), batch, seed, t
    (30, 16, 2, 2, 20264001, 65537),
    (60, 1 << 10, 3, 2, 20264002, 65537),
    (60, 1 << 13, 5, 1, 20264003, 786433),
]


def mult_section(R, O, p):
    """Section "bfv_p_over_q_mult": the RNS core of BFVrns EvalMult with multTech == HPSPOVERQ run by the reference's own
    DCRTPoly (oracle/ref_driver.cpp: ref_bfv_hps_p_over_q_mult); the ScaleAndRound tables are stored for the GPU box."""
    R.ref_scale_tables_get.argtypes = [vp, vp, u64, vp, vp]
    R.ref_bfv_hps_p_over_q_mult.argtypes = [vp, vp, vp, vp, vp, vp, u64, ctypes.POINTER(vp)]
    out = []
    for nbits, n, sq, batch, seed, t in MULT_CASES:
        q, psi = np.zeros(2 * sq, np.uint64), np.zeros(2 * sq, np.uint64)
        R.ref_gen_params(nbits, n, 2 * sq, p(q), p(psi))
        mk = lambda a, b: R.ref_ctx_create(n, b - a, p(q[a:b].copy()), p(psi[a:b].copy()))
        cQ, cR, cQR = mk(0, sq), mk(sq, 2 * sq), mk(0, 2 * sq)
        frac, modq = np.empty(sq, np.float64), np.empty((sq, sq + 1), np.uint64)
        R.ref_scale_tables_get(cR, cQ, t, p(frac), p(modq))
        ins = [O.fill_ring(batch, n, q[:sq].copy(), seed + k) for k in range(4)]
        hs = [R.ref_polys_load(cQ, p(v), batch, 0) for v in ins]
        outs = (vp * 3)()
        assert R.ref_bfv_hps_p_over_q_mult(hs[0], hs[1], hs[2], hs[3], cR, cQR, t, outs) == 0
        rec = {"nbits": nbits, "N": n, "sizeQ": sq, "batch": batch, "seed": seed, "t": t, "q": [int(v) for v in q],
               "psi": [int(v) for v in psi], "frac_hex": [float(v).hex() for v in frac],
               "modq": [int(v) for v in modq.reshape(-1)], "sha_in": [digest(v) for v in ins], "sha_out": []}
        for k in range(3):
            res = np.empty((batch, sq, n), np.uint64)
            R.ref_polys_store(outs[k], p(res))
            rec["sha_out"].append(digest(res))
            R.ref_polys_free(outs[k])
        for h in hs:
            R.ref_polys_free(h)
        for c in (cQ, cR, cQR):
            R.ref_ctx_destroy(c)
        out.append(rec)
        print("p-over-q mult case", n, sq, "done")
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
    R.ref_polys_expand_crt_basis_reverse.argtypes = [vp, vp, vp, ci]
    R.ref_pl_over_q_tables_get.argtypes = [vp, vp, vp, vp]
    R.ref_polys_fast_expand_pl_over_q.argtypes = [vp, vp, vp, vp]
    R.ref_polys_expand_ql_hat.argtypes = [vp, vp, vp]
    R.ref_polys_scale_round_p_over_q.argtypes = [vp, vp, vp]
    O = oracle_py.load()
    p = lambda a: a.ctypes.data_as(vp)
    out = []
    for nbits, n, sq, sl, batch, seed in CASES:
        T = sq + sl
        aq, apsi = np.zeros(T, np.uint64), np.zeros(T, np.uint64)
        R.ref_gen_params(nbits, n, T, p(aq), p(apsi))
        mk = lambda qq, pp: R.ref_ctx_create(n, len(qq), p(np.ascontiguousarray(qq, np.uint64)), p(np.ascontiguousarray(pp, np.uint64)))
        q, psi, pl, plpsi = aq[:sq].copy(), apsi[:sq].copy(), aq[sq:].copy(), apsi[sq:].copy()
        cQ, cQl, cPl = mk(q, psi), mk(q[:sl], psi[:sl]), mk(pl, plpsi)
        cQlPl = mk(np.concatenate([q[:sl], pl]), np.concatenate([psi[:sl], plpsi]))
        cPQ = mk(np.concatenate([pl, q]), np.concatenate([plpsi, psi]))
        x = O.fill_ring(batch, n, q, seed)
        x[0, :, 0] = q - 1
        rec = {"nbits": nbits, "N": n, "sizeQ": sq, "sizeQl": sl, "batch": batch, "seed": seed,
               "q": [int(v) for v in aq], "psi": [int(v) for v in apsi], "sha_x": digest(x)}
        for fmt in (0, 1):
            for rf in (0, 1):
                hx = R.ref_polys_load(cQ, p(x), batch, fmt)
                assert R.ref_polys_expand_crt_basis_reverse(hx, cPl, cPQ, rf) == 0
                res = np.empty((batch, T, n), np.uint64)
                R.ref_polys_store(hx, p(res))
                rec[f"sha_expand_reverse_f{fmt}_r{rf}"] = digest(res)
                R.ref_polys_free(hx)
        neg, qi = np.empty(sq, np.uint64), np.empty((sq, sl), np.uint64)
        R.ref_pl_over_q_tables_get(cQ, cPl, p(neg), p(qi))
        rec["neg_pl_qhat_inv_modq"] = [int(v) for v in neg]
        rec["qinv_modp"] = [int(v) for v in qi.reshape(-1)]
        hx = R.ref_polys_load(cQ, p(x), batch, 1)
        assert R.ref_polys_fast_expand_pl_over_q(hx, cQl, cPl, cQlPl) == 0
        res = np.empty((batch, 2 * sl, n), np.uint64)
        R.ref_polys_store(hx, p(res))
        rec["sha_fast_expand_pl_over_q"] = digest(res)
        R.ref_polys_free(hx)
        c = constants(O, q[:sl], seed + 1)
        y = O.fill_ring(batch, n, q[:sl].copy(), seed + 2)
        hy = R.ref_polys_load(cQl, p(y), batch, 0)
        assert R.ref_polys_expand_ql_hat(hy, cQ, p(c)) == 0
        res = np.empty((batch, sq, n), np.uint64)
        R.ref_polys_store(hy, p(res))
        rec["sha_expand_ql_hat"] = digest(res)
        R.ref_polys_free(hy)
        # ScaleAndRoundPOverQ with the extra tower below (pl[0]) and above (the chain's first prime) the q_i
        for name, qs, ps, ex, ep in (("below", q, psi, pl[0], plpsi[0]), ("above", aq[1:sq + 1], apsi[1:sq + 1], aq[0], apsi[0])):
            full, fpsi = np.append(qs, ex).astype(np.uint64), np.append(ps, ep).astype(np.uint64)
            cq, cf = mk(qs, ps), mk(full, fpsi)
            k = constants(O, qs, seed + 3)
            z = O.fill_ring(batch, n, full, seed + 4)
            z[0, :, 0] = full - 1
            hz = R.ref_polys_load(cf, p(z), batch, 1)
            assert R.ref_polys_scale_round_p_over_q(hz, cq, p(k)) == 0
            res = np.empty((batch, sq, n), np.uint64)
            R.ref_polys_store(hz, p(res))
            rec["sha_scale_round_p_over_q_" + name] = digest(res)
            R.ref_polys_free(hz)
            R.ref_ctx_destroy(cq); R.ref_ctx_destroy(cf)
        for cc in (cQ, cQl, cPl, cQlPl, cPQ):
            R.ref_ctx_destroy(cc)
        out.append(rec)
        print("p-over-q case", n, sq, sl, "done")
    path = os.path.join(ROOT, "tests", "golden", "ext_vectors.json")
    with open(path) as f:
        doc = json.load(f)
    doc["bfv_p_over_q"] = out
    doc["bfv_p_over_q_mult"] = mult_section(R, O, p)
    doc["bfv_p_over_q_generator"] = "tests/golden/make_golden_poverq.py over oracle/_ref/libopenfhe_ref.so (reference commit e70da11)"
    with open(path, "w") as f:
        json.dump(doc, f, indent=1)
    print("wrote", path)


if __name__ == "__main__":
    main()
