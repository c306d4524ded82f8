"""Oracle restatement vs the reference's own compiled sources (oracle/_ref), on fresh random inputs.
Only where oracle/_ref was built (the container that has /root/reference); elsewhere the committed
golden vectors (tests/test_oracle_golden.py) carry the same evidence."""
import ctypes
import os

import numpy as np
import pytest

from tests import oracle_py

pytestmark = pytest.mark.skipif(not os.path.exists(oracle_py.REF_SO), reason="oracle/_ref not built here")
vp, sz, ci, u32, u64 = ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int, ctypes.c_uint32, ctypes.c_uint64


@pytest.fixture(scope="module")
def ref():
    R = ctypes.CDLL(oracle_py.REF_SO)
    R.ref_gen_params.argtypes = [u32, u32, u32, vp, vp]
    R.ref_root_of_unity.restype = u64
    R.ref_root_of_unity.argtypes = [u32, u64]
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
    R.ref_polys_automorphism.restype = vp
    R.ref_polys_automorphism.argtypes = [vp, u32, ci]
    R.ref_precompute_automap.argtypes = [u32, u32, vp]
    R.ref_last_error.restype = ctypes.c_char_p
    R.ref_polys_approx_switch.restype = vp
    R.ref_polys_approx_switch.argtypes = [vp, vp]
    R.ref_polys_approx_mod_up.argtypes = [vp, vp, vp]
    R.ref_polys_approx_mod_down.restype = vp
    R.ref_polys_approx_mod_down.argtypes = [vp, vp, vp, u64]
    R.ref_polys_rescale.argtypes = [vp, vp, u64]
    R.ref_switch_tables_get.argtypes = [vp, vp, vp, vp]
    R.ref_rescale_tables_get.argtypes = [vp, vp, vp]
    R.ref_polys_switch_crt_basis.restype = vp
    R.ref_polys_switch_crt_basis.argtypes = [vp, vp]
    R.ref_polys_expand_crt_basis.argtypes = [vp, vp, vp, ci]
    R.ref_scale_tables_get.argtypes = [vp, vp, u64, vp, vp]
    R.ref_polys_scale_and_round.restype = vp
    R.ref_polys_scale_and_round.argtypes = [vp, vp, vp, u64]
    R.ref_polys_expand_crt_basis_reverse.argtypes = [vp, vp, vp, ci]
    R.ref_pl_over_q_tables_get.argtypes = [vp, vp, vp, vp]
    R.ref_polys_fast_expand_pl_over_q.argtypes = [vp, vp, vp, vp]
    R.ref_polys_expand_ql_hat.argtypes = [vp, vp, vp]
    R.ref_polys_scale_round_p_over_q.argtypes = [vp, vp, vp]
    R.ref_polys_times_towers.restype = vp
    R.ref_polys_times_towers.argtypes = [vp, vp]
    R.ref_bfv_hps_p_over_q_mult.argtypes = [vp, vp, vp, vp, vp, vp, u64, ctypes.POINTER(vp)]
    R.ref_bfv_hps_mult.argtypes = [vp, vp, vp, vp, vp, vp, u64, ctypes.POINTER(vp)]
    return R


def p(a):
    return a.ctypes.data_as(vp)


@pytest.mark.parametrize("nbits,n,towers", [(20, 16, 2), (40, 512, 3), (60, 1 << 13, 2), (60, 1 << 14, 3), (55, 1 << 16, 2)])
def test_params_and_ops_match_reference(oracle, ref, nbits, n, towers):
    q, psi = oracle.gen_params(nbits, n, towers)
    qr, pr = np.zeros(towers, np.uint64), np.zeros(towers, np.uint64)
    ref.ref_gen_params(nbits, n, towers, p(qr), p(pr))
    assert np.array_equal(q, qr) and np.array_equal(psi, pr)
    rng = np.random.default_rng(n + nbits)
    batch = 2
    a = np.stack([rng.integers(0, int(m), size=(batch, n), dtype=np.uint64) for m in q], axis=1)
    b = np.stack([rng.integers(0, int(m), size=(batch, n), dtype=np.uint64) for m in q], axis=1)
    a[0, :, 0], b[0, :, 0] = q - 1, q - 1
    octx = oracle.ctx(n, q, psi)
    ctx = ref.ref_ctx_create(n, towers, p(q), p(psi))
    res = np.empty_like(a)
    for op in range(3):
        ha, hb = ref.ref_polys_load(ctx, p(a), batch, 0), ref.ref_polys_load(ctx, p(b), batch, 0)
        # in-place forms only: the fork's out-of-place Plus/Times copy NativeVectors, whose PimData
        # member is copied shallowly (pim/PimData.h:36-44) and double-frees under the dpu.h stub
        ref.ref_polys_op(ha, hb, op, 1)                      # += / -= / *=
        ref.ref_polys_store(ha, p(res))
        assert np.array_equal(res, octx.elementwise(op, a, b)), f"in-place op {op}"
        for h in (ha, hb):
            ref.ref_polys_free(h)
    ha = ref.ref_polys_load(ctx, p(a), batch, 1)
    ref.ref_polys_switch_format(ha, 1)
    ref.ref_polys_store(ha, p(res))
    assert np.array_equal(res, octx.ntt(a, inverse=False))
    ref.ref_polys_switch_format(ha, 1)
    ref.ref_polys_store(ha, p(res))
    assert np.array_equal(res, a)
    ref.ref_polys_free(ha)
    hb = ref.ref_polys_load(ctx, p(b), batch, 0)
    ref.ref_polys_switch_format(hb, 1)
    ref.ref_polys_store(hb, p(res))
    assert np.array_equal(res, octx.ntt(b, inverse=True))
    ref.ref_polys_free(hb)
    ref.ref_ctx_destroy(ctx)


@pytest.mark.parametrize("n,towers", [(8, 2), (1 << 10, 3), (1 << 14, 2)])
def test_automorphism_matches_reference(oracle, ref, n, towers):
    """DCRTPoly::AutomorphismTransform(k) and (k, precomp), both formats."""
    q, psi = oracle.gen_params(60, n, towers)
    rng = np.random.default_rng(n)
    batch = 2
    a = np.stack([rng.integers(0, int(m), size=(batch, n), dtype=np.uint64) for m in q], axis=1)
    a[0, :, :3] = 0                                     # negated zeros: the reference stores q
    octx = oracle.ctx(n, q, psi)
    ctx = ref.ref_ctx_create(n, towers, p(q), p(psi))
    res = np.empty_like(a)
    for k in (3, 5, 2 * n - 1, n + 1, 25 % (2 * n) | 1):
        mr = np.empty(n, np.uint32)
        ref.ref_precompute_automap(n, k, p(mr))
        assert np.array_equal(mr, oracle.automap(n, k))
        for fmt in (0, 1):
            ha = ref.ref_polys_load(ctx, p(a), batch, fmt)
            hr = ref.ref_polys_automorphism(ha, k, 0)
            assert hr, ref.ref_last_error()
            ref.ref_polys_store(hr, p(res))
            assert np.array_equal(res, octx.automorphism(a, k, fmt)), (k, fmt)
            ref.ref_polys_free(hr)
            if fmt == 0:
                hr = ref.ref_polys_automorphism(ha, k, 1)
                assert hr, ref.ref_last_error()
                ref.ref_polys_store(hr, p(res))
                assert np.array_equal(res, octx.automorphism(a, k, 0)), (k, "precomp")
                ref.ref_polys_free(hr)
            ref.ref_polys_free(ha)
    ha = ref.ref_polys_load(ctx, p(a), batch, 0)
    assert not ref.ref_polys_automorphism(ha, 4, 0) and b"not odd" in ref.ref_last_error()
    ref.ref_polys_free(ha)
    ref.ref_ctx_destroy(ctx)


@pytest.mark.parametrize("nbits,n,sq,sp", [(30, 16, 2, 1), (60, 64, 3, 2), (60, 1 << 10, 5, 3), (45, 1 << 12, 2, 4)])
def test_basis_conversion_matches_reference(oracle, ref, nbits, n, sq, sp):
    """ApproxSwitchCRTBasis, ApproxModUp, ApproxModDown (CKKS and BGV forms),
    DropLastElementAndScale and ModReduce of the reference's DCRTPoly vs the restatement, and the tables
    against the BigInteger values the schemes' parameter code computes."""
    q, psi = oracle.gen_params(nbits, n, sq + sp)
    ob = oracle.basis(n, q, psi, sq)
    mk = lambda a, b: ref.ref_ctx_create(n, b - a, p(q[a:b].copy()), p(psi[a:b].copy()))
    cQ, cP, cQP, cR = mk(0, sq), mk(sq, sq + sp), mk(0, sq + sp), mk(0, sq + sp - 1)
    B = 2
    for rev, (ca, cb) in ((False, (cQ, cP)), (True, (cP, cQ))):
        hi, hm = ob.tables(rev)
        rhi, rhm = np.empty_like(hi), np.empty_like(hm)
        ref.ref_switch_tables_get(ca, cb, p(rhi), p(rhm))
        assert np.array_equal(hi, rhi) and np.array_equal(hm, rhm)
    a, b = oracle.rescale_tables(q)
    ra, rb = np.empty_like(a), np.empty_like(b)
    ref.ref_rescale_tables_get(cQP, p(ra), p(rb))
    assert np.array_equal(a, ra) and np.array_equal(b, rb)

    x = oracle.fill_ring(B, n, q[:sq].copy(), 11)
    x[0, :, 0] = q[:sq] - 1
    x[0, :, 1] = 0
    hx = ref.ref_polys_load(cQ, p(x), B, 1)
    hr = ref.ref_polys_approx_switch(hx, cP)
    res = np.empty((B, sp, n), np.uint64)
    ref.ref_polys_store(hr, p(res))
    assert np.array_equal(res, ob.approx_switch(x))
    ref.ref_polys_free(hr); ref.ref_polys_free(hx)
    for fmt in (0, 1):
        hx = ref.ref_polys_load(cQ, p(x), B, fmt)
        assert ref.ref_polys_approx_mod_up(hx, cP, cQP) == 0, ref.ref_last_error()
        res = np.empty((B, sq + sp, n), np.uint64)
        ref.ref_polys_store(hx, p(res))
        assert np.array_equal(res, ob.mod_up(x, fmt)), f"ApproxModUp format {fmt}"
        ref.ref_polys_free(hx)
    y = oracle.fill_ring(B, n, q, 12)
    y[0, :, 0] = q - 1
    for t in (0, 65537, 2):
        hy = ref.ref_polys_load(cQP, p(y), B, 0)
        hr = ref.ref_polys_approx_mod_down(hy, cQ, cP, t)
        assert hr, ref.ref_last_error()
        res = np.empty((B, sq, n), np.uint64)
        ref.ref_polys_store(hr, p(res))
        assert np.array_equal(res, ob.mod_down(y, t)), f"ApproxModDown t={t}"
        ref.ref_polys_free(hr); ref.ref_polys_free(hy)
        for fmt in (0, 1):
            hy = ref.ref_polys_load(cQP, p(y), B, fmt)
            assert ref.ref_polys_rescale(hy, cR, t) == 0, ref.ref_last_error()
            res = np.empty((B, sq + sp - 1, n), np.uint64)
            ref.ref_polys_store(hy, p(res))
            assert np.array_equal(res, oracle.rescale(n, q, psi, y, fmt, t)), f"rescale t={t} format {fmt}"
            ref.ref_polys_free(hy)
    for c in (cQ, cP, cQP, cR):
        ref.ref_ctx_destroy(c)


@pytest.mark.parametrize("nbits,n,sq,sp,t", [(30, 16, 2, 3, 65537), (60, 64, 3, 4, 65537), (60, 1 << 10, 5, 6, 786433),
                                               (55, 1 << 11, 20, 21, 65537), (60, 16, 80, 81, 65537)])
def test_bfv_hps_set_matches_reference(oracle, ref, nbits, n, sq, sp, t):
    """SURVEY §8(f) rank 3, BFV half: SwitchCRTBasis, ExpandCRTBasis and ScaleAndRound of the reference's DCRTPoly
    (IEEE-double overflow count included) vs the restatement; tables as bfvrns-cryptoparameters.cpp builds them."""
    q, psi = oracle.gen_params(nbits, n, sq + sp)
    ob = oracle.basis(n, q, psi, sq)
    mk = lambda a, b: ref.ref_ctx_create(n, b - a, p(q[a:b].copy()), p(psi[a:b].copy()))
    cQ, cP, cQP = mk(0, sq), mk(sq, sq + sp), mk(0, sq + sp)
    B = 2
    x = oracle.fill_ring(B, n, q[:sq].copy(), 61)
    x[0, :, 0] = q[:sq] - 1
    x[0, :, 1] = 0
    hx = ref.ref_polys_load(cQ, p(x), B, 1)
    hr = ref.ref_polys_switch_crt_basis(hx, cP)
    res = np.empty((B, sp, n), np.uint64)
    ref.ref_polys_store(hr, p(res))
    assert np.array_equal(res, ob.switch_exact(x))
    ref.ref_polys_free(hr); ref.ref_polys_free(hx)
    for fmt in (0, 1):
        for rf in (0, 1):
            hx = ref.ref_polys_load(cQ, p(x), B, fmt)
            assert ref.ref_polys_expand_crt_basis(hx, cP, cQP, rf) == 0, ref.ref_last_error()
            res = np.empty((B, sq + sp, n), np.uint64)
            ref.ref_polys_store(hx, p(res))
            assert np.array_equal(res, ob.expand(x, fmt, rf)), f"ExpandCRTBasis {fmt}->{rf}"
            ref.ref_polys_free(hx)
    frac, modr = np.empty(sq, np.float64), np.empty((sp, sq + 1), np.uint64)
    ref.ref_scale_tables_get(cQ, cP, t, p(frac), p(modr))
    y = oracle.fill_ring(B, n, q, 62)
    y[0, :, 0] = q - 1
    hy = ref.ref_polys_load(cQP, p(y), B, 1)
    hr = ref.ref_polys_scale_and_round(hy, cQ, cP, t)
    res = np.empty((B, sp, n), np.uint64)
    ref.ref_polys_store(hr, p(res))
    assert np.array_equal(res, ob.scale_and_round(y, frac, modr, 1))
    ref.ref_polys_free(hr); ref.ref_polys_free(hy)
    for c in (cQ, cP, cQP):
        ref.ref_ctx_destroy(c)


@pytest.mark.parametrize("nbits,n,sq,sr,t", [(30, 16, 2, 3, 65537), (60, 256, 3, 4, 65537), (60, 1 << 11, 4, 5, 786433)])
def test_bfv_hps_mult_core_matches_reference(oracle, ref, nbits, n, sq, sr, t):
    """The RNS core of BFVrns EvalMult (HPS), bfvrns-leveledshe.cpp:193-207, 322-356, executed by the reference's own
    DCRTPoly vs the composition of the restated pieces."""
    q, psi = oracle.gen_params(nbits, n, sq + sr)
    ob = oracle.basis(n, q, psi, sq)
    mk = lambda a, b: ref.ref_ctx_create(n, b - a, p(q[a:b].copy()), p(psi[a:b].copy()))
    cQ, cR, cQR = mk(0, sq), mk(sq, sq + sr), mk(0, sq + sr)
    B = 2
    ins = [oracle.fill_ring(B, n, q[:sq].copy(), 90 + k) for k in range(4)]
    hs = [ref.ref_polys_load(cQ, p(v), B, 0) for v in ins]
    outs = (vp * 3)()
    assert ref.ref_bfv_hps_mult(hs[0], hs[1], hs[2], hs[3], cR, cQR, t, outs) == 0, ref.ref_last_error()
    frac, modr = np.empty(sq, np.float64), np.empty((sr, sq + 1), np.uint64)
    ref.ref_scale_tables_get(cQ, cR, t, p(frac), p(modr))
    want = ob.bfv_hps_mult(ins[0], ins[1], ins[2], ins[3], frac, modr)
    for k in range(3):
        res = np.empty((B, sq, n), np.uint64)
        ref.ref_polys_store(outs[k], p(res))
        assert np.array_equal(res, want[k]), f"c{k}"
        ref.ref_polys_free(outs[k])
    for h in hs:
        ref.ref_polys_free(h)
    for c in (cQ, cR, cQR):
        ref.ref_ctx_destroy(c)


@pytest.mark.parametrize("nbits,n,sq,sp", [(30, 16, 2, 3), (60, 256, 3, 2), (60, 1 << 11, 4, 5)])
def test_expand_reverse_order_matches_reference(oracle, ref, nbits, n, sq, sp):
    """DCRTPolyImpl::ExpandCRTBasisReverseOrder (dcrtpoly-impl.h:1361-1410) on the reference's own DCRTPoly."""
    q, psi = oracle.gen_params(nbits, n, sq + sp)
    ob = oracle.basis(n, q, psi, sq)
    pq, ppsi = np.concatenate([q[sq:], q[:sq]]), np.concatenate([psi[sq:], psi[:sq]])
    cQ = ref.ref_ctx_create(n, sq, p(q[:sq].copy()), p(psi[:sq].copy()))
    cP = ref.ref_ctx_create(n, sp, p(q[sq:].copy()), p(psi[sq:].copy()))
    cPQ = ref.ref_ctx_create(n, sq + sp, p(pq), p(ppsi))
    B = 2
    x = oracle.fill_ring(B, n, q[:sq].copy(), 71)
    x[0, :, 0] = q[:sq] - 1
    for fmt in (0, 1):
        for rf in (0, 1):
            hx = ref.ref_polys_load(cQ, p(x), B, fmt)
            assert ref.ref_polys_expand_crt_basis_reverse(hx, cP, cPQ, rf) == 0, ref.ref_last_error()
            res = np.empty((B, sq + sp, n), np.uint64)
            ref.ref_polys_store(hx, p(res))
            assert np.array_equal(res, ob.expand_reverse(x, fmt, rf)), f"ExpandCRTBasisReverseOrder {fmt}->{rf}"
            ref.ref_polys_free(hx)
    for c in (cQ, cP, cPQ):
        ref.ref_ctx_destroy(c)


@pytest.mark.parametrize("nbits,n,sq,sl", [(30, 16, 2, 2), (60, 256, 3, 2), (60, 1 << 11, 5, 5), (45, 64, 4, 1)])
def test_hps_p_over_q_set_matches_reference(oracle, ref, nbits, n, sq, sl):
    """FastExpandCRTBasisPloverQ (tables and values), ExpandCRTBasisQlHat and ScaleAndRoundPOverQ
    (dcrtpoly-impl.h:1413-1468, 1514-1534, 2052-2066) on the reference's own DCRTPoly."""
    allq, allpsi = oracle.gen_params(nbits, n, sq + sl)
    q, psi, pl, plpsi = allq[:sq].copy(), allpsi[:sq].copy(), allq[sq:].copy(), allpsi[sq:].copy()
    ql, qlpsi = q[:sl].copy(), psi[:sl].copy()
    cQ = ref.ref_ctx_create(n, sq, p(q), p(psi))
    cQl = ref.ref_ctx_create(n, sl, p(ql), p(qlpsi))
    cPl = ref.ref_ctx_create(n, sl, p(pl), p(plpsi))
    qlpl, qlplpsi = np.concatenate([ql, pl]), np.concatenate([qlpsi, plpsi])
    cQlPl = ref.ref_ctx_create(n, 2 * sl, p(qlpl), p(qlplpsi))
    neg, qi = np.empty(sq, np.uint64), np.empty((sq, sl), np.uint64)
    ref.ref_pl_over_q_tables_get(cQ, cPl, p(neg), p(qi))
    oneg, oqi = oracle.pl_over_q_tables(q, pl)
    assert np.array_equal(neg, oneg) and np.array_equal(qi, oqi)
    B = 2
    x = oracle.fill_ring(B, n, q, 81)
    x[0, :, 0] = q - 1
    x[0, :, 1] = 0
    hx = ref.ref_polys_load(cQ, p(x), B, 1)
    assert ref.ref_polys_fast_expand_pl_over_q(hx, cQl, cPl, cQlPl) == 0, ref.ref_last_error()
    res = np.empty((B, 2 * sl, n), np.uint64)
    ref.ref_polys_store(hx, p(res))
    assert np.array_equal(res, oracle.fast_expand_pl_over_q(x, n, q, ql, pl))
    ref.ref_polys_free(hx)
    # ExpandCRTBasisQlHat: Ql -> Q with arbitrary reduced constants
    rng = np.random.default_rng(5)
    c = np.array([int(rng.integers(0, int(ql[i]))) for i in range(sl)], np.uint64)
    y = oracle.fill_ring(B, n, ql, 82)
    y[0, :, 0] = ql - 1
    hy = ref.ref_polys_load(cQl, p(y), B, 0)
    assert ref.ref_polys_expand_ql_hat(hy, cQ, p(c)) == 0, ref.ref_last_error()
    res = np.empty((B, sq, n), np.uint64)
    ref.ref_polys_store(hy, p(res))
    assert np.array_equal(res, oracle.expand_ql_hat(y, n, q, sl, c))
    ref.ref_polys_free(hy)
    # ScaleAndRoundPOverQ: the extra tower smaller than every q_i (the chain's next prime) and larger (the chain's first)
    for qs, ps, extra, epsi in ((q, psi, pl[0], plpsi[0]), (allq[1:sq + 1].copy(), allpsi[1:sq + 1].copy(), allq[0], allpsi[0])):
        full, fpsi = np.append(qs, extra).astype(np.uint64), np.append(ps, epsi).astype(np.uint64)
        cq = ref.ref_ctx_create(n, sq, p(qs), p(ps))
        cf = ref.ref_ctx_create(n, sq + 1, p(full), p(fpsi))
        k = np.array([int(rng.integers(0, int(qs[i]))) for i in range(sq)], np.uint64)
        z = oracle.fill_ring(B, n, full, 83)
        z[0, :, 0] = full - 1
        z[0, :, 1] = 0
        hz = ref.ref_polys_load(cf, p(z), B, 1)
        assert ref.ref_polys_scale_round_p_over_q(hz, cq, p(k)) == 0, ref.ref_last_error()
        res = np.empty((B, sq, n), np.uint64)
        ref.ref_polys_store(hz, p(res))
        assert np.array_equal(res, oracle.scale_round_p_over_q(z, n, qs, k))
        ref.ref_polys_free(hz)
        ref.ref_ctx_destroy(cq); ref.ref_ctx_destroy(cf)
    for cc in (cQ, cQl, cPl, cQlPl):
        ref.ref_ctx_destroy(cc)


@pytest.mark.parametrize("nbits,n,sq,t", [(30, 16, 2, 65537), (60, 256, 3, 65537), (60, 1 << 11, 4, 786433)])
def test_bfv_hps_p_over_q_mult_core_matches_reference(oracle, ref, nbits, n, sq, t):
    """The RNS core of BFVrns EvalMult (HPSPOVERQ), bfvrns-leveledshe.cpp:208-235, 322-334, 357-366, executed by the
    reference's own DCRTPoly vs the composition of the restated pieces; the t*Q/(QR) tables are the reference's
    recipe with Q and R exchanged."""
    q, psi = oracle.gen_params(nbits, n, 2 * sq)
    ob = oracle.basis(n, q, psi, sq)
    mk = lambda a, b: ref.ref_ctx_create(n, b - a, p(q[a:b].copy()), p(psi[a:b].copy()))
    cQ, cR, cQR = mk(0, sq), mk(sq, 2 * sq), mk(0, 2 * sq)
    B = 2
    ins = [oracle.fill_ring(B, n, q[:sq].copy(), 190 + k) for k in range(4)]
    hs = [ref.ref_polys_load(cQ, p(v), B, 0) for v in ins]
    outs = (vp * 3)()
    assert ref.ref_bfv_hps_p_over_q_mult(hs[0], hs[1], hs[2], hs[3], cR, cQR, t, outs) == 0, ref.ref_last_error()
    frac, modq = np.empty(sq, np.float64), np.empty((sq, sq + 1), np.uint64)
    ref.ref_scale_tables_get(cR, cQ, t, p(frac), p(modq))
    want = ob.bfv_hps_p_over_q_mult(ins[0], ins[1], ins[2], ins[3], frac, modq)
    for k in range(3):
        res = np.empty((B, sq, n), np.uint64)
        ref.ref_polys_store(outs[k], p(res))
        assert np.array_equal(res, want[k]), f"c{k}"
        ref.ref_polys_free(outs[k])
    for h in hs:
        ref.ref_polys_free(h)
    for c in (cQ, cR, cQR):
        ref.ref_ctx_destroy(c)


def test_times_tower_constants_matches_reference(oracle, ref):
    """DCRTPolyImpl::Times(const std::vector<NativeInteger>&) (dcrtpoly-impl.h:617-627) vs the restated per-tower product."""
    n, towers, B = 256, 4, 2
    q, psi = oracle.gen_params(60, n, towers)
    c = ref.ref_ctx_create(n, towers, p(q), p(psi))
    k = oracle.fill_ring(1, 1, q.copy(), 5).reshape(-1).copy()
    x = oracle.fill_ring(B, n, q, 6)
    x[0, :, 0] = q - 1
    hx = ref.ref_polys_load(c, p(x), B, 0)
    hr = ref.ref_polys_times_towers(hx, p(k))
    res = np.empty_like(x)
    ref.ref_polys_store(hr, p(res))
    assert np.array_equal(res, oracle.expand_ql_hat(x, n, q, towers, k))
    ref.ref_polys_free(hr); ref.ref_polys_free(hx); ref.ref_ctx_destroy(c)


@pytest.mark.parametrize("n,size_ql,size_p,alpha,t,nbits", [
    (64, 6, 2, 2, 0, 60), (64, 5, 2, 2, 0, 60), (128, 4, 3, 3, 0, 45), (64, 7, 3, 3, 65537, 55), (32, 3, 1, 1, 0, 30),
])
def test_hybrid_key_switching_matches_reference(oracle, ref, n, size_ql, size_p, alpha, t, nbits):
    """orc_keyswitch_* against the statements of KeySwitchHYBRID (keyswitch-hybrid.cpp:330-480) run on the reference's own
    DCRTPoly: digits (incl. a shorter last digit), the inner product with the key, both ApproxModDown (CKKS and BGV form)."""
    ref.ref_keyswitch_hybrid.argtypes = [vp, vp, u32, u32, vp, vp, u64, ctypes.POINTER(vp), vp]
    T, B = size_ql + size_p, 2
    q, psi = oracle.gen_params(nbits, n, T)
    cQl = ref.ref_ctx_create(n, size_ql, p(q[:size_ql].copy()), p(psi[:size_ql].copy()))
    cQlP = ref.ref_ctx_create(n, T, p(q), p(psi))
    qlp = oracle.ctx(n, q, psi)
    np_ = (size_ql + alpha - 1) // alpha
    c = oracle.fill_ring(B, n, q[:size_ql].copy(), 91)
    c[0, :, 0] = q[:size_ql] - 1
    evk_b = oracle.fill_ring(np_, n, q, 92)
    evk_a = oracle.fill_ring(np_, n, q, 93)
    hc = ref.ref_polys_load(cQl, p(c), B, 0)
    outs = (vp * 2)()
    digits = np.empty((np_, B, T, n), np.uint64)
    assert ref.ref_keyswitch_hybrid(hc, cQlP, size_ql, alpha, p(evk_b), p(evk_a), t, outs, p(digits)) == 0, ref.ref_last_error()
    assert np.array_equal(oracle.keyswitch_precompute(qlp, size_ql, alpha, c), digits), "digits"
    want = oracle.keyswitch_hybrid(qlp, size_ql, alpha, c, evk_b, evk_a, t)
    for k in range(2):
        res = np.empty((B, size_ql, n), np.uint64)
        ref.ref_polys_store(outs[k], p(res))
        assert np.array_equal(res, want[k]), f"ct{k}"
        ref.ref_polys_free(outs[k])
    ref.ref_polys_free(hc)
    ref.ref_ctx_destroy(cQl); ref.ref_ctx_destroy(cQlP)


# (bits of q, towers, t): one case per branch of dcrtpoly-impl.h:1537-1816, branch = 4*pow2 + 2*split + reduce
DECRYPT_CASES = [(40, 4, 1 << 10, 4), (40, 4, 1 << 30, 5), (60, 4, 1 << 16, 6), (60, 4, 1 << 40, 7),
                 (30, 4, 65537, 0), (40, 4, 65537, 1), (60, 4, 65537, 2), (60, 4, 33554467, 3), (60, 9, 65537, 2), (60, 40, 65537, 3), (55, 2, 2, 6)]


@pytest.mark.parametrize("nbits,towers,t,branch", DECRYPT_CASES)
def test_decrypt_scale_and_round_matches_reference(oracle, ref, nbits, towers, t, branch):
    """orc_scale_and_round_native and orc_decrypt_tables against DCRTPoly::ScaleAndRound(t, ...) -> NativePoly with the
    tables built by the reference's BigInteger as bfvrns-cryptoparameters.cpp:447-494 builds them; every branch."""
    ref.ref_polys_scale_and_round_native.argtypes = [vp, u64, vp, vp, vp, vp, vp]
    n, B = 256, 3
    q, psi = oracle.gen_params(nbits, n, towers)
    modt, bmodt, frac, bfrac, br = oracle.decrypt_tables(q, t)
    assert br == branch
    c = ref.ref_ctx_create(n, towers, p(q), p(psi))
    x = oracle.fill_ring(B, n, q, 211)
    x[0, :, 0] = q - 1
    x[0, :, 1] = 0
    x[1, :, 2] = q >> np.uint64(1)
    hx = ref.ref_polys_load(c, p(x), B, 1)
    out = np.empty((B, n), np.uint64)
    rm, rb = np.zeros(towers, np.uint64), np.zeros(towers, np.uint64)
    rf, rbf = np.zeros(towers, np.float64), np.zeros(towers, np.float64)
    assert ref.ref_polys_scale_and_round_native(hx, t, p(out), p(rm), p(rb), p(rf), p(rbf)) == 0, ref.ref_last_error()
    assert np.array_equal(rm, modt) and np.array_equal(rb, bmodt), "integer tables"
    assert np.array_equal(rf.view(np.uint64), frac.view(np.uint64)) and np.array_equal(rbf.view(np.uint64), bfrac.view(np.uint64)), "fractions"
    assert np.array_equal(oracle.scale_and_round_native(x, n, q, t), out)
    assert (out < t).all()
    ref.ref_polys_free(hx); ref.ref_ctx_destroy(c)


def behz_ring(oracle, nbits, n, nq, nB, msk_bits=0):
    """Q, B and msk the way bfvrns-cryptoparameters.cpp:665-711 chains them (PreviousPrime below the last modulus of
    Q); msk_bits > 0 replaces msk by a prime of that size (the scheme's `msk is large enough` loop, :700-706)."""
    q, psi = oracle.gen_params(nbits, n, nq + nB + 1)
    if msk_bits:
        q[-1] = oracle.L.orc_next_prime(oracle.L.orc_first_prime(msk_bits, 2 * n), 2 * n)
        psi[-1] = oracle.L.orc_root_of_unity(2 * n, int(q[-1]))
    return q, psi


# (bits, nq, |B|, t, msk bits): the scheme's shape |B| = nq, small and large t, msk above every q_i, a single tower
BEHZ_CASES = [(60, 3, 3, 65537, 0), (45, 2, 2, 2, 0), (30, 4, 4, 1 << 20, 0), (40, 2, 2, 65537, 50), (55, 1, 1, 786433, 0),
              (60, 5, 2, 65537, 0)]


@pytest.mark.parametrize("nbits,nq,nB,t,msk_bits", BEHZ_CASES)
def test_behz_matches_reference(oracle, ref, nbits, nq, nB, t, msk_bits):
    """BEHZ restatement (tables and the four methods) against the reference's own DCRTPoly: FastBaseConvqToBskMontgomery
    from either format, FastRNSFloorq (whole poly, incl. the twisted towers of Q it leaves behind), FastBaseConvSK and the
    decryption ScaleAndRound(moduliQ, t, tgamma, ...), dcrtpoly-impl.h:2005-2050, 2069-2420."""
    ref.ref_behz_create.restype = vp
    ref.ref_behz_create.argtypes = [vp, u32, u64]
    ref.ref_behz_destroy.argtypes = [vp]
    ref.ref_behz_table.restype = sz
    ref.ref_behz_table.argtypes = [vp, ci, vp]
    ref.ref_polys_behz_q_to_bsk.argtypes = [vp, vp]
    ref.ref_polys_behz_floor.argtypes = [vp, vp]
    ref.ref_polys_behz_conv_sk.argtypes = [vp, vp, vp]
    ref.ref_polys_behz_scale_and_round.argtypes = [vp, vp, vp]
    n, B, nb = 128, 3, nB + 1
    q, psi = behz_ring(oracle, nbits, n, nq, nB, msk_bits)
    ob = oracle.behz(n, q, psi, nq, t)
    cQBsk = ref.ref_ctx_create(n, nq + nb, p(q), p(psi))
    cQ = ref.ref_ctx_create(n, nq, p(q[:nq].copy()), p(psi[:nq].copy()))
    z = ref.ref_behz_create(cQBsk, nq, t)
    for k in range(ob.NTABLES):
        mine = ob.table(k)
        theirs = np.empty(ref.ref_behz_table(z, k, None), np.uint64)
        ref.ref_behz_table(z, k, p(theirs))
        assert np.array_equal(mine, theirs), f"table {k}"
    x = oracle.fill_ring(B, n, q[:nq].copy(), 311)
    x[0, :, 0] = q[:nq] - np.uint64(1)
    x[0, :, 1] = 0
    for fmt in (0, 1):
        h = ref.ref_polys_load(cQ, p(x), B, fmt)
        assert ref.ref_polys_behz_q_to_bsk(h, z) == 0, ref.ref_last_error()
        got = np.empty((B, nq + nb, n), np.uint64)
        ref.ref_polys_store(h, p(got))
        assert np.array_equal(ob.q_to_bsk_montgomery(x, fmt), got), f"FastBaseConvqToBskMontgomery format {fmt}"
        ref.ref_polys_free(h)
    y = oracle.fill_ring(B, n, q, 312)
    y[0, :, 0] = q - np.uint64(1)
    y[1, :, 3] = 0
    h = ref.ref_polys_load(cQBsk, p(y), B, 1)
    assert ref.ref_polys_behz_floor(h, z) == 0, ref.ref_last_error()
    got = np.empty((B, nq + nb, n), np.uint64)
    ref.ref_polys_store(h, p(got))
    mine = ob.fast_rns_floor_q(y)
    assert np.array_equal(mine, got), "FastRNSFloorq"
    assert ref.ref_polys_behz_conv_sk(h, z, cQ) == 0, ref.ref_last_error()
    got2 = np.empty((B, nq, n), np.uint64)
    ref.ref_polys_store(h, p(got2))
    assert np.array_equal(ob.fast_base_conv_sk(mine), got2), "FastBaseConvSK"
    ref.ref_polys_free(h)
    h = ref.ref_polys_load(cQ, p(x), B, 1)
    out = np.empty((B, n), np.uint64)
    assert ref.ref_polys_behz_scale_and_round(h, z, p(out)) == 0, ref.ref_last_error()
    assert np.array_equal(ob.scale_and_round(x), out), "ScaleAndRound(moduliQ, t, tgamma, ...)"
    ref.ref_polys_free(h)
    ref.ref_behz_destroy(z); ref.ref_ctx_destroy(cQBsk); ref.ref_ctx_destroy(cQ)


@pytest.mark.parametrize("nbits,n,towers,fmt", [(60, 256, 3, 0), (60, 256, 3, 1), (30, 128, 4, 0), (45, 64, 1, 0), (58, 512, 5, 1)])
def test_crt_decompose_matches_reference(oracle, ref, nbits, n, towers, fmt):
    """orc_crt_decompose against DCRTPoly::CRTDecompose(0) (dcrtpoly-impl.h:266-288), every digit, both input formats;
    moduli of mixed sizes exercise both branches of NativeVector::SwitchModulus."""
    ref.ref_polys_crt_decompose.restype = vp
    ref.ref_polys_crt_decompose.argtypes = [vp, u32]
    q, psi = oracle.gen_params(nbits, n, towers)
    if towers >= 4:                                   # a much smaller modulus among large ones: the `v %= nm` path
        q[1] = oracle.L.orc_next_prime(oracle.L.orc_first_prime(20, 2 * n), 2 * n)
        psi[1] = oracle.L.orc_root_of_unity(2 * n, int(q[1]))
    B = 2
    c = ref.ref_ctx_create(n, towers, p(q), p(psi))
    x = oracle.fill_ring(B, n, q, 511)
    x[0, :, 0] = q - np.uint64(1)
    x[0, :, 1] = q >> np.uint64(1)
    x[0, :, 2] = (q >> np.uint64(1)) + np.uint64(1)
    hx = ref.ref_polys_load(c, p(x), B, fmt)
    mine = oracle.ctx(n, q, psi).crt_decompose(x, fmt)
    for i in range(towers):
        hd = ref.ref_polys_crt_decompose(hx, i)
        got = np.empty((B, towers, n), np.uint64)
        ref.ref_polys_store(hd, p(got))
        assert np.array_equal(mine[i], got), f"digit {i}"
        ref.ref_polys_free(hd)
    ref.ref_polys_free(hx); ref.ref_ctx_destroy(c)
