"""BFV HPSPOVERQ / HPSPOVERQLEVELED set and the reverse-order expansion on the B200: ExpandCRTBasisReverseOrder,
FastExpandCRTBasisPloverQ, ExpandCRTBasisQlHat, ScaleAndRoundPOverQ (dcrtpoly-impl.h:1361-1468, 1514-1534, 2052-2066)
vs the oracle restatement (pinned against the compiled reference, tests/test_oracle_vs_ref.py) and the reference's
golden digests.  Bit-exact."""
import json
import os

import numpy as np
import pytest

from tests.test_gpu_basis import Dev, digest

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))
with open(os.path.join(HERE, "golden", "ext_vectors.json")) as f:
    EXT = json.load(f)


def constants(oracle, q, seed):
    return oracle.fill_ring(1, 1, np.ascontiguousarray(q, np.uint64), seed).reshape(-1).copy()


def run_set(pkg, backend, oracle, n, aq, apsi, sq, sl, batch, seed, check):
    """Runs the whole set on the device; check(name, device result as [B][T][N], inputs...) decides what to compare with."""
    T = sq + sl
    ring = pkg.Ring(backend, n, aq, apsi)
    rq, rql, rpl = pkg.RingView(ring, 0, sq), pkg.RingView(ring, 0, sl), pkg.RingView(ring, sq, sl)
    q, pl = aq[:sq].copy(), aq[sq:].copy()
    d = Dev(backend)
    x = oracle.fill_ring(batch, n, q, seed)
    x[0, :, 0] = q - 1
    dx, out = d.put(x), d.new(batch * T * n)
    # ExpandCRTBasisReverseOrder: Q -> [Pl | Q]
    q2p = pkg.Basis(rq, rpl)
    for fmt in (0, 1):
        for rf in (0, 1):
            q2p.expand_crt_basis_reverse_order(out, dx, batch, fmt, rf)
            check(f"expand_reverse_f{fmt}_r{rf}", backend.d2h(out, batch * T * n).reshape(batch, T, n), x)
    # FastExpandCRTBasisPloverQ: Q -> [Ql | Pl] with the scheme's tables
    neg, qi = oracle.pl_over_q_tables(q, pl)
    q2pl, pl2ql = pkg.Basis(rq, rpl, neg, qi), pkg.Basis(rpl, rql)
    q2pl.fast_expand_crt_basis_pl_over_q(pl2ql, out, dx, batch)
    check("fast_expand_pl_over_q", backend.d2h(out, batch * 2 * sl * n).reshape(batch, 2 * sl, n), x)
    assert np.array_equal(backend.d2h(dx, x.size).reshape(x.shape), x), "source modified"
    with pytest.raises(pkg.FheError):
        q2pl.fast_expand_crt_basis_pl_over_q(pl2ql, dx, dx, batch)       # in place is refused
    if sl != sq:
        with pytest.raises(pkg.FheError):
            q2pl.fast_expand_crt_basis_pl_over_q(q2p, out, dx, batch)    # second basis must start at Pl
    # ExpandCRTBasisQlHat: Ql -> Q
    c = constants(oracle, q[:sl], seed + 1)
    y = oracle.fill_ring(batch, n, q[:sl].copy(), seed + 2)
    full_c = np.zeros(sq, np.uint64)
    full_c[:sl] = c
    plan = pkg.TowerScale(rq, full_c)
    dy = d.put(y)
    plan.expand_crt_basis_ql_hat(out, dy, sl, batch)
    check("expand_ql_hat", backend.d2h(out, batch * sq * n).reshape(batch, sq, n), y, c)
    plan.close()
    for o in (q2p, q2pl, pl2ql, rq, rql, rpl):
        o.close()
    ring.close()
    # ScaleAndRoundPOverQ, extra tower below / above the q_i
    for name, qs, ps, ex, ep in (("below", q, apsi[:sq], pl[0], apsi[sq]), ("above", aq[1:sq + 1], apsi[1:sq + 1], aq[0], apsi[0])):
        full, fpsi = np.append(qs, ex).astype(np.uint64), np.append(ps, ep).astype(np.uint64)
        r2 = pkg.Ring(backend, n, full, fpsi)
        rq2 = pkg.RingView(r2, 0, sq)
        k = constants(oracle, qs, seed + 3)
        z = oracle.fill_ring(batch, n, full, seed + 4)
        z[0, :, 0] = full - 1
        plan = pkg.TowerScale(rq2, k)
        dz = d.put(z)
        plan.scale_and_round_p_over_q(out, dz, batch)
        check("scale_round_p_over_q_" + name, backend.d2h(out, batch * sq * n).reshape(batch, sq, n), z, np.ascontiguousarray(qs), k)
        plan.close(); rq2.close(); r2.close()
    d.close()


@pytest.mark.parametrize("nbits,n,sq,sl,batch", [(30, 16, 2, 2, 2), (60, 64, 3, 1, 3), (60, 1 << 10, 5, 4, 2),
                                                 (45, 1 << 12, 4, 4, 1), (60, 1 << 13, 12, 12, 1), (60, 1 << 14, 6, 6, 2)])
def test_p_over_q_set_vs_oracle(pkg, backend, oracle, nbits, n, sq, sl, batch):
    aq, apsi = oracle.gen_params(nbits, n, sq + sl)
    q, psi, pl = aq[:sq].copy(), apsi[:sq].copy(), aq[sq:].copy()
    ob = oracle.basis(n, aq, apsi, sq)

    def check(name, got, *ins):
        if name.startswith("expand_reverse"):
            fmt, rf = int(name[-4]), int(name[-1])
            want = ob.expand_reverse(ins[0], fmt, rf)
        elif name == "fast_expand_pl_over_q":
            want = oracle.fast_expand_pl_over_q(ins[0], n, q, q[:sl].copy(), pl)
        elif name == "expand_ql_hat":
            want = oracle.expand_ql_hat(ins[0], n, q, sl, ins[1])
        else:
            want = oracle.scale_round_p_over_q(ins[0], n, ins[1], ins[2])
        assert np.array_equal(got, want), name

    run_set(pkg, backend, oracle, n, aq, apsi, sq, sl, batch, 4100 + n, check)


@pytest.mark.parametrize("case", EXT["bfv_p_over_q"], ids=lambda c: f"N{c['N']}_Q{c['sizeQ']}_L{c['sizeQl']}")
def test_p_over_q_set_reference_golden(pkg, backend, oracle, case):
    n, sq, sl, batch = case["N"], case["sizeQ"], case["sizeQl"], case["batch"]
    aq, apsi = np.array(case["q"], np.uint64), np.array(case["psi"], np.uint64)
    neg, qi = oracle.pl_over_q_tables(aq[:sq].copy(), aq[sq:].copy())
    assert neg.tolist() == case["neg_pl_qhat_inv_modq"] and qi.reshape(-1).tolist() == case["qinv_modp"]

    def check(name, got, *ins):
        if name == "expand_reverse_f0_r0":
            assert digest(ins[0]) == case["sha_x"], "input generator drifted"
        assert digest(got) == case["sha_" + name], name

    run_set(pkg, backend, oracle, n, aq, apsi, sq, sl, batch, case["seed"], check)


def test_tower_scale_edge_cases(pkg, backend, oracle):
    """N = 1 (scalar kernels), empty batch, and the argument checks."""
    n, sq = 1, 3
    q = oracle.gen_params(60, 16, sq + 1)[0]
    ring = pkg.Ring(backend, n, q, None)
    rq = pkg.RingView(ring, 0, sq)
    d = Dev(backend)
    k = constants(oracle, q[:sq], 7)
    plan = pkg.TowerScale(rq, k)
    z = oracle.fill_ring(5, n, q, 8)
    dz, out = d.put(z), d.new(5 * sq)
    plan.scale_and_round_p_over_q(out, dz, 5)
    assert np.array_equal(backend.d2h(out, 5 * sq).reshape(5, sq, 1), oracle.scale_round_p_over_q(z, n, q[:sq].copy(), k))
    y = oracle.fill_ring(5, n, q[:2].copy(), 9)
    dy = d.put(y)
    plan.expand_crt_basis_ql_hat(out, dy, 2, 5)
    assert np.array_equal(backend.d2h(out, 5 * sq).reshape(5, sq, 1), oracle.expand_ql_hat(y, n, q[:sq].copy(), 2, k[:2]))
    plan.times(out, dy, 0)
    d3 = d.put(z[:, :sq].copy())
    plan.times(d3, d3, 5)                                                # Times(vector<NativeInteger>), in place
    assert np.array_equal(backend.d2h(d3, 5 * sq).reshape(5, sq, 1), oracle.expand_ql_hat(z[:, :sq].copy(), n, q[:sq].copy(), sq, k))
    plan.scale_and_round_p_over_q(out, dz, 0)                            # empty batch: nothing to do
    with pytest.raises(pkg.FheError):
        plan.expand_crt_basis_ql_hat(out, dy, 4, 5)                      # more source towers than the ring has
    with pytest.raises(pkg.FheError):
        plan.scale_and_round_p_over_q(out, out, 5)                       # in place
    with pytest.raises(pkg.FheError):
        pkg.TowerScale(rq, q[:sq])                                       # constants must be reduced
    plan.close(); rq.close(); d.close(); ring.close()


@pytest.mark.parametrize("case", EXT["bfv_p_over_q_mult"], ids=lambda c: f"N{c['N']}_Q{c['sizeQ']}")
def test_bfv_hps_p_over_q_mult_reference_golden(pkg, backend, oracle, case):
    """The RNS core of BFVrns EvalMult (HPSPOVERQ) composed on the device vs the reference's digests and the oracle; the
    host table recipes (hps_p_over_q_tables, hps_scale_tables with Q and R exchanged) vs the reference's tables."""
    n, sq, batch, t = case["N"], case["sizeQ"], case["batch"], case["t"]
    q, psi = np.array(case["q"], np.uint64), np.array(case["psi"], np.uint64)
    frac = np.array([float.fromhex(v) for v in case["frac_hex"]], np.float64)
    modq = np.array(case["modq"], np.uint64).reshape(sq, sq + 1)
    f2, m2 = pkg.hps_scale_tables(q[sq:], q[:sq], t)
    assert np.array_equal(f2, frac) and np.array_equal(m2, modq)
    neg, qinv = pkg.hps_p_over_q_tables(q[:sq], q[sq:])
    oneg, oqinv = oracle.pl_over_q_tables(q[:sq].copy(), q[sq:].copy())
    assert np.array_equal(neg, oneg) and np.array_equal(qinv, oqinv)
    ring = pkg.Ring(backend, n, q, psi)
    mult = pkg.BfvHpsPOverQMult(ring, sq, neg, qinv, frac, modq)
    ins = [oracle.fill_ring(batch, n, q[:sq].copy(), case["seed"] + k) for k in range(4)]
    assert [digest(v) for v in ins] == case["sha_in"]
    d = Dev(backend)
    din = [d.put(v) for v in ins]
    dout = [d.new(batch * sq * n) for _ in range(3)]
    mult(dout, din[0:2], din[2:4], batch)
    got = [backend.d2h(o, batch * sq * n).reshape(batch, sq, n) for o in dout]
    assert [digest(v) for v in got] == case["sha_out"]
    want = oracle.basis(n, q, psi, sq).bfv_hps_p_over_q_mult(ins[0], ins[1], ins[2], ins[3], frac, modq)
    for k in range(3):
        assert np.array_equal(got[k], want[k]), f"c{k}"
    for k in range(4):
        assert np.array_equal(backend.d2h(din[k], ins[k].size).reshape(ins[k].shape), ins[k]), "input modified"
    mult.close(); d.close(); ring.close()


@pytest.mark.parametrize("nbits,n,towers,batch", [(60, 1 << 12, 5, 3), (45, 1 << 14, 9, 2)])
def test_times_tower_constants_vs_oracle(pkg, backend, oracle, nbits, n, towers, batch):
    """DCRTPolyImpl::Times(const std::vector<NativeInteger>&) (dcrtpoly-impl.h:617-627), out of place and in place."""
    q, psi = oracle.gen_params(nbits, n, towers)
    ring = pkg.Ring(backend, n, q, psi)
    k = constants(oracle, q, 31)
    plan = pkg.TowerScale(ring, k)
    x = oracle.fill_ring(batch, n, q, 32)
    x[0, :, 0] = q - 1
    want = oracle.expand_ql_hat(x, n, q, towers, k)
    for a in range(batch):
        for i in range(0, towers, 2):
            assert int(want[a, i, 0]) == int(x[a, i, 0]) * int(k[i]) % int(q[i])
    d = Dev(backend)
    dx, out = d.put(x), d.new(x.size)
    plan.times(out, dx, batch)
    assert np.array_equal(backend.d2h(out, x.size).reshape(x.shape), want)
    plan.times(dx, dx, batch)
    assert np.array_equal(backend.d2h(dx, x.size).reshape(x.shape), want)
    plan.close(); d.close(); ring.close()
