"""SURVEY §8(f) rank 3 on the B200 (csrc/basis.cu): ApproxSwitchCRTBasis, ApproxModUp, ApproxModDown (CKKS and BGV
forms), DropLastElementAndScale, ModReduce — vs the oracle restatement of dcrtpoly-impl.h:746-812, 1034-1175 (itself
pinned against the compiled reference) and the reference's golden digests.  Bit-exact."""
import hashlib
import json
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))
with open(os.path.join(HERE, "golden", "ext_vectors.json")) as f:
    EXT = json.load(f)


def digest(a):
    return hashlib.sha256(np.ascontiguousarray(a, np.uint64).tobytes()).hexdigest()


class Dev:
    """a few device buffers that free themselves"""

    def __init__(self, be):
        self.be, self.ptrs = be, []

    def put(self, arr):
        p = self.be.alloc(max(arr.nbytes, 8))
        self.be.h2d(p, arr)
        self.ptrs.append(p)
        return p

    def new(self, nwords):
        p = self.be.alloc(max(nwords * 8, 8))
        self.ptrs.append(p)
        return p

    def close(self):
        for p in self.ptrs:
            self.be.free(p)


def setup(pkg, backend, oracle, nbits, n, sq, sp):
    q, psi = oracle.gen_params(nbits, n, sq + sp)
    ring = pkg.Ring(backend, n, q, psi)
    rq, rp = pkg.RingView(ring, 0, sq), pkg.RingView(ring, sq, sp)
    return q, psi, ring, rq, rp, oracle.basis(n, q, psi, sq)


@pytest.mark.parametrize("nbits,n,sq,sp,batch", [(30, 16, 2, 1, 2), (60, 64, 3, 2, 3), (60, 1 << 10, 5, 3, 2),
                                                 (45, 1 << 12, 2, 4, 1), (60, 1 << 13, 12, 9, 1), (60, 1 << 14, 4, 2, 2)])
def test_basis_conversion_vs_oracle(pkg, backend, oracle, nbits, n, sq, sp, batch):
    q, psi, ring, rq, rp, ob = setup(pkg, backend, oracle, nbits, n, sq, sp)
    d = Dev(backend)
    q2p, p2q = pkg.Basis(rq, rp), pkg.Basis(rp, rq)
    for bs, rev in ((q2p, False), (p2q, True)):
        hi, hm = bs.tables()
        ohi, ohm = ob.tables(rev)
        assert np.array_equal(hi, ohi) and np.array_equal(hm, ohm)
    x = oracle.fill_ring(batch, n, q[:sq].copy(), 11)
    x[0, :, 0] = q[:sq] - 1
    x[0, :, 1] = 0
    dx, dp = d.put(x), d.new(batch * sp * n)
    q2p.approx_switch(dp, dx, batch)
    assert np.array_equal(backend.d2h(dp, batch * sp * n).reshape(batch, sp, n), ob.approx_switch(x))
    dqp = d.new(batch * (sq + sp) * n)
    for fmt in (0, 1):
        q2p.approx_mod_up(dqp, dx, batch, fmt)
        got = backend.d2h(dqp, batch * (sq + sp) * n).reshape(batch, sq + sp, n)
        assert np.array_equal(got, ob.mod_up(x, fmt)), f"ApproxModUp format {fmt}"
        assert np.array_equal(backend.d2h(dx, x.size).reshape(x.shape), x), "source modified"
    y = oracle.fill_ring(batch, n, q, 12)
    y[0, :, 0] = q - 1
    dy, dq = d.put(y), d.new(batch * sq * n)
    for t in (0, 65537, 2, 0):
        p2q.approx_mod_down(dq, dy, batch, t)
        got = backend.d2h(dq, batch * sq * n).reshape(batch, sq, n)
        assert np.array_equal(got, ob.mod_down(y, t)), f"ApproxModDown t={t}"
    # explicit tables, as a scheme would hand them over
    hi, hm = ob.tables(False)
    b2 = pkg.Basis(rq, rp, hi, hm)
    b2.approx_switch(dp, dx, batch)
    assert np.array_equal(backend.d2h(dp, batch * sp * n).reshape(batch, sp, n), ob.approx_switch(x))
    for o in (b2, q2p, p2q, rq, rp):
        o.close()
    d.close()
    ring.close()


@pytest.mark.parametrize("nbits,n,towers,batch", [(30, 16, 2, 2), (60, 64, 4, 3), (60, 1 << 11, 3, 2), (50, 1 << 14, 6, 1)])
def test_rescale_vs_oracle(pkg, backend, oracle, nbits, n, towers, batch):
    q, psi = oracle.gen_params(nbits, n, towers)
    ring = pkg.Ring(backend, n, q, psi)
    plan = pkg.Rescale(ring)
    a, b = plan.tables()
    oa, obb = oracle.rescale_tables(q)
    assert np.array_equal(a, oa) and np.array_equal(b, obb)
    d = Dev(backend)
    y = oracle.fill_ring(batch, n, q, 21)
    y[0, :, 0] = q - 1
    y[0, :, 1] = 0
    dy, do = d.put(y), d.new(batch * (towers - 1) * n)
    for t in (0, 65537, 0, 2):
        for fmt in (0, 1):
            if t:
                plan.mod_reduce(do, dy, batch, t, fmt)
            else:
                plan.drop_last_element_and_scale(do, dy, batch, fmt)
            got = backend.d2h(do, batch * (towers - 1) * n).reshape(batch, towers - 1, n)
            assert np.array_equal(got, oracle.rescale(n, q, psi, y, fmt, t)), f"rescale t={t} format {fmt}"
    assert np.array_equal(backend.d2h(dy, y.size).reshape(y.shape), y), "source modified"
    plan.close()
    d.close()
    ring.close()


def test_basis_errors(pkg, backend, oracle):
    q, psi = oracle.gen_params(60, 16, 3)
    ring = pkg.Ring(backend, 16, q, psi)
    N = pkg._native
    with pytest.raises(pkg.FheError):
        pkg.RingView(ring, 2, 2)
    one = pkg.RingView(ring, 0, 1)
    with pytest.raises(pkg.FheError, match="renders it invalid"):
        pkg.Rescale(one)                                       # the reference's config_error (dcrtpoly-impl.h:722-723)
    plan = pkg.Rescale(ring)
    assert N.lib.fhe_mod_reduce(plan.h, 8, 8, 1, 0, 0, 0) == N.FHE_EINVAL
    dup = pkg.Ring(backend, 16, np.array([q[0], q[0]], np.uint64), np.array([psi[0], psi[0]], np.uint64))
    with pytest.raises(pkg.FheError, match="coprime"):
        pkg.Basis(dup, one)
    for o in (plan, one, dup):
        o.close()
    ring.close()


def test_mod_up_then_mod_down_is_identity_on_q(pkg, backend, oracle):
    """Full-size property (cfg-4 ring size): ModDown(P * x in QP) = x, and ModDown(ModUp(x)) is within the approximate
    conversion's error of x / P -> checked through linearity instead: ModDown(a + b) - ModDown(a) - ModDown(b) is in
    {-1, 0, 1} * small per coefficient is not exact, so only the exact identity is asserted."""
    n, sq, sp, batch = 1 << 16, 6, 2, 2
    q, psi, ring, rq, rp, ob = setup(pkg, backend, oracle, 60, n, sq, sp)
    p2q = pkg.Basis(rp, rq)
    d = Dev(backend)
    x = oracle.fill_ring(batch, n, q[:sq].copy(), 31)
    # y = P * x over QP: Q towers hold x * [P]_{q_i}, P towers hold 0
    y = np.zeros((batch, sq + sp, n), np.uint64)
    for i in range(sq):
        pm = 1
        for pj in q[sq:]:
            pm = pm * int(pj) % int(q[i])
        y[:, i, :] = oracle.vec_scalar(2, x[:, i, :].reshape(-1), pm, int(q[i])).reshape(batch, n)
    dy, dq = d.put(y), d.new(batch * sq * n)
    p2q.approx_mod_down(dq, dy, batch, 0)
    assert np.array_equal(backend.d2h(dq, batch * sq * n).reshape(batch, sq, n), x)
    for o in (p2q, rq, rp):
        o.close()
    d.close()
    ring.close()


@pytest.mark.parametrize("case", EXT["basis"], ids=lambda c: f"N{c['N']}_Q{c['sizeQ']}_P{c['sizeP']}")
def test_basis_conversion_reference_golden(pkg, backend, oracle, case):
    n, sq, sp, batch = case["N"], case["sizeQ"], case["sizeP"], case["batch"]
    q, psi = np.array(case["q"], np.uint64), np.array(case["psi"], np.uint64)
    ring = pkg.Ring(backend, n, q, psi)
    rq, rp = pkg.RingView(ring, 0, sq), pkg.RingView(ring, sq, sp)
    q2p, p2q, plan = pkg.Basis(rq, rp), pkg.Basis(rp, rq), pkg.Rescale(ring)
    x = oracle.fill_ring(batch, n, q[:sq].copy(), case["seed"])
    y = oracle.fill_ring(batch, n, q, case["seed"] + 1)
    assert digest(x) == case["sha_x"] and digest(y) == case["sha_y"]
    d = Dev(backend)
    dx, dy = d.put(x), d.put(y)
    out = d.new(batch * (sq + sp) * n)
    q2p.approx_switch(out, dx, batch)
    assert digest(backend.d2h(out, batch * sp * n)) == case["sha_switch"]
    for fmt in (0, 1):
        q2p.approx_mod_up(out, dx, batch, fmt)
        assert digest(backend.d2h(out, batch * (sq + sp) * n)) == case[f"sha_modup_f{fmt}"]
    for t in (0, 65537):
        p2q.approx_mod_down(out, dy, batch, t)
        assert digest(backend.d2h(out, batch * sq * n)) == case[f"sha_moddown_t{t}"]
        for fmt in (0, 1):
            if t:
                plan.mod_reduce(out, dy, batch, t, fmt)
            else:
                plan.drop_last_element_and_scale(out, dy, batch, fmt)
            assert digest(backend.d2h(out, batch * (sq + sp - 1) * n)) == case[f"sha_rescale_t{t}_f{fmt}"]
    for o in (q2p, p2q, plan, rq, rp):
        o.close()
    d.close()
    ring.close()


# ---- BFV (HPS) half of rank 3: SwitchCRTBasis, ExpandCRTBasis, ScaleAndRound ---------------------------------
@pytest.mark.parametrize("nbits,n,sq,sp,batch", [(30, 16, 2, 3, 2), (60, 64, 3, 4, 3), (60, 1 << 10, 5, 6, 2),
                                                 (55, 1 << 12, 9, 12, 1), (60, 1 << 14, 4, 5, 2)])
def test_exact_switch_and_expand_vs_oracle(pkg, backend, oracle, nbits, n, sq, sp, batch):
    q, psi, ring, rq, rp, ob = setup(pkg, backend, oracle, nbits, n, sq, sp)
    d = Dev(backend)
    q2p = pkg.Basis(rq, rp)
    x = oracle.fill_ring(batch, n, q[:sq].copy(), 71)
    x[0, :, 0] = q[:sq] - 1
    x[0, :, 1] = 0
    dx, dp, dqp = d.put(x), d.new(batch * sp * n), d.new(batch * (sq + sp) * n)
    q2p.switch_crt_basis(dp, dx, batch)
    assert np.array_equal(backend.d2h(dp, batch * sp * n).reshape(batch, sp, n), ob.switch_exact(x))
    for fmt in (0, 1):
        for rf in (0, 1):
            q2p.expand_crt_basis(dqp, dx, batch, fmt, rf)
            got = backend.d2h(dqp, batch * (sq + sp) * n).reshape(batch, sq + sp, n)
            assert np.array_equal(got, ob.expand(x, fmt, rf)), f"ExpandCRTBasis {fmt}->{rf}"
    assert np.array_equal(backend.d2h(dx, x.size).reshape(x.shape), x), "source modified"
    q2p.approx_switch(dp, dx, batch)                       # the approximate form still works on the same basis object
    assert np.array_equal(backend.d2h(dp, batch * sp * n).reshape(batch, sp, n), ob.approx_switch(x))
    for o in (q2p, rq, rp):
        o.close()
    d.close()
    ring.close()


@pytest.mark.parametrize("case", EXT["bfv"], ids=lambda c: f"N{c['N']}_Q{c['sizeQ']}_P{c['sizeP']}")
def test_bfv_set_reference_golden(pkg, backend, oracle, case):
    """Against the reference's digests, with the reference's own ScaleAndRound tables (stored in the fixture)."""
    n, sq, sp, batch = case["N"], case["sizeQ"], case["sizeP"], case["batch"]
    q, psi = np.array(case["q"], np.uint64), np.array(case["psi"], np.uint64)
    ring = pkg.Ring(backend, n, q, psi)
    rq, rp = pkg.RingView(ring, 0, sq), pkg.RingView(ring, sq, sp)
    q2p = pkg.Basis(rq, rp)
    frac = np.array([float.fromhex(v) for v in case["frac_hex"]], np.float64)
    modr = np.array(case["modr"], np.uint64).reshape(sp, sq + 1)
    plan = pkg.ScaleRound(ring, sp, frac, modr, out_last=True)
    x = oracle.fill_ring(batch, n, q[:sq].copy(), case["seed"])
    y = oracle.fill_ring(batch, n, q, case["seed"] + 1)
    assert digest(x) == case["sha_x"] and digest(y) == case["sha_y"]
    d = Dev(backend)
    dx, dy, out = d.put(x), d.put(y), d.new(batch * (sq + sp) * n)
    q2p.switch_crt_basis(out, dx, batch)
    assert digest(backend.d2h(out, batch * sp * n)) == case["sha_switch_exact"]
    for fmt in (0, 1):
        for rf in (0, 1):
            q2p.expand_crt_basis(out, dx, batch, fmt, rf)
            assert digest(backend.d2h(out, batch * (sq + sp) * n)) == case[f"sha_expand_f{fmt}_r{rf}"], (fmt, rf)
    plan.scale_and_round(out, dy, batch)
    got = backend.d2h(out, batch * sp * n)
    assert digest(got) == case["sha_scale_round"]
    assert np.array_equal(got.reshape(batch, sp, n), oracle.basis(n, q, psi, sq).scale_and_round(y, frac, modr, 1))
    for o in (plan, q2p, rq, rp):
        o.close()
    d.close()
    ring.close()


def test_scale_and_round_to_first_towers(pkg, backend, oracle):
    """The other orientation (output = the FIRST towers, dcrtpoly-impl.h:1892-1895) with arbitrary tables."""
    n, sq, sp, batch = 1 << 11, 3, 4, 2
    q, psi, ring, rq, rp, ob = setup(pkg, backend, oracle, 60, n, sq, sp)
    rng = np.random.default_rng(9)
    frac = rng.random(sp)
    modo = np.stack([rng.integers(0, int(m), size=sp + 1, dtype=np.uint64) for m in q[:sq]])
    plan = pkg.ScaleRound(ring, sq, frac, modo, out_last=False)
    y = oracle.fill_ring(batch, n, q, 81)
    d = Dev(backend)
    dy, out = d.put(y), d.new(batch * sq * n)
    plan.scale_and_round(out, dy, batch)
    assert np.array_equal(backend.d2h(out, batch * sq * n).reshape(batch, sq, n), ob.scale_and_round(y, frac, modo, 0))
    for o in (plan, rq, rp):
        o.close()
    d.close()
    ring.close()


@pytest.mark.parametrize("case", [c for c in EXT["bfv"] if "sha_mult_out" in c], ids=lambda c: f"N{c['N']}_Q{c['sizeQ']}")
def test_bfv_hps_mult_core_reference_golden(pkg, backend, oracle, case):
    """BASELINE configs[4] as a real EvalMult core: ExpandCRTBasis x4 -> tensor product over QR -> INTT -> ScaleAndRound ->
    SwitchCRTBasis on the device (pkg.BfvHpsMult) against the digests of the reference running the same sequence on its
    own DCRTPoly (bfvrns-leveledshe.cpp:193-207, 322-356)."""
    n, sq, sr, batch = case["N"], case["sizeQ"], case["sizeP"], case["batch"]
    q, psi = np.array(case["q"], np.uint64), np.array(case["psi"], np.uint64)
    ring = pkg.Ring(backend, n, q, psi)
    frac = np.array([float.fromhex(v) for v in case["frac_hex"]], np.float64)
    modr = np.array(case["modr"], np.uint64).reshape(sr, sq + 1)
    mult = pkg.BfvHpsMult(ring, sq, frac, modr)
    ins = [oracle.fill_ring(batch, n, q[:sq].copy(), case["seed"] + 10 + k) for k in range(4)]
    d = Dev(backend)
    din = [d.put(v) for v in ins]
    dout = [d.new(batch * sq * n) for _ in range(3)]
    for _ in range(2):                                  # twice: scratch buffers and cached tables are reused
        mult(dout, din[0:2], din[2:4], batch)
        got = [backend.d2h(p_, batch * sq * n) for p_ in dout]
        assert [digest(v) for v in got] == case["sha_mult_out"]
    for v, dv in zip(ins, din):
        assert np.array_equal(backend.d2h(dv, v.size).reshape(v.shape), v), "inputs modified"
    mult.close()
    d.close()
    ring.close()


@pytest.mark.parametrize("wide", [False, True], ids=["split30", "mac128"])
@pytest.mark.parametrize("n,sq,sp,batch", [(256, 33, 11, 2), (512, 17, 20, 1), (128, 40, 7, 1), (64, 8, 10, 2),
                                           (64, 9, 30, 1)])
def test_basis_accumulate_forms_many_towers(pkg, backend, oracle, monkeypatch, wide, n, sq, sp, batch):
    """Both accumulate forms of basis_switch_kernel / scale_round_kernel (carry-free 30-bit split limbs, normalised every
    8 terms; 128-bit multiply-accumulate, FHE_B200_BASIS_WIDE=1) with more source towers than one normalisation period,
    odd and even counts, every column-block width, and the largest operands the moduli allow."""
    if wide:
        monkeypatch.setenv("FHE_B200_BASIS_WIDE", "1")
    else:
        monkeypatch.delenv("FHE_B200_BASIS_WIDE", raising=False)
    q, psi, ring, rq, rp, ob = setup(pkg, backend, oracle, 60, n, sq, sp)
    d = Dev(backend)
    q2p = pkg.Basis(rq, rp)
    x = oracle.fill_ring(batch, n, q[:sq].copy(), 301)
    x[0, :, 0] = q[:sq] - 1
    x[0, :, 1] = 0
    x[0, :, 2] = 1
    dx, dp = d.put(x), d.new(batch * sp * n)
    q2p.approx_switch(dp, dx, batch)
    assert np.array_equal(backend.d2h(dp, batch * sp * n).reshape(batch, sp, n), ob.approx_switch(x))
    q2p.switch_crt_basis(dp, dx, batch)
    assert np.array_equal(backend.d2h(dp, batch * sp * n).reshape(batch, sp, n), ob.switch_exact(x))
    # ScaleAndRound: the inputs themselves are the factors; tables at their largest values
    rng = np.random.default_rng(sq * 100 + sp)
    frac = rng.random(sq)
    modr = np.stack([rng.integers(0, int(m), size=sq + 1, dtype=np.uint64) for m in q[sq:]])
    modr[:, ::2] = (q[sq:] - 1)[:, None]
    plan = pkg.ScaleRound(ring, sp, frac, modr, out_last=True)
    y = oracle.fill_ring(batch, n, q, 302)
    y[0, :, 0] = q - 1
    y[0, :, 1] = q - 1
    y[0, :, 2] = 0
    dy, out = d.put(y), d.new(batch * sp * n)
    plan.scale_and_round(out, dy, batch)
    assert np.array_equal(backend.d2h(out, batch * sp * n).reshape(batch, sp, n), ob.scale_and_round(y, frac, modr, 1))
    for o in (plan, q2p, rq, rp):
        o.close()
    d.close()
    ring.close()
