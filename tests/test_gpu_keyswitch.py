"""Hybrid key switching (KeySwitchHYBRID, src/pke/lib/keyswitch/keyswitch-hybrid.cpp:324-480) as a device composition,
against the CPU oracle (pinned against the reference's own DCRTPoly in tests/test_oracle_vs_ref.py) — bit exact."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("logn,size_ql,size_p,alpha,t,nbits,batch", [
    (6, 6, 2, 2, 0, 60, 2), (8, 5, 2, 2, 0, 60, 3), (10, 4, 3, 3, 65537, 45, 2), (12, 7, 3, 3, 0, 55, 2),
    (13, 3, 1, 1, 0, 30, 1), (14, 8, 3, 3, 0, 60, 2), (12, 9, 2, 1, 0, 60, 1), (16, 4, 2, 2, 0, 60, 1),
])
def test_keyswitch_hybrid_vs_oracle(pkg, backend, oracle, logn, size_ql, size_p, alpha, t, nbits, batch):
    n, T = 1 << logn, size_ql + size_p
    q, psi = oracle.gen_params(nbits, n, T)
    qlp = oracle.ctx(n, q, psi)
    ring = pkg.Ring(backend, n, q, psi)
    ks = pkg.KeySwitch(ring, size_ql, alpha)
    npart = (size_ql + alpha - 1) // alpha
    assert ks.digits == npart
    c = oracle.fill_ring(batch, n, q[:size_ql].copy(), 191)
    c[0, :, 0] = q[:size_ql] - 1
    evk_b, evk_a = oracle.fill_ring(npart, n, q, 192), oracle.fill_ring(npart, n, q, 193)
    dc, db, da = backend.alloc(c.nbytes), backend.alloc(evk_b.nbytes), backend.alloc(evk_a.nbytes)
    backend.h2d(dc, c); backend.h2d(db, evk_b); backend.h2d(da, evk_a)
    words_d = npart * batch * T * n
    dd = backend.alloc(words_d * 8)
    ks.precompute(dd, dc, batch)
    want_d = oracle.keyswitch_precompute(qlp, size_ql, alpha, c)
    assert np.array_equal(backend.d2h(dd, words_d).reshape(want_d.shape), want_d), "digits"
    d0, d1 = backend.alloc(batch * T * n * 8), backend.alloc(batch * T * n * 8)
    ks.inner(d0, d1, dd, db, da, batch)
    w0, w1 = oracle.keyswitch_inner(qlp, want_d, evk_b, evk_a)
    assert np.array_equal(backend.d2h(d0, w0.size).reshape(w0.shape), w0), "cTilda0"
    assert np.array_equal(backend.d2h(d1, w1.size).reshape(w1.shape), w1), "cTilda1"
    o0, o1 = backend.alloc(c.nbytes), backend.alloc(c.nbytes)
    ks(o0, o1, dc, db, da, batch, t)
    want = oracle.keyswitch_hybrid(qlp, size_ql, alpha, c, evk_b, evk_a, t)
    assert np.array_equal(backend.d2h(o0, c.size).reshape(c.shape), want[0]), "ct0"
    assert np.array_equal(backend.d2h(o1, c.size).reshape(c.shape), want[1]), "ct1"
    f0, f1 = backend.alloc(c.nbytes), backend.alloc(c.nbytes)
    ks.fast(f0, f1, dd, db, da, batch, t)                   # EvalFastKeySwitchCore on the digits computed above
    assert np.array_equal(backend.d2h(f0, c.size).reshape(c.shape), want[0]), "fast ct0"
    assert np.array_equal(backend.d2h(f1, c.size).reshape(c.shape), want[1]), "fast ct1"
    # acc += key switch (what the schemes do with KeySwitchCore's result), accumulators adjacent and apart
    qcol = q[:size_ql][None, :, None]                      # moduli below 2^60: the sums fit a word
    for apart in (False, True):
        r0, r1 = oracle.fill_ring(batch, n, q[:size_ql].copy(), 194), oracle.fill_ring(batch, n, q[:size_ql].copy(), 195)
        blk = backend.alloc(3 * c.nbytes)
        a0, a1 = blk, blk + (2 * c.nbytes if apart else c.nbytes)
        backend.h2d(a0, r0); backend.h2d(a1, r1)
        ks.add_to(a0, a1, dc, db, da, batch, t)
        w0, w1 = (r0 + want[0]) % qcol, (r1 + want[1]) % qcol
        assert np.array_equal(backend.d2h(a0, c.size).reshape(c.shape), w0), "acc0"
        assert np.array_equal(backend.d2h(a1, c.size).reshape(c.shape), w1), "acc1"
        backend.free(blk)
    for pbuf in (dc, db, da, dd, d0, d1, o0, o1, f0, f1):
        backend.free(pbuf)
    ks.close()
    ring.close()


@pytest.mark.parametrize("logn,size_ql,size_p,alpha,t", [(11, 6, 2, 2, 0), (15, 5, 2, 2, 65537)])
def test_keyswitch_hybrid_replayed(pkg, backend, oracle, logn, size_ql, size_p, alpha, t):
    """The same plan called again and again: the first call runs its launches directly, the second records them (a captured
    graph per key, csrc/runtime.cu keyswitch_graphed), later ones replay.  Fresh inputs, fresh outputs, a second key, a
    larger batch (the plan's buffers move) and the first key again: every call bit-equal to the oracle."""
    n, T = 1 << logn, size_ql + size_p
    q, psi = oracle.gen_params(58, n, T)
    qlp = oracle.ctx(n, q, psi)
    ring = pkg.Ring(backend, n, q, psi)
    ks = pkg.KeySwitch(ring, size_ql, alpha)
    npart = ks.digits
    keys = []
    for seed in (401, 402):
        kb, ka = oracle.fill_ring(npart, n, q, seed), oracle.fill_ring(npart, n, q, seed + 50)
        db, da = backend.alloc(kb.nbytes), backend.alloc(ka.nbytes)
        backend.h2d(db, kb); backend.h2d(da, ka)
        keys.append((kb, ka, db, da))
    launches = []
    for call, (which, batch) in enumerate([(0, 1), (0, 1), (0, 1), (1, 1), (0, 1), (0, 2), (0, 2), (1, 2), (0, 1)]):
        kb, ka, db, da = keys[which]
        c = oracle.fill_ring(batch, n, q[:size_ql].copy(), 500 + call)
        # the two results far apart (two final launches) on even calls, adjacent halves of one block (one launch) on odd ones
        dc, blk = backend.alloc(c.nbytes), backend.alloc(3 * c.nbytes)
        o0, o1 = blk, blk + (c.nbytes if call % 2 else 2 * c.nbytes)
        backend.h2d(dc, c)
        before = pkg._native.lib.fhe_launch_count()
        ks(o0, o1, dc, db, da, batch, t)
        launches.append(pkg._native.lib.fhe_launch_count() - before)
        want = oracle.keyswitch_hybrid(qlp, size_ql, alpha, c, kb, ka, t)
        assert np.array_equal(backend.d2h(o0, c.size).reshape(c.shape), want[0]), f"ct0, call {call}"
        assert np.array_equal(backend.d2h(o1, c.size).reshape(c.shape), want[1]), f"ct1, call {call}"
        for pbuf in (dc, blk):
            backend.free(pbuf)
    assert launches[0] == launches[2] == launches[4], launches       # a replay counts its kernels like a direct run
    assert launches[1] == launches[3] == launches[0] - 1, launches   # adjacent results: one final launch instead of two
    for _, _, db, da in keys:
        backend.free(db); backend.free(da)
    ks.close()
    ring.close()


@pytest.mark.parametrize("logn,size_ql,size_p,alpha,t", [(12, 7, 3, 3, 0), (14, 5, 2, 2, 65537)])
def test_keyswitch_hybrid_on_the_single_pass_ntt(pkg, backend, oracle, monkeypatch, logn, size_ql, size_p, alpha, t):
    """The composition's transforms read and write towers of wider polys (source and destination strides differ): the same
    through the single-pass NTT kernel (FHE_B200_NTT_FUSED=2)."""
    monkeypatch.setenv("FHE_B200_NTT_FUSED", "2")
    n, T, batch = 1 << logn, size_ql + size_p, 2
    q, psi = oracle.gen_params(57, n, T)
    qlp = oracle.ctx(n, q, psi)
    ring = pkg.Ring(backend, n, q, psi)
    ks = pkg.KeySwitch(ring, size_ql, alpha)
    npart = ks.digits
    c = oracle.fill_ring(batch, n, q[:size_ql].copy(), 611)
    evk_b, evk_a = oracle.fill_ring(npart, n, q, 612), oracle.fill_ring(npart, n, q, 613)
    dc, db, da = backend.alloc(c.nbytes), backend.alloc(evk_b.nbytes), backend.alloc(evk_a.nbytes)
    backend.h2d(dc, c); backend.h2d(db, evk_b); backend.h2d(da, evk_a)
    o0, o1 = backend.alloc(c.nbytes), backend.alloc(c.nbytes)
    want = oracle.keyswitch_hybrid(qlp, size_ql, alpha, c, evk_b, evk_a, t)
    for call in range(3):                                     # direct, captured, replayed
        ks(o0, o1, dc, db, da, batch, t)
        assert np.array_equal(backend.d2h(o0, c.size).reshape(c.shape), want[0]), f"ct0, call {call}"
        assert np.array_equal(backend.d2h(o1, c.size).reshape(c.shape), want[1]), f"ct1, call {call}"
    for pbuf in (dc, db, da, o0, o1):
        backend.free(pbuf)
    ks.close()
    ring.close()


def test_keyswitch_argument_checks(pkg, backend, oracle):
    n = 64
    q, psi = oracle.gen_params(60, n, 4)
    ring = pkg.Ring(backend, n, q, psi)
    for size_ql, alpha in ((0, 1), (4, 1), (2, 0)):
        with pytest.raises(pkg.FheError):
            pkg.KeySwitch(ring, size_ql, alpha)
    ks = pkg.KeySwitch(ring, 3, 2)
    x = backend.alloc(4 * 4 * n * 8)
    with pytest.raises(pkg.FheError):
        ks(x, x, x, x, x, 1)                        # the two outputs must be distinct
    ks(x, x, x, x, x, 0)                            # an empty batch is a no-op
    with pytest.raises(pkg.FheError):
        ks.add_to(x, x + 4 * n * 8, x, x, x, 1)     # an accumulator must not be the input
    with pytest.raises(pkg.FheError):
        ks.fast(x, x, x, x, x, 1)
    with pytest.raises(pkg.FheError):
        ks.fast(x, x + 8, x, x, x, 1)               # 16-byte alignment
    with pytest.raises(pkg.FheError):
        ks.fast(x, 0, x, x, x, 1)
    ks.fast(x, x, x, x, x, 0)
    backend.free(x)
    ks.close()
    ring.close()


# ---- BFV decryption: ScaleAndRound(t, ...) -> NativePoly (dcrtpoly-impl.h:1537-1816), every branch ---------------
@pytest.mark.parametrize("nbits,towers,t,branch", [
    (40, 4, 1 << 10, 4), (40, 4, 1 << 30, 5), (60, 4, 1 << 16, 6), (60, 4, 1 << 40, 7),
    (30, 4, 65537, 0), (40, 4, 65537, 1), (60, 4, 65537, 2), (60, 4, 33554467, 3), (60, 40, 65537, 3), (55, 2, 2, 6),
    (60, 16, 65537, 3), (60, 30, 1 << 20, 6),
])
def test_decrypt_scale_and_round_vs_oracle(pkg, backend, oracle, nbits, towers, t, branch):
    n, batch = 1 << 11, 3
    q, psi = oracle.gen_params(nbits, n, towers)
    ring = pkg.Ring(backend, n, q, psi)
    want_tables = oracle.decrypt_tables(q, t)
    assert want_tables[4] == branch
    plan = pkg.DecryptScale(ring, t)                         # tables derived from the moduli
    got = plan.tables()
    assert got[4] == branch
    for a, b in zip(got[:2], want_tables[:2]):
        assert np.array_equal(a, b)
    for a, b in zip(got[2:4], want_tables[2:4]):
        assert np.array_equal(a.view(np.uint64), b.view(np.uint64))
    x = oracle.fill_ring(batch, n, q, 311)
    x[0, :, 0] = q - 1
    x[0, :, 1] = 0
    x[1, :, 2] = q >> np.uint64(1)
    dx, dy = backend.alloc(x.nbytes), backend.alloc(batch * n * 8)
    backend.h2d(dx, x)
    plan(dy, dx, batch)
    want = oracle.scale_and_round_native(x, n, q, t)
    assert np.array_equal(backend.d2h(dy, batch * n).reshape(batch, n), want)
    plan2 = pkg.DecryptScale(ring, t, want_tables[:4])       # the scheme's own tables passed in
    plan2(dy, dx, batch)
    assert np.array_equal(backend.d2h(dy, batch * n).reshape(batch, n), want)
    plan.close(); plan2.close()
    backend.free(dx); backend.free(dy)
    ring.close()


def test_decrypt_scale_recovers_a_scaled_message(pkg, backend, oracle):
    """Property at full size (N = 2^15, 16 towers): x = round(Q/t * m) + small noise decrypts to m."""
    n, towers, t = 1 << 15, 16, 65537
    q, psi = oracle.gen_params(60, n, towers)
    Q = 1
    for qi in q:
        Q *= int(qi)
    rng = np.random.default_rng(5)
    m = rng.integers(0, t, size=n, dtype=np.uint64)
    noise = rng.integers(-1000, 1000, size=n)
    x = np.empty((1, towers, n), np.uint64)
    big = [((Q * int(mi) + t // 2) // t + int(ei)) % Q for mi, ei in zip(m[:64], noise[:64])]
    for i, qi in enumerate(q):
        x[0, i, :64] = [b % int(qi) for b in big]
    m[64:] = 0
    x[0, :, 64:] = 0
    ring = pkg.Ring(backend, n, q, psi)
    plan = pkg.DecryptScale(ring, t)
    dx, dy = backend.alloc(x.nbytes), backend.alloc(n * 8)
    backend.h2d(dx, x)
    plan(dy, dx, 1)
    assert np.array_equal(backend.d2h(dy, n), m)
    plan.close()
    backend.free(dx); backend.free(dy)
    ring.close()


# ---- BV key switching's digits: DCRTPolyImpl::CRTDecompose(0) (dcrtpoly-impl.h:266-288) ---------------------------
@pytest.mark.parametrize("logn,nbits,towers,batch,fmt", [
    (6, 60, 3, 2, 0), (6, 60, 3, 2, 1), (10, 30, 4, 3, 0), (11, 45, 1, 1, 0), (12, 58, 5, 2, 1), (14, 60, 8, 1, 0), (16, 60, 4, 1, 0),
])
def test_crt_decompose_vs_oracle(pkg, backend, oracle, logn, nbits, towers, batch, fmt):
    n = 1 << logn
    q, psi = oracle.gen_params(nbits, n, towers)
    if towers >= 4:                                   # a much smaller modulus among large ones: the `v %= nm` path
        q[1] = oracle.L.orc_next_prime(oracle.L.orc_first_prime(20, 2 * n), 2 * n)
        psi[1] = oracle.L.orc_root_of_unity(2 * n, int(q[1]))
    ring = pkg.Ring(backend, n, q, psi)
    plan = pkg.CrtDecompose(ring)
    x = oracle.fill_ring(batch, n, q, 611)
    x[0, :, 0] = q - np.uint64(1)
    x[0, :, 1] = q >> np.uint64(1)
    x[0, :, 2] = (q >> np.uint64(1)) + np.uint64(1)
    dx, dd = backend.alloc(x.nbytes), backend.alloc(towers * x.nbytes)
    backend.h2d(dx, x)
    for _ in range(2):
        plan(dd, dx, batch, fmt)
        got = backend.d2h(dd, towers * x.size).reshape(towers, batch, towers, n)
        assert np.array_equal(got, oracle.ctx(n, q, psi).crt_decompose(x, fmt))
    assert np.array_equal(backend.d2h(dx, x.size).reshape(x.shape), x), "input modified"
    with pytest.raises(pkg.FheError):
        plan(dx, dx, batch, fmt)
    backend.free(dx); backend.free(dd)
    plan.close()
    ring.close()
