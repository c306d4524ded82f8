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
    for pbuf in (dc, db, da, dd, d0, d1, o0, o1):
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
    backend.free(x)
    ks.close()
    ring.close()
