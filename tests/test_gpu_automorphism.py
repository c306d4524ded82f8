"""DCRTPoly::AutomorphismTransform on the B200 (csrc/automorph.cu) vs the oracle
restatement of poly-impl.h:312-380 / nbtheory2.cpp:274-286 and the reference's golden digests. Bit-exact."""
import hashlib
import json
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))
EVALUATION, COEFFICIENT = 0, 1
with open(os.path.join(HERE, "golden", "ext_vectors.json")) as f:
    EXT = json.load(f)


def digest(a):
    return hashlib.sha256(np.ascontiguousarray(a, np.uint64).tobytes()).hexdigest()


@pytest.mark.parametrize("n,towers,batch", [(1, 2, 2), (2, 3, 1), (4, 1, 3), (8, 2, 2), (256, 3, 2), (1 << 11, 2, 3),
                                            (1 << 14, 4, 2), (1 << 16, 2, 1)])
def test_automorphism_vs_oracle(pkg, backend, oracle, n, towers, batch):
    if n >= 8:
        q, psi = oracle.gen_params(60, n, towers)
    else:                                       # no NTT needed: any moduli do
        q, psi = np.array([1152921504606748673, 1152921504606683137, 97][:towers], np.uint64), None
    ring = pkg.Ring(backend, n, q, psi)
    octx = oracle.ctx(n, q, psi if psi is not None else np.ones(towers, np.uint64))
    a = oracle.fill_ring(batch, n, q, 77 + n)
    a.reshape(-1)[:: max(1, a.size // 7)] = 0    # negated zeros must come back as q, as in the reference
    for k in sorted({1, 3, 5, 2 * n - 1, (n + 1) | 1, (5 ** 7) % (2 * n) | 1}):
        for fmt in (EVALUATION, COEFFICIENT):
            d = pkg.DCRTPolyBatch(ring, batch, fmt, a)
            r = d.AutomorphismTransform(k)
            assert r.format == fmt
            assert np.array_equal(r.to_host(), octx.automorphism(a, k, fmt)), (k, fmt)
            assert np.array_equal(d.to_host(), a), "source was modified"
        # two-argument form with the PrecomputeAutoMap table resident in HBM
        m = np.empty(n, np.uint32)
        pkg._native.check(pkg._native.lib.fhe_precompute_automap(n, k, m.ctypes.data))
        assert np.array_equal(m, oracle.automap(n, k))
        dm = backend.alloc(max(n * 4, 8))
        pkg._native.check(pkg._native.lib.fhe_h2d(backend.h, dm, m.ctypes.data, n * 4, 0))
        d = pkg.DCRTPolyBatch(ring, batch, EVALUATION, a)
        assert np.array_equal(d.AutomorphismTransform(k, precomp=dm).to_host(), octx.automorphism(a, k, 0))
        backend.free(dm)
    ring.close()


def test_automorphism_errors(pkg, backend, oracle):
    q, psi = oracle.gen_params(60, 16, 2)
    ring = pkg.Ring(backend, 16, q, psi)
    d = pkg.DCRTPolyBatch(ring, 1, EVALUATION, oracle.fill_ring(1, 16, q, 1))
    with pytest.raises(RuntimeError, match="not odd"):
        d.AutomorphismTransform(4)
    N = pkg._native
    assert N.lib.fhe_automorphism(ring.h, d.dptr, d.dptr, 1, 3, 0, 0) == N.FHE_EINVAL      # in place is refused
    assert N.lib.fhe_automorphism(ring.h, d.dptr, d.dptr + 8, 1, 2, 0, 0) == N.FHE_EINVAL  # even index
    assert b"not odd" in N.lib.fhe_last_error()
    del d
    ring.close()


@pytest.mark.parametrize("case", EXT["automorphism"], ids=lambda c: f"N{c['N']}_T{c['towers']}")
def test_automorphism_reference_golden(pkg, backend, oracle, case):
    n, towers, batch = case["N"], case["towers"], case["batch"]
    q, psi = np.array(case["q"], np.uint64), np.array(case["psi"], np.uint64)
    ring = pkg.Ring(backend, n, q, psi)
    a = oracle.fill_ring(batch, n, q, case["seed"])
    a[0, :, :2] = 0
    assert digest(a) == case["sha_a"]
    for ks, e in case["k"].items():
        for fmt, name in ((EVALUATION, "eval"), (COEFFICIENT, "coef")):
            r = pkg.DCRTPolyBatch(ring, batch, fmt, a).AutomorphismTransform(int(ks)).to_host()
            assert digest(r) == e["sha_" + name], (ks, name)
    ring.close()


def test_automorphism_composition_and_ntt_commute(pkg, backend, oracle):
    """Properties at a full-size ring: sigma_k1 . sigma_k2 = sigma_(k1 k2 mod 2N), sigma_k^-1 undoes sigma_k, and the
    EVALUATION-format permutation equals NTT(sigma_k(INTT(x))) (non-zero coefficients: the reference's COEFFICIENT
    form leaves a negated zero as q)."""
    n, towers, batch = 1 << 16, 3, 2
    q, psi = oracle.gen_params(60, n, towers)
    ring = pkg.Ring(backend, n, q, psi)
    x = pkg.DCRTPolyBatch(ring, batch, EVALUATION)
    ring.fill_uniform(x.dptr, batch, 4242)
    k1, k2 = 5, 2 * n - 1
    a = x.AutomorphismTransform(k1).AutomorphismTransform(k2).to_host()
    b = x.AutomorphismTransform((k1 * k2) % (2 * n)).to_host()
    assert np.array_equal(a, b)
    kinv = pow(k1, -1, 2 * n)
    assert np.array_equal(x.AutomorphismTransform(k1).AutomorphismTransform(kinv).to_host(), x.to_host())
    c = x.clone()
    c.SwitchFormat()                                     # -> COEFFICIENT
    ch = c.to_host()
    if (ch == 0).any():
        pytest.skip("a zero coefficient: q - 0 = q is not canonical in the reference's COEFFICIENT form")
    cc = c.AutomorphismTransform(k1)
    cc.SwitchFormat()
    assert np.array_equal(cc.to_host(), x.AutomorphismTransform(k1).to_host())
    ring.close()
