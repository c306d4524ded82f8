"""BEHZ on the device (fhe_behz_*, csrc/behz.cu + the basis kernels) against the CPU oracle, which
tests/test_oracle_vs_ref.py::test_behz_matches_reference pins against the reference's own DCRTPoly
(dcrtpoly-impl.h:2005-2050, 2069-2420; tables bfvrns-cryptoparameters.cpp:662-885) — bit exact."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def behz_ring(oracle, nbits, n, nq, nB, msk_bits=0):
    q, psi = oracle.gen_params(nbits, n, nq + nB + 1)
    if msk_bits:
        q[-1] = oracle.L.orc_next_prime(oracle.L.orc_first_prime(msk_bits, 2 * n), 2 * n)
        psi[-1] = oracle.L.orc_root_of_unity(2 * n, int(q[-1]))
    return q, psi


def get(backend, dptr, shape):
    return backend.d2h(dptr, int(np.prod(shape))).reshape(shape)


# (log N, bits, nq, |B|, t, msk bits, batch): the scheme's shape |B| = nq at several sizes (below and above the ring size
# from which the conversions run on the tensor cores), msk above every q_i, one tower of Q, more towers of Q than of B
CASES = [(6, 60, 3, 3, 65537, 0, 2), (5, 45, 2, 2, 2, 0, 3), (10, 30, 4, 4, 1 << 20, 0, 2), (8, 40, 2, 2, 65537, 50, 2),
         (9, 55, 1, 1, 786433, 0, 1), (12, 60, 5, 2, 65537, 0, 2), (13, 60, 8, 8, 65537, 0, 2), (14, 55, 4, 4, 1032193, 0, 1),
         (16, 60, 3, 3, 65537, 0, 1)]


@pytest.mark.parametrize("logn,nbits,nq,nB,t,msk_bits,batch", CASES)
def test_behz_vs_oracle(pkg, backend, oracle, logn, nbits, nq, nB, t, msk_bits, batch):
    n, nb = 1 << logn, nB + 1
    T = nq + nb
    q, psi = behz_ring(oracle, nbits, n, nq, nB, msk_bits)
    ob = oracle.behz(n, q, psi, nq, t)
    ring = pkg.Ring(backend, n, q, psi)
    z = pkg.Behz(ring, nq, t)
    for k in range(z.NTABLES):
        assert np.array_equal(z.table(k), ob.table(k)), f"table {k}"
    x = oracle.fill_ring(batch, n, q[:nq].copy(), 411)
    x[0, :, 0] = q[:nq] - np.uint64(1)
    x[0, :, 1] = 0
    x[0, :, 2] = 1
    dx, dq = backend.alloc(x.nbytes), backend.alloc(x.nbytes)
    dbig = backend.alloc(batch * T * n * 8)
    backend.h2d(dx, x)
    for fmt in (0, 1):
        z.q_to_bsk_montgomery(dbig, dx, batch, fmt)
        assert np.array_equal(get(backend, dbig, (batch, T, n)), ob.q_to_bsk_montgomery(x, fmt)), f"FastBaseConvqToBskMontgomery, format {fmt}"
        assert np.array_equal(get(backend, dx, x.shape), x), "input modified"
    y = oracle.fill_ring(batch, n, q, 412)
    y[0, :, 0] = q - np.uint64(1)
    y[0, :, 3] = 0
    want = ob.fast_rns_floor_q(y)
    backend.h2d(dbig, y)
    z.fast_rns_floor_q(dbig, batch, update_q=False)
    got = get(backend, dbig, (batch, T, n))
    assert np.array_equal(got[:, nq:], want[:, nq:]), "FastRNSFloorq, towers of Bsk"
    assert np.array_equal(got[:, :nq], y[:, :nq]), "update_q = 0 leaves the towers of Q alone"
    backend.h2d(dbig, y)
    z.fast_rns_floor_q(dbig, batch)
    assert np.array_equal(get(backend, dbig, (batch, T, n)), want), "FastRNSFloorq as the reference leaves the poly"
    z.fast_base_conv_sk(dq, dbig, batch)
    assert np.array_equal(get(backend, dq, x.shape), ob.fast_base_conv_sk(want)), "FastBaseConvSK"
    assert np.array_equal(get(backend, dbig, (batch, T, n)), want), "FastBaseConvSK modified its input"
    dout = backend.alloc(batch * n * 8)
    z.scale_and_round(dout, dx, batch)
    out = get(backend, dout, (batch, n))
    assert np.array_equal(out, ob.scale_and_round(x)), "ScaleAndRound(moduliQ, t, tgamma, ...)"
    assert (out <= t).all()
    for pbuf in (dx, dq, dbig, dout):
        backend.free(pbuf)
    z.close()
    ring.close()


def oracle_behz_mult(oracle, ob, octx, a, b):
    """bfvrns-leveledshe.cpp:275-297, 322-334, 367-388 on the oracle: expand, tensor product over QBsk, scale, convert back."""
    e = [ob.q_to_bsk_montgomery(v, 0) for v in (a[0], a[1], b[0], b[1])]
    c0 = octx.elementwise(2, e[0], e[2])
    c1 = octx.elementwise(0, octx.elementwise(2, e[0], e[3]), octx.elementwise(2, e[1], e[2]))
    c2 = octx.elementwise(2, e[1], e[3])
    return [ob.fast_base_conv_sk(ob.fast_rns_floor_q(octx.ntt(c, inverse=True))) for c in (c0, c1, c2)]


@pytest.mark.parametrize("logn,nbits,nq,t,batch", [(8, 60, 3, 65537, 2), (12, 55, 4, 65537, 2), (13, 60, 6, 786433, 1)])
def test_bfv_behz_mult_core_vs_oracle(pkg, backend, oracle, logn, nbits, nq, t, batch):
    """The BEHZ EvalMult core as one device composition (pkg.BfvBehzMult) against the same sequence on the oracle."""
    n = 1 << logn
    q, psi = behz_ring(oracle, nbits, n, nq, nq)
    ob, octx = oracle.behz(n, q, psi, nq, t), oracle.ctx(n, q, psi)
    ring = pkg.Ring(backend, n, q, psi)
    mult = pkg.BfvBehzMult(ring, nq, t)
    ins = [oracle.fill_ring(batch, n, q[:nq].copy(), 421 + k) for k in range(4)]
    din = [backend.alloc(v.nbytes) for v in ins]
    for dv, v in zip(din, ins):
        backend.h2d(dv, v)
    dout = [backend.alloc(ins[0].nbytes) for _ in range(3)]
    want = oracle_behz_mult(oracle, ob, octx, ins[0:2], ins[2:4])
    for _ in range(2):                                  # twice: scratch buffers are reused
        mult(dout, din[0:2], din[2:4], batch)
        for k in range(3):
            assert np.array_equal(get(backend, dout[k], ins[0].shape), want[k]), f"c{k}"
    for pbuf in din + dout:
        backend.free(pbuf)
    mult.close()
    ring.close()


def test_behz_decryption_recovers_the_message(pkg, backend, oracle):
    """Meaning, not only agreement: for x = round(Q/t * m) + e (a fresh BFV phase with small noise) the BEHZ
    ScaleAndRound returns m modulo t."""
    n, nq, t, batch = 1 << 10, 3, 65537, 2
    q, psi = behz_ring(oracle, 50, n, nq, nq)
    ring = pkg.Ring(backend, n, q, psi)
    z = pkg.Behz(ring, nq, t)
    Q = 1
    for m in q[:nq]:
        Q *= int(m)
    rng = np.random.default_rng(5)
    msg = rng.integers(0, t, size=(batch, n))
    noise = rng.integers(-1000, 1001, size=(batch, n))
    x = np.empty((batch, nq, n), np.uint64)
    for b in range(batch):
        for k in range(n):
            v = ((Q * int(msg[b, k]) + t // 2) // t + int(noise[b, k])) % Q
            for i in range(nq):
                x[b, i, k] = v % int(q[i])
    dx, dout = backend.alloc(x.nbytes), backend.alloc(batch * n * 8)
    backend.h2d(dx, x)
    z.scale_and_round(dout, dx, batch)
    out = get(backend, dout, (batch, n))
    assert np.array_equal(out % np.uint64(t), msg.astype(np.uint64))
    backend.free(dx); backend.free(dout)
    z.close()
    ring.close()


def test_behz_argument_checks(pkg, backend, oracle):
    n = 64
    q, psi = oracle.gen_params(60, n, 5)
    ring = pkg.Ring(backend, n, q, psi)
    for size_q, t in ((0, 65537), (4, 65537), (5, 65537), (2, 1), (2, 1 << 33)):
        with pytest.raises(pkg.FheError):
            pkg.Behz(ring, size_q, t)
    r2 = pkg.Ring(backend, n, np.array([q[0], q[1], 1 << 20, q[3]], np.uint64), None)      # an even modulus in Bsk
    with pytest.raises(pkg.FheError):
        pkg.Behz(r2, 2, 65537)
    r2.close()
    z = pkg.Behz(ring, 2, 65537)
    x = backend.alloc(5 * n * 8 + 16)
    with pytest.raises(pkg.FheError):
        z.q_to_bsk_montgomery(x, x, 1, 0)               # dst aliases src
    with pytest.raises(pkg.FheError):
        z.q_to_bsk_montgomery(x, x + 8, 1, 0)           # misaligned
    with pytest.raises(pkg.FheError):
        z.fast_base_conv_sk(x, x, 1)
    z.fast_rns_floor_q(x, 0)                            # an empty batch is a no-op
    backend.free(x)
    z.close()
    ring.close()
