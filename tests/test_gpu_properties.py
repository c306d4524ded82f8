"""Size-independent properties at BASELINE.json's full sizes (where the CPU oracle would take minutes):
round trips, commutativity, linearity of the transform, product/transform consistency, and digests
that must agree between independent routes.  Everything stays on the device; only 16-byte digests
and small samples cross PCIe.  Samples of the full-size results are also checked against the oracle.
"""
import importlib

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ADD, SUB, MUL = 0, 1, 2


class Dev:
    def __init__(self, pkg, n, towers, batch, opts="arena_mb=2048"):
        self.pkg, self.n, self.towers, self.batch = pkg, n, towers, batch
        self.be = pkg.Backend(0, opts)
        self.q, self.psi = pkg.gen_params(60, n, towers)
        self.ring = pkg.Ring(self.be, n, self.q, self.psi)
        self.words = self.ring.words(batch)
        self.bufs = []

    def buf(self, seed=None):
        p = self.be.alloc(self.words * 8)
        self.bufs.append(p)
        if seed is not None:
            self.ring.fill_uniform(p, self.batch, seed)
        return p

    def digest(self, p):
        return self.be.checksum(p, self.words)

    def sample(self, p, b, t):
        """one tower of one batch item, to the host"""
        off = ((b * self.towers + t) * self.n) * 8
        return self.be.d2h(p + off, self.n)

    def close(self):
        for p in self.bufs:
            self.be.free(p)
        self.ring.close()
        self.be.close()


def check_common(d, oracle):
    ring, be, B = d.ring, d.be, d.batch
    a, b, c, t1, t2 = d.buf(11), d.buf(12), d.buf(13), d.buf(), d.buf()
    da, db = d.digest(a), d.digest(b)
    # (a + b) - b == a ; a + b == b + a
    ring.op(ADD, t1, a, b, B); ring.op(ADD, t2, b, a, B)
    assert d.digest(t1) == d.digest(t2)
    ring.op(SUB, t1, t1, b, B)
    assert d.digest(t1) == da
    # a * b == b * a ; (a + b) * c == a*c + b*c
    ring.op(MUL, t1, a, b, B); ring.op(MUL, t2, b, a, B)
    assert d.digest(t1) == d.digest(t2)
    ring.op(ADD, t1, a, b, B); ring.op(MUL, t1, t1, c, B)
    ring.op(MUL, t2, a, c, B)
    a_c = d.digest(t2)
    ring.op(MUL, c, b, c, B)                     # c <- b*c
    ring.op(ADD, t2, t2, c, B)
    assert d.digest(t1) == d.digest(t2) and a_c != d.digest(t2)
    # inputs untouched, values in range
    assert d.digest(a) == da and d.digest(b) == db
    # forward then inverse transform is the identity; transform is linear
    ring.op(ADD, t1, a, b, B)                    # t1 = a + b (coefficient domain)
    ring.ntt(t1, B, False)
    ring.ntt(a, B, False); ring.ntt(b, B, False)
    ring.op(ADD, t2, a, b, B)
    assert d.digest(t1) == d.digest(t2)
    ring.ntt(a, B, True); ring.ntt(b, B, True)
    assert d.digest(a) == da and d.digest(b) == db
    # a sample of the full-size results against the oracle (last batch item, last tower)
    bi, ti = B - 1, d.towers - 1
    o = oracle.ctx(d.n, d.q[ti:ti + 1], d.psi[ti:ti + 1])
    xa, xb = d.sample(a, bi, ti), d.sample(b, bi, ti)
    first = (bi * d.towers + ti) * d.n
    assert np.array_equal(xa, oracle.fill_uniform(d.n, d.q[ti], 11, first))
    ring.op(MUL, t1, a, b, B)
    assert np.array_equal(d.sample(t1, bi, ti), o.elementwise(MUL, xa, xb).reshape(-1))
    ring.ntt(t1, B, False)
    assert np.array_equal(d.sample(t1, bi, ti), o.ntt(o.elementwise(MUL, xa, xb), inverse=False).reshape(-1))


def test_config2_full_size(pkg, oracle):
    """BASELINE configs[1]: N=2^14, 8 towers, batch 256 (268 MB per operand)."""
    d = Dev(pkg, 1 << 14, 8, 256)
    try:
        check_common(d, oracle)
    finally:
        d.close()


def test_config3_shape_ntt_round_trip(pkg, oracle):
    """BASELINE configs[2]: N=2^16, 30 towers; batch 256 of the 1024 (4 GB in place per buffer)."""
    d = Dev(pkg, 1 << 16, 30, 256)
    try:
        check_common(d, oracle)
    finally:
        d.close()


def test_config5_pipeline_is_negacyclic_product(pkg, oracle):
    """BASELINE configs[4]: N=2^15, 16 towers.  transform=1 pipeline of ct_multiply against the same
    product built from the separate calls, and a sample against the oracle's convolution route."""
    d = Dev(pkg, 1 << 15, 16, 64)
    try:
        ring, B = d.ring, d.batch
        a0, a1, b0, b1 = d.buf(1), d.buf(2), d.buf(3), d.buf(4)
        c = [d.buf(), d.buf(), d.buf()]
        e = [d.buf(), d.buf(), d.buf()]
        # separate calls: poly_mul transforms its inputs in place, so feed it copies via ADD with zero shift
        ring.ct_multiply(c, [a0, a1], [b0, b1], B, 1)          # inputs are now in EVALUATION format
        ring.op(MUL, e[0], a0, b0, B); ring.op(MUL, e[2], a1, b1, B)
        ring.op(MUL, e[1], a1, b0, B); t = d.buf(); ring.op(MUL, t, a0, b1, B); ring.op(ADD, e[1], e[1], t, B)
        for p in e:
            ring.ntt(p, B, True)
        for x, y in zip(c, e):
            assert d.digest(x) == d.digest(y)
        # sample: c0 = INTT(NTT(a0) .* NTT(b0)) for one tower through the oracle
        bi, ti = B // 2, 5
        first = (bi * d.towers + ti) * d.n
        o = oracle.ctx(d.n, d.q[ti:ti + 1], d.psi[ti:ti + 1])
        xa, xb = oracle.fill_uniform(d.n, d.q[ti], 1, first), oracle.fill_uniform(d.n, d.q[ti], 3, first)
        exp = o.ntt(o.elementwise(MUL, o.ntt(xa), o.ntt(xb)), inverse=True).reshape(-1)
        assert np.array_equal(d.sample(c[0], bi, ti), exp)
    finally:
        d.close()


def test_device_checksum_matches_host(pkg, backend):
    sh = importlib.import_module("upmem--openfhe_b200.sharding")
    x = np.random.default_rng(3).integers(0, 2**64 - 1, size=100003, dtype=np.uint64)
    p = backend.alloc(x.nbytes)
    backend.h2d(p, x)
    assert backend.checksum(p, x.size) == sh.checksum(x)
    assert backend.checksum(p, 0) == (0, 0)
    backend.free(p)


def test_concurrent_callers_on_different_streams(pkg, backend, oracle):
    """SURVEY §8(b) "Threading": the reference's callers are OpenMP loops, so the C ABI must accept concurrent calls on
    different buffers (locked arena, per-call stream).  Eight threads allocate, run add / mul / NTT round trips on their
    own stream index and free; every result is checked."""
    import threading
    n, towers, batch = 1 << 12, 3, 4
    q, psi = oracle.gen_params(60, n, towers)
    octx = oracle.ctx(n, q, psi)
    ring = pkg.Ring(backend, n, q, psi)
    nstreams = pkg._native.lib.fhe_num_streams(backend.h)
    errors = []

    def worker(tid):
        try:
            st = tid % nstreams
            for rep in range(6):
                a = oracle.fill_ring(batch, n, q, 100 + 10 * tid + rep)
                b = oracle.fill_ring(batch, n, q, 500 + 10 * tid + rep)
                da = pkg.DCRTPolyBatch(ring, batch, pkg.EVALUATION, a, stream=st)
                db = pkg.DCRTPolyBatch(ring, batch, pkg.EVALUATION, b, stream=st)
                dc = da * db
                dc += da
                dc.SwitchFormat()
                dc.SwitchFormat()
                want = octx.elementwise(0, octx.elementwise(2, a, b), a)
                if not np.array_equal(dc.to_host(), want):
                    errors.append((tid, rep))
                del da, db, dc
        except Exception as ex:                       # noqa: BLE001
            errors.append((tid, repr(ex)))

    threads = [threading.Thread(target=worker, args=(t,)) for t in range(8)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    assert not errors, errors
    ring.close()
