"""GPU parity: the CUDA path, called through the C ABI, against the CPU oracle — bit exact.

Shapes follow the reference's own tests (tiny rings with fixed vectors) and
BASELINE.json's configs (N = 2^14..2^16 with 60-bit moduli) at batch sizes the oracle finishes in
seconds; full-size configs are covered by size-independent properties in test_gpu_properties.py.
"""
import ctypes

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ADD, SUB, MUL = 0, 1, 2


def rand_polys(rng, batch, n, q):
    return np.stack([rng.integers(0, int(qt), size=(batch, n), dtype=np.uint64) for qt in q], axis=1)


def params(oracle, nbits, n, towers):
    return oracle.gen_params(nbits, n, towers)


# ---- reference KATs replayed on the GPU --------------------------------------------------------
def test_kat_crt_polynomial_mult(pkg, backend):
    """UnitTestTransform.cpp:59-98: q=113, N=4, a=b=[1,2,4,1] -> [94,109,11,18]."""
    ring = pkg.Ring(backend, 4, [113], [18])
    a = pkg.DCRTPolyBatch(ring, 1, pkg.COEFFICIENT, np.array([1, 2, 4, 1], np.uint64))
    b = a.clone()
    a.SwitchFormat()
    assert a.to_host().reshape(-1).tolist() == [46, 62, 104, 18]
    b.SwitchFormat()
    c = a * b
    c.SwitchFormat()
    assert c.to_host().reshape(-1).tolist() == [94, 109, 11, 18]
    ring.close()


def test_kat_dcrt_arithmetic(pkg, backend):
    """UnitTestDCRTElements.cpp:285-406: 3 towers {8353,8369,8513}, N=4."""
    q = [8353, 8369, 8513]
    ring = pkg.Ring(backend, 4, q)
    a = pkg.DCRTPolyBatch(ring, 1, pkg.EVALUATION, np.tile(np.array([2, 4, 3, 2], np.uint64), 3))
    b = pkg.DCRTPolyBatch(ring, 1, pkg.EVALUATION, np.tile(np.array([2, 1, 2, 0], np.uint64), 3))
    for res, exp in [(a + b, [4, 5, 5, 2]), (a - b, [0, 3, 1, 2]), (a * b, [4, 4, 6, 0])]:
        assert res.to_host().reshape(3, 4).tolist() == [exp] * 3
    c = a.clone(); c += b
    assert c.to_host().reshape(3, 4).tolist() == [[4, 5, 5, 2]] * 3
    c = a.clone(); c -= b
    assert c.to_host().reshape(3, 4).tolist() == [[0, 3, 1, 2]] * 3
    c = a.clone(); c *= b
    assert c.to_host().reshape(3, 4).tolist() == [[4, 4, 6, 0]] * 3
    ring.close()


def test_times_requires_evaluation(pkg, backend):
    """dcrtpoly.h:188-189: Times throws outside EVALUATION format."""
    ring = pkg.Ring(backend, 4, [113], [18])
    a = pkg.DCRTPolyBatch(ring, 1, pkg.COEFFICIENT, np.array([1, 2, 4, 1], np.uint64))
    with pytest.raises(RuntimeError):
        a.Times(a)
    ring.close()


# ---- element-wise ops --------------------------------------------------------------------------
@pytest.mark.parametrize("n,towers,batch,nbits", [
    (4, 3, 1, 14), (8, 2, 5, 28), (2, 1, 1, 20), (1024, 3, 2, 60), (2048, 2, 3, 60),
    (1 << 14, 2, 2, 60), (1 << 14, 8, 4, 60), (1 << 15, 4, 1, 59), (1 << 16, 3, 2, 60), (1 << 12, 5, 7, 50),
])
def test_elementwise_vs_oracle(pkg, backend, oracle, n, towers, batch, nbits):
    q, psi = params(oracle, nbits, n, towers)
    octx = oracle.ctx(n, q, psi)
    ring = pkg.Ring(backend, n, q)
    rng = np.random.default_rng(1000 + n + towers)
    a, b = rand_polys(rng, batch, n, q), rand_polys(rng, batch, n, q)
    # edge vectors: zeros, q-1, alternating
    a[0, :, : min(n, 8)] = 0
    b[0, :, : min(n, 8)] = (q - 1)[:, None]
    a[-1, :, -min(n, 8):] = (q - 1)[:, None]
    b[-1, :, -min(n, 8):] = (q - 1)[:, None]
    da = pkg.DCRTPolyBatch(ring, batch, pkg.EVALUATION, a)
    db = pkg.DCRTPolyBatch(ring, batch, pkg.EVALUATION, b)
    for op, res in [(ADD, da + db), (SUB, da - db), (MUL, da * db)]:
        exp = octx.elementwise(op, a, b)
        assert np.array_equal(res.to_host(), exp), f"op {op}"
    for op in (ADD, SUB, MUL):                       # in-place forms
        c = da.clone()
        if op == ADD: c += db
        elif op == SUB: c -= db
        else: c *= db
        assert np.array_equal(c.to_host(), octx.elementwise(op, a, b))
    ring.close()


# ---- tables ------------------------------------------------------------------------------------
@pytest.mark.parametrize("n,nbits", [(4, 10), (256, 30), (1 << 12, 60), (1 << 16, 60)])
def test_twiddle_tables_vs_oracle(pkg, backend, oracle, n, nbits):
    q, psi = params(oracle, nbits, n, 2)
    ring = pkg.Ring(backend, n, q, psi)
    for t in range(2):
        W, Wp, WI, WIp, _, _ = oracle.precompute(psi[t], n, q[t])
        for which, exp in enumerate([W, Wp, WI, WIp]):
            assert np.array_equal(ring.table(t, which), exp)
    ring.close()


# ---- NTT ---------------------------------------------------------------------------------------
@pytest.mark.parametrize("logn,towers,batch,nbits", [
    (1, 1, 1, 10), (2, 2, 3, 12), (3, 2, 1, 22), (5, 1, 2, 30), (8, 3, 2, 60),
    (9, 2, 3, 60), (10, 1, 1, 60), (11, 3, 17, 55), (12, 2, 16, 60), (13, 2, 5, 60),
    (14, 2, 2, 60), (14, 4, 33, 60), (15, 3, 4, 60), (16, 2, 3, 60), (16, 3, 16, 60), (16, 1, 1, 60),
    # two-pass kernels with short moduli (range tracking and the float quotient estimate at every size)
    (9, 2, 4, 14), (10, 2, 3, 20), (11, 2, 2, 25), (12, 2, 2, 33), (13, 1, 2, 41), (14, 1, 2, 48), (16, 1, 1, 21),
])
def test_ntt_vs_oracle(pkg, backend, oracle, logn, towers, batch, nbits):
    n = 1 << logn
    q, psi = params(oracle, nbits, n, towers)
    octx = oracle.ctx(n, q, psi)
    ring = pkg.Ring(backend, n, q, psi)
    rng = np.random.default_rng(7 + logn * 31 + batch)
    x = rand_polys(rng, batch, n, q)
    x[0, :, 0] = (q - 1)
    x[-1, :, -1] = 0
    d = pkg.DCRTPolyBatch(ring, batch, pkg.COEFFICIENT, x)
    d.SwitchFormat()
    fwd = d.to_host()
    exp = octx.ntt(x, inverse=False)
    assert np.array_equal(fwd, exp), "forward"
    d.SwitchFormat()
    assert np.array_equal(d.to_host(), x), "round trip"
    # inverse on its own, from arbitrary evaluation-domain data
    y = rand_polys(rng, batch, n, q)
    e = pkg.DCRTPolyBatch(ring, batch, pkg.EVALUATION, y)
    e.SwitchFormat()
    assert np.array_equal(e.to_host(), octx.ntt(y, inverse=True)), "inverse"
    ring.close()


@pytest.mark.parametrize("logn,nbits", [(9, 60), (12, 60), (14, 60), (16, 60), (16, 59), (13, 30)])
def test_ntt_extreme_values(pkg, backend, oracle, logn, nbits):
    """Inputs that push the lazily reduced intermediates to the top of their ranges: all q-1, alternating
    q-1 / 0, a single q-1 — forward and inverse, against the oracle."""
    n = 1 << logn
    q, psi = params(oracle, nbits, n, 2)
    octx = oracle.ctx(n, q, psi)
    ring = pkg.Ring(backend, n, q, psi)
    x = np.zeros((4, 2, n), np.uint64)
    x[0] = (q - 1)[:, None]
    x[1, :, ::2] = (q - 1)[:, None]
    x[2, :, 1::2] = (q - 1)[:, None]
    x[3, :, n // 2] = q - 1
    for fmt, inv in ((pkg.COEFFICIENT, False), (pkg.EVALUATION, True)):
        d = pkg.DCRTPolyBatch(ring, 4, fmt, x)
        d.SwitchFormat()
        got = d.to_host()
        assert np.array_equal(got, octx.ntt(x, inverse=inv))
        assert (got < q[None, :, None]).all()
    ring.close()


@pytest.mark.parametrize("logn", [9, 10, 12])
def test_ntt_small_kernel_matches_two_pass(pkg, backend, oracle, logn):
    n = 1 << logn
    q, psi = params(oracle, 60, n, 2)
    ring = pkg.Ring(backend, n, q, psi)
    x = rand_polys(np.random.default_rng(3), 3, n, q)
    d1 = pkg.DCRTPolyBatch(ring, 3, pkg.COEFFICIENT, x)
    d2 = d1.clone()
    ring.ntt(d1.dptr, 3, inverse=False, algo=0)
    ring.ntt(d2.dptr, 3, inverse=False, algo=1)
    assert np.array_equal(d1.to_host(), d2.to_host())
    ring.ntt(d1.dptr, 3, inverse=True, algo=1)
    ring.ntt(d2.dptr, 3, inverse=True, algo=0)
    assert np.array_equal(d1.to_host(), x) and np.array_equal(d2.to_host(), x)
    ring.close()


def test_trivial_root_is_identity(pkg, backend):
    """transformnat-impl.h:580-582: psi in {0,1} leaves the element untouched."""
    ring = pkg.Ring(backend, 8, [113, 97], [1, 0])
    x = np.arange(16, dtype=np.uint64)
    d = pkg.DCRTPolyBatch(ring, 1, pkg.COEFFICIENT, x)
    d.SwitchFormat()
    assert np.array_equal(d.to_host().reshape(-1), x)
    ring.close()


def test_poly_mul_is_negacyclic_convolution(pkg, backend, oracle):
    n, towers = 64, 2
    q, psi = params(oracle, 40, n, towers)
    ring = pkg.Ring(backend, n, q, psi)
    rng = np.random.default_rng(11)
    a, b = rand_polys(rng, 2, n, q), rand_polys(rng, 2, n, q)
    da = pkg.DCRTPolyBatch(ring, 2, pkg.COEFFICIENT, a)
    db = pkg.DCRTPolyBatch(ring, 2, pkg.COEFFICIENT, b)
    dc = pkg.DCRTPolyBatch(ring, 2, pkg.COEFFICIENT)
    ring.poly_mul(dc.dptr, da.dptr, db.dptr, 2)
    got = dc.to_host()
    for bi in range(2):
        for t in range(towers):
            qq = int(q[t])
            exp = [0] * n
            for i in range(n):
                for j in range(n):
                    k, v = i + j, int(a[bi, t, i]) * int(b[bi, t, j])
                    if k >= n: exp[k - n] = (exp[k - n] - v) % qq
                    else: exp[k] = (exp[k] + v) % qq
            assert got[bi, t].tolist() == exp
    ring.close()


# ---- flat vectors (PimData surface) --------------------------------------------------------------
@pytest.mark.parametrize("n", [1, 2, 3, 255, 1024, 4097, 100001])
@pytest.mark.parametrize("q", [73, 8353, 1152921504606748673])
def test_vec_ops_vs_oracle(pkg, backend, oracle, n, q):
    rng = np.random.default_rng(n)
    a = rng.integers(0, q, size=n, dtype=np.uint64)
    b = rng.integers(0, q, size=n, dtype=np.uint64)
    da, db, dd = backend.alloc(n * 8), backend.alloc(n * 8), backend.alloc(n * 8)
    backend.h2d(da, a); backend.h2d(db, b)
    for op in (ADD, SUB, MUL):
        backend.vec_op(op, dd, da, db, n, q)
        assert np.array_equal(backend.d2h(dd, n), oracle.vec(op, a, b, q))
        s = int(rng.integers(0, 2**63))
        backend.vec_op_scalar(op, dd, da, s, n, q)
        assert np.array_equal(backend.d2h(dd, n), oracle.vec_scalar(op, a, s, q))
    for p in (da, db, dd):
        backend.free(p)


# ---- host-operand (end-to-end) entry points -------------------------------------------------------
def test_host_pipeline_vs_oracle(pkg, backend, oracle):
    n, towers, batch = 1 << 14, 4, 80           # 80 items of 512 KB -> several 32 MB chunks
    q, psi = params(oracle, 60, n, towers)
    octx = oracle.ctx(n, q, psi)
    ring = pkg.Ring(backend, n, q, psi)
    rng = np.random.default_rng(5)
    a, b = rand_polys(rng, batch, n, q), rand_polys(rng, batch, n, q)
    out = np.empty_like(a)
    for op in (ADD, MUL):
        ring.op_host(op, out, a, b, batch)
        assert np.array_equal(out, octx.elementwise(op, a, b))
    x = a.copy()
    ring.ntt_host(x, batch, inverse=False)
    assert np.array_equal(x, octx.ntt(a, inverse=False))
    ring.ntt_host(x, batch, inverse=True)
    assert np.array_equal(x, a)
    # the reference's in-place sequence a += b; a *= b; a -= b (dcrtpoly-impl.h:398-416) on host operands
    x = a.copy()
    ring.chain_host([ADD, MUL, SUB], x, b, batch)
    want = octx.elementwise(SUB, octx.elementwise(MUL, octx.elementwise(ADD, a, b), b), b)
    assert np.array_equal(x, want)
    ring.close()


def test_fill_uniform_matches_oracle_generator(pkg, backend, oracle):
    n, towers, batch = 1 << 10, 3, 4
    q, psi = params(oracle, 60, n, towers)
    ring = pkg.Ring(backend, n, q)
    d = pkg.DCRTPolyBatch(ring, batch)
    ring.fill_uniform(d.dptr, batch, 20260001)
    got = d.to_host()
    assert np.array_equal(got, oracle.fill_ring(batch, n, q, 20260001))
    assert (got < q[None, :, None]).all()
    ring.close()


# ---- mod-ops of the PimData surface -------------------------------------------
@pytest.mark.parametrize("mod", [73, 8353, (1 << 40) + 15, 1152921504606748673, (1 << 63) + 29])
def test_modops_vs_oracle(pkg, backend, oracle, mod):
    n = 10007
    rng = np.random.default_rng(mod % 1000)
    x = rng.integers(0, mod, size=n, dtype=np.uint64)
    x[:4] = [mod - 1, mod >> 1, (mod >> 1) + 1, 1]
    dx, dy = backend.alloc(n * 8), backend.alloc(n * 8)
    cases = [(0, mod + 1000, 0), (0, max(3, mod // 3), 0), (1, 2, 0), (2, max(5, mod // 7), 0), (3, 0, 0), (4, 0, 0),
             (5, 65537, 0), (6, mod + 5, 0), (9, 12345, max(7, mod // 5)), (10, mod - 1, mod), (11, 7, 0),
             (12, max(2, mod // 1000), 0)]
    if mod >= (1 << 63):                       # new modulus above 2^64 is not representable; skip the growing switch
        cases = cases[1:]
    for kind, p, q in cases:
        backend.h2d(dx, x)
        rc, exp = oracle.modop(kind, x, mod, p, q)
        assert rc == 0
        backend.vec_modop(kind, dy, dx, n, mod, p, q)
        assert np.array_equal(backend.d2h(dy, n), exp), f"kind {kind}"
        backend.vec_modop(kind, dx, dx, n, mod, p, q)          # in place
        assert np.array_equal(backend.d2h(dx, n), exp), f"kind {kind} in place"
    if mod < (1 << 63):                        # ModInverse works on signed 64-bit values in the reference
        xi = x.copy()
        xi[xi == 0] = 1
        if mod in (73, 8353, 1152921504606748673):               # prime: every non-zero element is invertible
            backend.h2d(dx, xi)
            rc, exp = oracle.modop(7, xi, mod)
            assert rc == 0
            backend.vec_modop(7, dy, dx, n, mod)
            assert np.array_equal(backend.d2h(dy, n), exp)
        xi[5] = 0
        backend.h2d(dx, xi)
        with pytest.raises(pkg.FheError):
            backend.vec_modop(8, dx, dx, n, mod)
    with pytest.raises(pkg.FheError):
        backend.vec_modop(9, dy, dx, n, mod, 5, 0)             # divide by zero, as the reference throws
    backend.free(dx); backend.free(dy)


def test_cpp_pim_surface_binary():
    """include/pim/{PimManager,PimData,PimDCRTPoly}.h driven from C++ (tests/cpp/pim_surface_test.cpp)."""
    import os
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    exe = os.path.join(root, "tests", "cpp", "bin", "pim_surface_test")
    if not os.path.exists(exe):
        subprocess.check_call(["make", "-C", root, "host"])
    r = subprocess.run([exe], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "all ok" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]


# ---- ciphertext tensor product and the multiplication pipelines (BASELINE configs 4 and 5) --------------
def _oracle_tensor(octx, a0, a1, b0, b1):
    """LeveledSHEBase::EvalMultCore, 2x2 case (base-leveledshe.cpp:667-672), op by op on the oracle."""
    c2 = octx.elementwise(MUL, a1, b1)
    c1 = octx.elementwise(MUL, a1, b0)
    c0 = octx.elementwise(MUL, b0, a0)
    c1 = octx.elementwise(ADD, c1, octx.elementwise(MUL, a0, b1))
    return c0, c1, c2


@pytest.mark.parametrize("n,towers,batch,nbits", [(4, 3, 1, 14), (512, 2, 3, 30), (1 << 12, 3, 2, 60), (1 << 14, 4, 3, 60),
                                                   (1 << 16, 2, 1, 60)])
@pytest.mark.parametrize("transform", [0, 1, 2])
def test_ct_multiply_vs_oracle(pkg, backend, oracle, n, towers, batch, nbits, transform):
    q, psi = params(oracle, nbits, n, towers)
    octx = oracle.ctx(n, q, psi)
    ring = pkg.Ring(backend, n, q, psi)
    rng = np.random.default_rng(n + transform)
    host = [rand_polys(rng, batch, n, q) for _ in range(4)]
    host[0][0, :, :2] = (q - 1)[:, None]
    host[2][0, :, :2] = (q - 1)[:, None]
    fmt = pkg.EVALUATION if transform != 1 else pkg.COEFFICIENT
    dev = [pkg.DCRTPolyBatch(ring, batch, fmt, h) for h in host]
    out = [pkg.DCRTPolyBatch(ring, batch, fmt) for _ in range(3)]
    ring.ct_multiply([o.dptr for o in out], [dev[0].dptr, dev[1].dptr], [dev[2].dptr, dev[3].dptr], batch, transform)
    x = host
    if transform:
        x = [octx.ntt(h, inverse=(transform == 2)) for h in host]
    exp = _oracle_tensor(octx, *x)
    if transform:
        exp = [octx.ntt(e, inverse=(transform == 1)) for e in exp]
    for o, e in zip(out, exp):
        assert np.array_equal(o.to_host(), e)
    ring.close()


def test_ct_multiply_in_place_and_squaring(pkg, backend, oracle):
    n, towers, batch = 1 << 13, 2, 2
    q, psi = params(oracle, 60, n, towers)
    octx = oracle.ctx(n, q, psi)
    ring = pkg.Ring(backend, n, q, psi)
    rng = np.random.default_rng(99)
    a0, a1 = rand_polys(rng, batch, n, q), rand_polys(rng, batch, n, q)
    d0, d1 = pkg.DCRTPolyBatch(ring, batch, pkg.EVALUATION, a0), pkg.DCRTPolyBatch(ring, batch, pkg.EVALUATION, a1)
    d2 = pkg.DCRTPolyBatch(ring, batch, pkg.EVALUATION)
    # square a ciphertext, writing c0 over a0 and c1 over a1
    ring.ct_multiply([d0.dptr, d1.dptr, d2.dptr], [d0.dptr, d1.dptr], [d0.dptr, d1.dptr], batch, 0)
    e0, e1, e2 = _oracle_tensor(octx, a0, a1, a0, a1)
    assert np.array_equal(d0.to_host(), e0) and np.array_equal(d1.to_host(), e1) and np.array_equal(d2.to_host(), e2)
    with pytest.raises(pkg.FheError):
        ring.ct_multiply([d0.dptr, d0.dptr, d2.dptr], [d0.dptr, d1.dptr], [d0.dptr, d1.dptr], batch, 0)
    ring.close()


# ---- round 2: the bench step through the host-operand call, stream-ordered frees, aliasing forms ----------------
def test_chain_host_with_switch_format_vs_oracle(pkg, backend, oracle):
    """a += b; a *= b; a.SwitchFormat() (-> COEFFICIENT); a.SwitchFormat() (-> EVALUATION): bench.py's e2e call."""
    n, towers, batch = 1 << 14, 4, 80           # several 32 MB chunks over the stream pool
    q, psi = params(oracle, 60, n, towers)
    octx = oracle.ctx(n, q, psi)
    ring = pkg.Ring(backend, n, q, psi)
    rng = np.random.default_rng(11)
    a, b = rand_polys(rng, batch, n, q), rand_polys(rng, batch, n, q)
    prod = octx.elementwise(MUL, octx.elementwise(ADD, a, b), b)
    x = a.copy()
    ring.chain_host([ADD, MUL, pkg.OP_INTT], x, b, batch)
    assert np.array_equal(x, octx.ntt(prod, inverse=True))
    x = a.copy()
    ring.chain_host([ADD, MUL, pkg.OP_INTT, pkg.OP_NTT], x, b, batch)
    assert np.array_equal(x, prod)
    # transform-only chain: b is not read and may be absent
    x = a.copy()
    n_ops = (ctypes.c_int * 1)(pkg.OP_NTT)
    pkg._native.check(pkg._native.lib.fhe_dcrt_chain_host(ring.h, n_ops, 1, x.ctypes.data_as(ctypes.c_void_p), None, batch))
    assert np.array_equal(x, octx.ntt(a, inverse=False))
    with pytest.raises(pkg.FheError):
        ring.chain_host([5], x, b, batch)
    ring.close()


def test_free_is_stream_ordered(pkg, backend, oracle):
    """A block freed while launches that write it are still pending on another stream must not be handed out again
    before they finish (the arena used to put it straight back on the free list)."""
    n, towers, batch = 1 << 14, 8, 128          # 128 MB
    q, psi = params(oracle, 60, n, towers)
    ring = pkg.Ring(backend, n, q, psi)
    nbytes = ring.words(batch) * 8
    pattern = np.arange(4096, dtype=np.uint64) * np.uint64(0x9E3779B97F4A7C15)
    for _ in range(3):
        tmp = backend.alloc(nbytes)
        for k in range(40):                      # a few ms of pending writes to tmp on stream 2
            ring.fill_uniform(tmp, batch, 1000 + k, stream=2)
        backend.free(tmp)                        # enqueue-and-free, as the live intercept does with temporaries
        blk = backend.alloc(nbytes)              # first fit would return tmp's block
        pkg._native.check(pkg._native.lib.fhe_h2d(backend.h, blk, pattern.ctypes.data_as(ctypes.c_void_p), pattern.nbytes, 0))
        backend.sync()                           # everything, including the pending fills
        assert np.array_equal(backend.d2h(blk, 4096), pattern), "a freed block was reused while writes to it were pending"
        backend.free(blk)
    # a ring created right after such a free uploads its tables with a blocking copy: they must survive too
    tmp = backend.alloc(nbytes)
    for k in range(40):
        ring.fill_uniform(tmp, batch, 2000 + k, stream=3)
    backend.free(tmp)
    ring2 = pkg.Ring(backend, n, q, psi)
    backend.sync()
    for t in (0, towers - 1):
        assert np.array_equal(ring2.table(t, 0), ring.table(t, 0))
    ring2.close()
    with pytest.raises(pkg.FheError):
        backend.free(0xdead00)                   # not an arena block
    p = backend.alloc(256)
    ring.fill_uniform(p, 0, 1, stream=1)
    backend.free(p)
    backend.sync()
    ring.close()


@pytest.mark.parametrize("opts,grows", [("arena_mb=32,arena_grow_mb=0", False), ("arena_mb=32,arena_grow_mb=4096", True)])
def test_arena_waits_or_grows_when_full(pkg, oracle, opts, grows):
    """An allocation that finds no free block while freed blocks are still behind unfinished work: with arena_grow_mb=0 it
    waits for them (the arena stays at its first slab), below the limit it takes another slab instead; either way a block is
    never handed out while earlier writes to it are pending, from several streams, with the frees sealed in groups."""
    n, towers, batch = 1 << 12, 4, 64                 # 8 MB per block, four of them fill a 32 MB slab
    q, psi = params(oracle, 60, n, towers)
    pattern = np.arange(2048, dtype=np.uint64) * np.uint64(0xD6E8FEB86659FD93)
    with pkg.Backend(0, opts + ",streams=8") as be:
        ring = pkg.Ring(be, n, q, psi)
        nbytes = ring.words(batch) * 8
        reserved = ctypes.c_size_t()
        for rnd in range(12):
            blocks = [be.alloc(nbytes) for _ in range(3)]
            for i, blk in enumerate(blocks):
                for k in range(6):                    # pending writes on a stream of its own
                    ring.fill_uniform(blk, batch, 10 * rnd + k, stream=1 + (i + rnd) % 7)
            for blk in blocks:
                be.free(blk)
            fresh = be.alloc(nbytes)                  # the slab is full of blocks that are freed but still being written
            pkg._native.check(pkg._native.lib.fhe_h2d(be.h, fresh, pattern.ctypes.data_as(ctypes.c_void_p), pattern.nbytes, 0))
            be.sync()
            assert np.array_equal(be.d2h(fresh, 2048), pattern), f"round {rnd}: a block was reused under pending writes"
            be.free(fresh)
        pkg._native.check(pkg._native.lib.fhe_arena_stats(be.h, ctypes.byref(reserved), None))
        if grows:
            assert reserved.value >= (64 << 20), reserved.value      # a second slab was taken
        else:
            assert reserved.value < (64 << 20), reserved.value       # still the first slab: it waited instead of growing
        ring.close()


def test_misaligned_polynomial_pointers_are_refused(pkg, backend, oracle):
    """The polynomial kernels move 128-bit words: a pointer that is not 16-byte aligned is refused by the C ABI
    (FHE_EINVAL) instead of reaching a kernel (fhe_b200.h, "Alignment"); the flat vector calls accept it."""
    n, towers, batch = 1 << 10, 2, 1
    q, psi = params(oracle, 60, n, towers)
    ring = pkg.Ring(backend, n, q, psi)
    blk = backend.alloc((ring.words(batch) + 2) * 8)
    dst = backend.alloc((ring.words(batch) + 2) * 8)
    odd, odd_dst = blk + 8, dst + 8
    L = pkg._native.lib
    for rc in (L.fhe_ntt_forward(ring.h, odd, batch, 0),
               L.fhe_ntt_inverse(ring.h, odd, batch, 0),
               L.fhe_automorphism(ring.h, odd_dst, blk, batch, 5, 0, 0),
               L.fhe_automorphism(ring.h, dst, odd, batch, 5, 1, 0),
               L.fhe_dcrt_op(ring.h, pkg.OP_MUL, dst, odd, blk, batch, 0)):
        assert rc != 0
        assert b"align" in L.fhe_last_error().lower() or b"misaligned" in L.fhe_last_error().lower()
    x = np.arange(1, 102, dtype=np.uint64)
    backend.h2d(odd, x)
    backend.vec_op(pkg.OP_ADD, odd_dst, odd, odd, x.size, int(q[0]))
    assert np.array_equal(backend.d2h(odd_dst, x.size), (x + x) % q[0])
    backend.sync()
    backend.free(blk); backend.free(dst)
    ring.close()


def test_flat_vector_eq_forms_in_place(pkg, backend, oracle):
    """VECTOR_EQ / SCALAR_EQ descriptor kinds run with dst == a (pim/common.h:14-21)."""
    q = 1152921504606748673
    for n in (3, 4097, 100001):
        rng = np.random.default_rng(n)
        a = rng.integers(0, q, size=n, dtype=np.uint64)
        b = rng.integers(0, q, size=n, dtype=np.uint64)
        da, db = backend.alloc(n * 8), backend.alloc(n * 8)
        for op in (ADD, SUB, MUL):
            backend.h2d(da, a); backend.h2d(db, b)
            backend.vec_op(op, da, da, db, n, q)
            assert np.array_equal(backend.d2h(da, n), oracle.vec(op, a, b, q))
            backend.h2d(da, a)
            backend.vec_op(op, db, da, db, n, q)             # dst == b
            assert np.array_equal(backend.d2h(db, n), oracle.vec(op, a, b, q))
            backend.h2d(da, a)
            backend.vec_op_scalar(op, da, da, 12345678901234567, n, q)
            assert np.array_equal(backend.d2h(da, n), oracle.vec_scalar(op, a, 12345678901234567, q))
        backend.free(da); backend.free(db)


# ---- the warp-specialised persistent NTT kernels (csrc/ntt_ws.cu), forced with algo = 3 -------------------------
@pytest.mark.parametrize("logn,towers,batch,nbits", [
    (14, 2, 1, 60), (14, 3, 8, 60), (14, 2, 13, 48), (14, 4, 200, 60),
    (15, 2, 1, 60), (15, 3, 9, 60), (15, 2, 64, 33),
    (16, 1, 1, 60), (16, 2, 7, 60), (16, 3, 16, 60), (16, 2, 41, 21), (16, 30, 8, 60),
])
def test_ntt_ws_kernels_vs_oracle(pkg, backend, oracle, logn, towers, batch, nbits):
    """Every batch size takes the ring of slots through partial groups (batch % 8 != 0), CTAs with fewer tiles than
    slots, and (batch 200 / 64 / 41) several laps of the ring."""
    n = 1 << logn
    q, psi = params(oracle, nbits, n, towers)
    octx = oracle.ctx(n, q, psi)
    ring = pkg.Ring(backend, n, q, psi)
    rng = np.random.default_rng(100 + logn * 31 + batch)
    x = rand_polys(rng, batch, n, q)
    x[0, :, 0] = (q - 1)
    x[-1, :, -1] = 0
    x[0, 0, :] = q[0] - 1                      # a whole tower at the top of the range
    words = ring.words(batch)
    d = backend.alloc(words * 8)
    backend.h2d(d, x)
    ring.ntt(d, batch, False, algo=3)
    fwd = backend.d2h(d, words).reshape(x.shape)
    assert np.array_equal(fwd, octx.ntt(x, inverse=False)), "forward"
    backend.h2d(d, x)                          # the one-tile-per-CTA kernels agree on the same input
    ring.ntt(d, batch, False, algo=2)
    assert np.array_equal(backend.d2h(d, words).reshape(x.shape), fwd), "forward, algo 2"
    ring.ntt(d, batch, True, algo=3)
    assert np.array_equal(backend.d2h(d, words).reshape(x.shape), x), "round trip"
    y = rand_polys(rng, batch, n, q)
    backend.h2d(d, y)
    ring.ntt(d, batch, True, algo=3)
    assert np.array_equal(backend.d2h(d, words).reshape(x.shape), octx.ntt(y, inverse=True)), "inverse"
    backend.free(d)
    ring.close()


@pytest.mark.parametrize("logn,towers,batch,nbits", [(12, 3, 5, 60), (13, 2, 9, 45), (14, 3, 8, 60), (15, 2, 64, 33), (16, 2, 7, 60), (16, 30, 8, 60)])
def test_ntt_prefetching_strided_pass_vs_oracle(pkg, backend, oracle, monkeypatch, logn, towers, batch, nbits):
    """The persistent strided pass with register prefetch (ntt_strided_pf_kernel, FHE_B200_NTT_PF=2 forces it; off by
    default): CTAs walk over several tiles (more tiles than CTAs at the larger shapes), both directions."""
    monkeypatch.setenv("FHE_B200_NTT_PF", "2")
    n = 1 << logn
    q, psi = params(oracle, nbits, n, towers)
    octx = oracle.ctx(n, q, psi)
    ring = pkg.Ring(backend, n, q, psi)
    rng = np.random.default_rng(700 + logn * 31 + batch)
    x = rand_polys(rng, batch, n, q)
    x[0, :, 0] = (q - 1)
    x[0, 0, :] = q[0] - 1
    words = ring.words(batch)
    d = backend.alloc(words * 8)
    backend.h2d(d, x)
    ring.ntt(d, batch, False, algo=2)
    assert np.array_equal(backend.d2h(d, words).reshape(x.shape), octx.ntt(x, inverse=False)), "forward"
    ring.ntt(d, batch, True, algo=2)
    assert np.array_equal(backend.d2h(d, words).reshape(x.shape), x), "round trip"
    y = rand_polys(rng, batch, n, q)
    backend.h2d(d, y)
    ring.ntt(d, batch, True, algo=2)
    assert np.array_equal(backend.d2h(d, words).reshape(x.shape), octx.ntt(y, inverse=True)), "inverse"
    backend.free(d)
    ring.close()


@pytest.mark.parametrize("logn,towers,batch,nbits", [(9, 3, 5, 60), (10, 2, 9, 45), (11, 4, 3, 30), (12, 3, 5, 60), (13, 2, 9, 59),
                                                     (14, 3, 8, 60), (14, 8, 40, 60), (13, 16, 37, 50)])
def test_ntt_single_pass_vs_oracle(pkg, backend, oracle, monkeypatch, logn, towers, batch, nbits):
    """Both passes in one kernel, the whole tower in shared memory (ntt_fused_kernel, N <= 2^14; FHE_B200_NTT_FUSED=2 forces
    it at every shape): forward and inverse against the oracle, worst-case operands, more CTAs than the device holds at once."""
    monkeypatch.setenv("FHE_B200_NTT_FUSED", "2")
    n = 1 << logn
    q, psi = params(oracle, nbits, n, towers)
    octx = oracle.ctx(n, q, psi)
    ring = pkg.Ring(backend, n, q, psi)
    rng = np.random.default_rng(900 + logn * 31 + batch)
    x = rand_polys(rng, batch, n, q)
    x[0, :, 0] = (q - 1)
    x[0, 0, :] = q[0] - 1
    x[-1, :, :] = (q - 1)[:, None]
    words = ring.words(batch)
    d = backend.alloc(words * 8)
    backend.h2d(d, x)
    before = pkg._native.lib.fhe_launch_count()
    ring.ntt(d, batch, False)
    assert pkg._native.lib.fhe_launch_count() - before == 1          # one launch, not two
    assert np.array_equal(backend.d2h(d, words).reshape(x.shape), octx.ntt(x, inverse=False)), "forward"
    ring.ntt(d, batch, True)
    assert np.array_equal(backend.d2h(d, words).reshape(x.shape), x), "round trip"
    y = rand_polys(rng, batch, n, q)
    y[0, :, :] = (q - 1)[:, None]
    backend.h2d(d, y)
    ring.ntt(d, batch, True)
    assert np.array_equal(backend.d2h(d, words).reshape(x.shape), octx.ntt(y, inverse=True)), "inverse"
    backend.free(d)
    ring.close()


def test_ntt_ws_full_size_round_trip_property(pkg, backend, oracle):
    """BASELINE configs[2] ring at a batch the default dispatch sends to the persistent kernels: forward then inverse
    is the identity, and the forward result equals the one-tile-per-CTA kernels' (which the oracle pins at small batch)."""
    n, towers, batch = 1 << 16, 30, 64
    q, psi = params(oracle, 60, n, towers)
    ring = pkg.Ring(backend, n, q, psi)
    words = ring.words(batch)
    a, b = backend.alloc(words * 8), backend.alloc(words * 8)
    ring.fill_uniform(a, batch, 424242)
    backend.d2d(b, a, words * 8)
    before = backend.checksum(a, words)
    ring.ntt(a, batch, False)                  # auto -> persistent kernels
    ring.ntt(b, batch, False, algo=2)
    assert backend.checksum(a, words) == backend.checksum(b, words)
    assert np.array_equal(backend.d2h(a, 1 << 20), backend.d2h(b, 1 << 20))
    ring.ntt(a, batch, True)
    ring.ntt(b, batch, True, algo=2)
    assert backend.checksum(a, words) == before and backend.checksum(b, words) == before
    backend.free(a); backend.free(b)
    ring.close()


@pytest.mark.parametrize("logn,towers,batch", [(4, 3, 2), (11, 8, 3), (14, 30, 1), (16, 4, 2)])
def test_dcrt_op_scalars_vs_numpy(pkg, backend, oracle, logn, towers, batch):
    """fhe_dcrt_op_scalars: DCRTPolyImpl::Plus / Minus(std::vector<Integer>) (dcrtpoly-impl.h:556-563, 577-584) - one
    constant per tower, reduced first as NativeVector::ModAdd / ModSub do, on every coefficient; in place and out of place."""
    n = 1 << logn
    q, psi = params(oracle, 60, n, towers)
    ring = pkg.Ring(backend, n, q, psi)
    rng = np.random.default_rng(900 + logn)
    x = rand_polys(rng, batch, n, q)
    x[0, :, 0] = q - 1
    x[0, :, 1] = 0
    s = rng.integers(0, 1 << 63, size=towers, dtype=np.uint64) * np.uint64(2) + np.uint64(1)      # larger than the moduli
    s[0] = 0
    words = ring.words(batch)
    da, dd = backend.alloc(words * 8), backend.alloc(words * 8)
    backend.h2d(da, x)
    L = pkg._native.lib
    qi = [int(v) for v in q]
    for op, f in ((pkg.OP_ADD, lambda v, c, m: (v + c) % m), (pkg.OP_SUB, lambda v, c, m: (v - c) % m)):
        want = np.empty_like(x)
        for t in range(towers):
            c = int(s[t]) % qi[t]
            want[:, t, :] = np.array([[f(int(v), c, qi[t]) for v in row] for row in x[:, t, :min(n, 64)]], np.uint64) if n <= 64 else \
                ((x[:, t, :].astype(object) + (c if op == pkg.OP_ADD else -c)) % qi[t]).astype(np.uint64)
        pkg._native.check(L.fhe_dcrt_op_scalars(ring.h, op, dd, da, s.ctypes.data_as(ctypes.c_void_p), batch, 0))
        assert np.array_equal(backend.d2h(dd, words).reshape(x.shape), want)
        backend.d2d(dd, da, words * 8)
        pkg._native.check(L.fhe_dcrt_op_scalars(ring.h, op, dd, dd, s.ctypes.data_as(ctypes.c_void_p), batch, 0))
        assert np.array_equal(backend.d2h(dd, words).reshape(x.shape), want), "in place"
    assert L.fhe_dcrt_op_scalars(ring.h, pkg.OP_MUL, dd, da, s.ctypes.data_as(ctypes.c_void_p), batch, 0) != 0
    backend.free(da); backend.free(dd)
    ring.close()
