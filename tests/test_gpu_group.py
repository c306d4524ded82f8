"""The multi-device part of the C ABI (fhe_group_*, fhe_comm_*) and of the kept surface (PimManager with `gpus=k`,
PimDCRTPoly over several devices), against the CPU oracle — bit exact.  Runs with however many B200s the box has:
with one, the group of one device takes the same code paths except NCCL; with two or more (`gpurun --gpus 2`) the
batch is really split and the gather really crosses NVLink."""
import os
import subprocess

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ADD, SUB, MUL = 0, 1, 2
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def rand_polys(rng, batch, n, q):
    return np.stack([rng.integers(0, int(qt), size=(batch, n), dtype=np.uint64) for qt in q], axis=1)


@pytest.fixture(scope="module")
def ndev(pkg):
    g = pkg.Group(0)
    n = g.size
    g.close()
    return n


@pytest.mark.parametrize("want", [1, 2, 4, 8])
def test_group_hot_path_vs_oracle(pkg, oracle, ndev, want):
    if want > ndev:
        pytest.skip(f"{ndev} device(s) visible")
    n, towers = 1 << 14, 3
    q, psi = oracle.gen_params(60, n, towers)
    octx = oracle.ctx(n, q, psi)
    rng = np.random.default_rng(want)
    with pkg.Group(want, opts="streams=2") as g:
        assert g.size == want
        ring = pkg.GroupRing(g, n, q, psi)
        for batch in (want * 3, want * 2 + 1, 1):
            # the partition is sharding.plan's: contiguous blocks, the first ranks take the remainder
            covered = 0
            for r in range(want):
                first, count = g.partition(batch, r)
                assert first == covered and count in (batch // want, batch // want + 1)
                covered += count
            assert covered == batch
            a, b = rand_polys(rng, batch, n, q), rand_polys(rng, batch, n, q)
            da, db, dc = ring.alloc(batch), ring.alloc(batch), ring.alloc(batch)
            ring.scatter(da, a, batch); ring.scatter(db, b, batch)
            ring.op(ADD, dc, da, db, batch)
            assert np.array_equal(ring.gather(dc, batch), octx.elementwise(ADD, a, b))
            ring.op(MUL, dc, da, db, batch)
            prod = octx.elementwise(MUL, a, b)
            assert np.array_equal(ring.gather(dc, batch), prod)
            ring.ntt(dc, batch, inverse=True)
            assert np.array_equal(ring.gather(dc, batch), octx.ntt(prod, inverse=True))
            ring.ntt(dc, batch, inverse=False)
            assert np.array_equal(ring.gather(dc, batch), prod)
            # the optional gather of results: every device receives the whole batch (NCCL when want > 1)
            full = [g.backend(r).alloc(ring.words(batch) * 8) for r in range(want)]
            ring.all_gather(full, dc, batch)
            g.sync()
            for r in range(want):
                assert np.array_equal(g.backend(r).d2h(full[r], ring.words(batch)).reshape(prod.shape), prod), f"device {r}"
                g.backend(r).free(full[r])
            # host operands over all devices: a += b; a *= b; SwitchFormat x2
            x = a.copy()
            ring.chain_host([ADD, MUL, pkg.OP_INTT, pkg.OP_NTT], x, b, batch)
            assert np.array_equal(x, octx.elementwise(MUL, octx.elementwise(ADD, a, b), b))
            for bufs in (da, db, dc):
                ring.free(bufs)
        ring.close()


def test_comm_of_one_rank(pkg, backend):
    """fhe_comm_* (the one-process-per-GPU form bench.py uses under torchrun) with a single rank: all-gather = copy."""
    try:
        uid = pkg.Comm.unique_id()
    except pkg.FheError as e:
        pytest.skip(f"NCCL not loadable here: {e}")
    comm = pkg.Comm(backend, 1, 0, uid)
    x = np.arange(100000, dtype=np.uint64) * np.uint64(0x9E3779B97F4A7C15)
    src, dst = backend.alloc(x.nbytes), backend.alloc(x.nbytes)
    backend.h2d(src, x)
    comm.all_gather(dst, src, x.size)
    assert np.array_equal(backend.d2h(dst, x.size), x)
    backend.free(src); backend.free(dst)
    comm.close()


def test_cpp_surface_over_several_devices(ndev):
    """tests/cpp/pim_multi_gpu_test.cpp: PimManager("gpus=k"), PimData and PimDCRTPoly from C++ on min(2, visible) GPUs."""
    exe = os.path.join(ROOT, "tests", "cpp", "bin", "pim_multi_gpu_test")
    if not os.path.exists(exe):
        subprocess.check_call(["make", "-C", ROOT, "host"])
    for k in sorted({1, min(2, ndev), ndev}):
        r = subprocess.run([exe, str(k)], capture_output=True, text=True, timeout=600)
        assert r.returncode == 0 and f"all ok on {k} device(s)" in r.stdout, r.stdout[-3000:] + r.stderr[-2000:]
