"""Multi-GPU host logic on CPU: the shard plan, and a world-size-2 `gloo` run in which each rank
computes its shard (with the oracle standing in for the device) and the gathered / checksummed
result equals the unsharded one."""
import importlib
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def sharding():
    return importlib.import_module("upmem--openfhe_b200.sharding")


@pytest.mark.parametrize("batch,towers,world", [(256, 8, 1), (256, 8, 8), (10, 3, 4), (3, 8, 8), (1, 30, 8), (2, 3, 8),
                                                (0, 4, 2), (7, 1, 8), (8, 8, 8), (1024, 30, 2)])
def test_plan_covers_every_unit_exactly_once(batch, towers, world):
    sh = sharding()
    shards = sh.plan(batch, towers, world)
    assert len(shards) == world and [s.rank for s in shards] == list(range(world))
    seen = np.zeros((batch, towers), np.int32)
    for s in shards:
        seen[s.b0:s.b1, s.t0:s.t1] += 1
    assert (seen == 1).all()
    units = [s.units for s in shards]
    if batch >= world:
        assert max(units) - min(units) <= towers          # batch split: at most one item of imbalance
        assert all(s.towers == towers for s in shards if s.units)


def test_checksums_add_up():
    sh = sharding()
    rng = np.random.default_rng(0)
    x = rng.integers(0, 2**63, size=(6, 3, 64), dtype=np.uint64)
    parts = [sh.checksum(sh.local_slice(x, s)) for s in sh.plan(6, 3, 4)]
    assert sh.combine_checksums(parts) == sh.checksum(x)


def _worker(rank, world, port, batch, towers, n, q_out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    import torch.distributed as dist
    sys.path.insert(0, ROOT)
    from tests import oracle_py
    sh = importlib.import_module("upmem--openfhe_b200.sharding")
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        O = oracle_py.load()
        q, psi = O.gen_params(60, n, towers)
        a, b = O.fill_ring(batch, n, q, 1), O.fill_ring(batch, n, q, 2)
        shards = sh.plan(batch, towers, world)
        me = shards[rank]
        if me.units:
            ctx = O.ctx(n, q[me.t0:me.t1], psi[me.t0:me.t1])
            la = np.ascontiguousarray(sh.local_slice(a, me))
            lb = np.ascontiguousarray(sh.local_slice(b, me))
            mul = ctx.elementwise(2, la, lb, threads=1)
            fwd = ctx.ntt(la, inverse=False, threads=1)
        else:
            mul = fwd = np.zeros((0, 0, n), np.uint64)
        full_mul = sh.gather_results(mul, shards, n, dist)
        full_fwd = sh.gather_results(fwd, shards, n, dist)
        sums = [None] * world
        dist.all_gather_object(sums, sh.checksum(mul))
        if rank == 0:
            ref = O.ctx(n, q, psi)
            exp_mul, exp_fwd = ref.elementwise(2, a, b, threads=1), ref.ntt(a, inverse=False, threads=1)
            ok = np.array_equal(full_mul, exp_mul) and np.array_equal(full_fwd, exp_fwd)
            ok = ok and sh.combine_checksums(sums) == sh.checksum(exp_mul)
            q_out.put(bool(ok))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("batch,towers", [(5, 3), (1, 4)])
def test_world_size_2_gloo_matches_unsharded(batch, towers):
    import torch.multiprocessing as mp
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q_out = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, batch, towers, 256, q_out)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(180)
        assert p.exitcode == 0
    assert q_out.get(timeout=5) is True
