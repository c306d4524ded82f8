"""Partitioning of a batch of DCRTPolys over the GPUs of one node.

Every (batch item, tower) unit of modadd / modmul / NTT is independent, so ranks take disjoint
units and no collective runs inside the path.  Contiguous blocks of the batch go to each rank;
towers are split only when there are fewer batch items than ranks.  `gather_results` is the one
optional exchange (north_star: "NCCL over NVLink only to gather results"); `checksum` gives the
cheap alternative when the gathered result would not fit one GPU (cfg 4).

The reference's counterpart is the scatter in PimManager::copy_to_pim (`size / nr_dpus` words per
DPU, no remainder handling — src/core/pim/host/PimManager.cpp:15-24) and the commented-out
per-tower split of PimManager-Impl.cpp:46-91; here remainders are spread over the first ranks.
"""
from dataclasses import dataclass

import numpy as np


@dataclass(frozen=True)
class Shard:
    rank: int
    b0: int      # batch range [b0, b1)
    b1: int
    t0: int      # tower range [t0, t1)
    t1: int

    @property
    def batch(self):
        return self.b1 - self.b0

    @property
    def towers(self):
        return self.t1 - self.t0

    @property
    def units(self):
        return self.batch * self.towers


def _split(n, parts):
    """parts contiguous ranges covering [0, n), sizes differing by at most one."""
    base, rem = divmod(n, parts)
    out, lo = [], 0
    for i in range(parts):
        hi = lo + base + (1 if i < rem else 0)
        out.append((lo, hi))
        lo = hi
    return out


def plan(batch, towers, world):
    """One Shard per rank.  batch >= world: split the batch, all towers everywhere.  Otherwise each
    batch item is given to a group of ranks that split its towers; ranks left over get empty shards."""
    if batch < 0 or towers < 1 or world < 1:
        raise ValueError("plan: need batch >= 0, towers >= 1, world >= 1")
    if batch >= world or batch == 0:
        return [Shard(r, lo, hi, 0, towers) for r, (lo, hi) in enumerate(_split(batch, world))]
    shards, r = [], 0
    for b, (g0, g1) in enumerate(_split(world, batch)):          # ranks [g0, g1) share batch item b
        group = g1 - g0
        for (t0, t1) in _split(towers, group):                    # empty tower ranges if group > towers
            shards.append(Shard(r, b, b + 1, t0, t1) if t1 > t0 else Shard(r, b, b, 0, 0))
            r += 1
    return shards


def local_slice(data, shard):
    """View of a host array laid out [batch][tower][N] restricted to one shard."""
    return data[shard.b0:shard.b1, shard.t0:shard.t1, :]


def checksum(words):
    """Order-independent 128-bit digest of a uint64 array: (sum mod 2^64, xor).  Sums of shard
    checksums equal the checksum of the whole, so ranks can compare results without gathering them."""
    w = np.ascontiguousarray(words, dtype=np.uint64).reshape(-1)
    s = int(np.add.reduce(w, dtype=np.uint64)) if w.size else 0
    x = int(np.bitwise_xor.reduce(w)) if w.size else 0
    return s, x


def combine_checksums(parts):
    s, x = 0, 0
    for ps, px in parts:
        s = (s + ps) & 0xFFFFFFFFFFFFFFFF
        x ^= px
    return s, x


def gather_results(local, shards, n, dist, device=None):
    """all-gather the per-rank results (uint64 arrays shaped like their shard) into the full
    [batch][tower][N] array on every rank.  `dist` is torch.distributed (nccl on GPUs, gloo on CPU);
    data travels as int64 bit patterns."""
    import torch
    world = dist.get_world_size()
    batch = max(s.b1 for s in shards)
    towers = max((s.t1 for s in shards), default=0)
    maxw = max(s.units for s in shards) * n
    buf = torch.zeros(maxw, dtype=torch.int64, device=device)
    flat = np.ascontiguousarray(local, dtype=np.uint64).reshape(-1)
    if flat.size:
        buf[:flat.size] = torch.from_numpy(flat.view(np.int64)).to(buf.device)
    out = torch.empty(world * maxw, dtype=torch.int64, device=device)
    dist.all_gather_into_tensor(out, buf)
    host = out.cpu().numpy().view(np.uint64).reshape(world, maxw)
    full = np.zeros((batch, towers, n), np.uint64)
    for s in shards:
        if s.units:
            full[s.b0:s.b1, s.t0:s.t1, :] = host[s.rank, :s.units * n].reshape(s.batch, s.towers, n)
    return full


class _DeviceWords:
    """A device buffer of the C ABI (plain address) seen by torch through the CUDA array interface, as int64."""

    def __init__(self, dptr, nwords):
        self.__cuda_array_interface__ = {"shape": (int(nwords),), "typestr": "<i8", "data": (int(dptr), False), "version": 2}


def gather_device(dptr, nwords, dist, device, out=None):
    """all-gather every rank's `nwords` result words, HBM to HBM over NCCL (NVLink / NVSwitch), without a host copy:
    returns an int64 torch tensor [world][nwords] on `device` (`out`, when given, is reused as that tensor).  The caller synchronises its own stream first —
    torch enqueues the collective behind ITS current stream, not the C ABI's."""
    import torch
    src = torch.as_tensor(_DeviceWords(dptr, nwords), device=device)
    if out is None:
        out = torch.empty(dist.get_world_size() * int(nwords), dtype=torch.int64, device=device)
    dist.all_gather_into_tensor(out.view(-1), src)
    return out.view(dist.get_world_size(), int(nwords))
