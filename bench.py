#!/usr/bin/env python
"""bench.py — RNS coefficient-ops/s of the B200 backend on BASELINE.json's metric: add, mult AND the NTT.

Workload (configs[2]'s ring, the largest single-GPU configuration): N = 2^16, 30 towers of 60-bit primes, batch 1024
polynomials per GPU (2.01 G coefficients = 16.1 GB per operand, two operands resident).  One step is the reference
arm's own operator sequence on every polynomial of the batch, in place as the reference applies it:

    a += b;  a *= b;  a.SwitchFormat();  a.SwitchFormat()        (EVALUATION -> COEFFICIENT -> EVALUATION)

= one modadd launch, one modmul launch, one inverse and one forward NTT (two launches each) over the whole batch:
4 coefficient-ops per coefficient slot (one op = one output coefficient of modadd, of modmul, or one
coefficient through one NTT / INTT), 8.05e9 per step per GPU.  The buffers are far larger than the 126 MB L2, so no
flush is needed between iterations.

  value    : device-resident throughput, CUDA events on the launching stream, max over ranks
  e2e      : the same step through the host-operand C-ABI call fhe_dcrt_chain_host(ADD, MUL, INTT, NTT): pinned host
             buffers, H2D of a and b and D2H of a inside the timed region
  roofline : the dominant kernel family of the step (the forward NTT's two passes) against the integer-pipe roof
             (Shoup products/s, measured live by fhe_imad_probe), with the element-wise HBM roofline beside it
  cpu_baseline : the reference's own CPU path (oracle/_ref) on the host cores, rank 0, N=1 only, bounded sample

`--impl reference` times only the CPU reference on the same config/metric (a bounded sample of the batch per step).
Multi-GPU: one process per GPU under torchrun; the global batch (1024 x N polynomials) is sharded by
upmem--openfhe_b200/sharding.py (weak scaling: 1024 polynomials per GPU), no data-path collective; NCCL carries the
barrier, the max-over-ranks reduction and the optional gather of a result block.
"""
import argparse
import importlib
import json
import os
import statistics
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

N_RING, TOWERS, BATCH, NBITS = 1 << 16, 30, 1024, 60
OPS_PER_SLOT = 4                       # modadd, modmul, INTT, NTT
WORKLOAD = ("dcrtpoly a+=b; a*=b; SwitchFormat x2 (INTT, NTT), in place, N=2^16 T=30 batch=1024 per GPU, "
            "60-bit moduli (BASELINE configs[2] ring; the metric's add, mult and NTT in one step)")
METRIC = "rns_coeff_ops_per_sec"
UNIT = "coeff-ops/s"
CPU_SAMPLE_BATCH = 64                  # polynomials per CPU step (of the 1024): ~0.3 s per step on 16 cores
# identical in both arms (the driver compares the dicts key for key)
CONFIG = {"workload": WORKLOAD, "N": N_RING, "towers": TOWERS, "batch_per_gpu": BATCH, "moduli_bits": NBITS,
          "ops_per_step": ["modadd", "modmul", "intt", "ntt"], "coeff_ops_per_step_per_gpu": OPS_PER_SLOT * BATCH * TOWERS * N_RING,
          "l2": "operands 16.1 GB each >> 126 MB L2: no flush needed"}


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-extra", action="store_true", help="skip the NTT / probe extras")
    return ap.parse_args()


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return json.load(f), "measured"
    return {"hbm_gbs": 6650.0}, "fallback"


class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons through NVML while the timed regions run."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.stop_flag, self.samples, self.reasons, self.max_mhz = index, False, [], set(), None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def run(self):
        if not self.nv:
            return
        nv = self.nv
        names = {
            getattr(nv, "nvmlClocksThrottleReasonHwSlowdown", 0x8): "hw_slowdown",
            getattr(nv, "nvmlClocksThrottleReasonHwThermalSlowdown", 0x40): "hw_thermal_slowdown",
            getattr(nv, "nvmlClocksThrottleReasonSwThermalSlowdown", 0x20): "sw_thermal_slowdown",
            getattr(nv, "nvmlClocksThrottleReasonSwPowerCap", 0x4): "sw_power_cap",
        }
        while not self.stop_flag:
            try:
                util = nv.nvmlDeviceGetUtilizationRates(self.h).gpu
                mhz = nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)
                mask = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                self.samples.append((mhz, util))
                for bit, name in names.items():
                    if mask & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            time.sleep(0.005)

    def summary(self):
        busy = [m for m, u in self.samples if u > 0] or [m for m, _ in self.samples]
        return {"sm_mhz": statistics.median(busy) if busy else None, "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(self.samples)}


def dist_setup(ngpus):
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    dist = None
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    return world, rank, local, dist


def barrier(dist, local):
    if dist is not None:
        import torch
        dist.barrier(device_ids=[local])
        torch.cuda.synchronize()


def max_over_ranks(dist, local, x):
    if dist is None:
        return x
    import torch
    t = torch.tensor([x], dtype=torch.float64, device=f"cuda:{local}")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def cpu_sample_note(batch, steps, what):
    return (f"{batch} of the {BATCH} polynomials per step (same ring, same four operators: += , *= , SwitchFormat x2 of the "
            f"reference's DCRTPoly, in place), {what}, {steps} steps")


def claim_host_threads():
    """Before libgomp is loaded: torchrun exports OMP_NUM_THREADS=1, and the reference sizes its thread pool from
    omp_get_max_threads() in a static constructor (utils/parallel.h:52).  Returns (host threads, the old value)."""
    was = os.environ.get("OMP_NUM_THREADS")
    try:
        n = max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        n = max(1, os.cpu_count() or 1)
    os.environ["OMP_NUM_THREADS"] = str(n)
    return n, was


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads, omp_was = claim_host_threads()
    from oracle import cpu_baseline
    r = cpu_baseline.time_step(N_RING, TOWERS, CPU_SAMPLE_BATCH, NBITS, steps=args.steps, warmup=args.warmup, threads=threads)
    total = sum(r["step_seconds"])
    value = r["coeff_ops_per_step"] * len(r["step_seconds"]) / total
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * total / len(r["step_seconds"]), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u64", "data": "synthetic", "impl": "reference",
        "config": CONFIG,
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": r["cores"], "kind": r["kind"],
                         "host_threads": r["host_threads"], "omp_num_threads_env_was": omp_was,
                         "sample": cpu_sample_note(CPU_SAMPLE_BATCH, args.steps, "omp-for over the batch on every host thread")},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)


_JSON_FD = None


def claim_stdout():
    """Keep stdout for the ONE JSON line: everything else this process (or a library inside it, e.g. NCCL's version
    banner) writes to fd 1 goes to stderr instead."""
    global _JSON_FD
    sys.stdout.flush()
    _JSON_FD = os.dup(1)
    os.dup2(2, 1)


def emit(line):
    data = (json.dumps(line) + "\n").encode()
    if _JSON_FD is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        os.write(_JSON_FD, data)


def host_mem_available():
    """Bytes of host memory this process can still take (MemAvailable, capped by the cgroup limit if there is one)."""
    avail = None
    try:
        with open("/proc/meminfo") as f:
            for ln in f:
                if ln.startswith("MemAvailable:"):
                    avail = int(ln.split()[1]) * 1024
    except OSError:
        pass
    for lim, cur in (("/sys/fs/cgroup/memory.max", "/sys/fs/cgroup/memory.current"),
                     ("/sys/fs/cgroup/memory/memory.limit_in_bytes", "/sys/fs/cgroup/memory/memory.usage_in_bytes")):
        try:
            v = open(lim).read().strip()
            if v != "max":
                room = int(v) - int(open(cur).read().strip())
                avail = room if avail is None else min(avail, room)
        except (OSError, ValueError):
            pass
    return avail


def run_group(args):
    """`--gpus N` without torchrun: ONE process drives N devices through the library's group API (fhe_group_*), the
    form a C++ caller of PimManager("gpus=N") gets.  Same step, same metric; device time = max over the devices."""
    host_threads, omp_was = claim_host_threads()
    pkg = importlib.import_module("upmem--openfhe_b200")
    n = args.gpus
    g = pkg.Group(n, opts="streams=4,arena_mb=1024")
    q, psi = pkg.gen_params(NBITS, N_RING, TOWERS)
    ring = pkg.GroupRing(g, N_RING, q, psi)
    batch = BATCH * n
    bes = [g.backend(r) for r in range(n)]
    da, db = ring.alloc(batch), ring.alloc(batch)
    ring.fill_uniform(da, batch, 20260002)
    ring.fill_uniform(db, batch, 20261002)
    g.sync()
    ADD, MUL = pkg.OP_ADD, pkg.OP_MUL
    warmup = max(args.warmup, 3)
    sampler = ClockSampler(0)
    sampler.start()

    def step():
        ring.op(ADD, da, da, db, batch)
        ring.op(MUL, da, da, db, batch)
        ring.ntt(da, batch, True)
        ring.ntt(da, batch, False)

    for _ in range(warmup):
        step()
    g.sync()
    l0 = pkg._native.lib.fhe_launch_count()
    for be in bes:
        be.timer_start(0)
    for _ in range(args.steps):
        step()
    ms = max(be.timer_stop(0) for be in bes)
    launches = pkg._native.lib.fhe_launch_count() - l0
    words = ring.words(batch)
    value = OPS_PER_SLOT * words * args.steps / (ms * 1e-3)
    # end to end on host operands: every device streams its block of the pinned host batch through its own pipeline
    avail = host_mem_available()
    e2e_batch = batch
    while e2e_batch > 32 * n and avail is not None and 2 * e2e_batch * TOWERS * N_RING * 8 > 0.4 * avail:
        e2e_batch //= 2
    e2e_words = e2e_batch * TOWERS * N_RING
    ha, pa = bes[0].host_alloc(e2e_words)
    hb, pb = bes[0].host_alloc(e2e_words)
    seed_ring = pkg.Ring(bes[0], N_RING, q, psi)
    tmp = bes[0].alloc(64 * TOWERS * N_RING * 8)
    for off in range(0, e2e_batch, 64):                 # synthetic residues, 64 polynomials at a time
        cnt = min(64, e2e_batch - off)
        for hbuf, sd in ((ha, 1), (hb, 2)):
            seed_ring.fill_uniform(tmp, cnt, 20260002 + 1000 * sd + off)
            bes[0].d2h(tmp, cnt * TOWERS * N_RING, out=hbuf[off * TOWERS * N_RING:(off + cnt) * TOWERS * N_RING])
    bes[0].free(tmp)
    seed_ring.close()
    OPS = [ADD, MUL, pkg.OP_INTT, pkg.OP_NTT]
    ring.chain_host(OPS, ha, hb, e2e_batch)
    e2e_steps = max(2, min(args.steps, 3))
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        ring.chain_host(OPS, ha, hb, e2e_batch)
    e2e_s = time.perf_counter() - t0
    sampler.stop_flag = True
    sampler.join()
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": n, "steps": args.steps, "warmup": warmup,
        "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u64",
        "data": "synthetic", "config": CONFIG,
        "parallelism": f"ONE process, {n} devices through fhe_group_* (what PimManager(\"gpus={n}\") drives), batch-sharded, "
                       "no data-path collective",
        "roofline": None, "cpu_baseline": None,
        "e2e": {"value": OPS_PER_SLOT * e2e_words * e2e_steps / e2e_s, "unit": UNIT, "h2d_bytes_per_step": 2 * e2e_words * 8,
                "d2h_bytes_per_step": e2e_words * 8, "steps": e2e_steps, "ms_per_step": 1e3 * e2e_s / e2e_steps,
                "batch_total": e2e_batch, "api": "fhe_group_chain_host(ADD, MUL, INTT, NTT): one host thread per device"},
        "gpu_launches": int(launches), "clocks": sampler.summary(),
    }
    emit(line)
    ring.free(da); ring.free(db)
    ring.close()
    g.close()


def main():
    args = parse()
    claim_stdout()
    if args.impl == "reference":
        return run_reference(args)
    if args.gpus > 1 and "WORLD_SIZE" not in os.environ:
        return run_group(args)
    host_threads, omp_was = claim_host_threads()
    world, rank, local, dist = dist_setup(args.gpus)
    pkg = importlib.import_module("upmem--openfhe_b200")
    be = pkg.Backend(local, "streams=4,arena_mb=1024")
    q, psi = pkg.gen_params(NBITS, N_RING, TOWERS)
    ring = pkg.Ring(be, N_RING, q, psi)
    sharding = importlib.import_module("upmem--openfhe_b200.sharding")
    shard = sharding.plan(BATCH * world, TOWERS, world)[rank]      # this rank's block of the global batch
    assert shard.batch == BATCH and shard.towers == TOWERS
    words = ring.words(BATCH)
    coeffs = words                                   # coefficient slots per GPU
    lib, ADD, MUL = pkg._native.lib, pkg.OP_ADD, pkg.OP_MUL
    warmup = max(args.warmup, 3)

    da, db = be.alloc(words * 8), be.alloc(words * 8)
    ring.fill_uniform(da, BATCH, 20260002 + shard.b0 * 7919)
    ring.fill_uniform(db, BATCH, 20261002 + shard.b0 * 7919)
    be.sync()

    sampler = ClockSampler(local)
    sampler.start()

    # integer-pipe roof, measured before the long loops (the HBM-bound kernels run into the power cap, and a probe
    # taken right after them sees the lowered clock for a few ms) and again after them; the roof is the better one
    blocks, threads, iters = 148 * 8, 256, 4096
    be.imad_probe(blocks, threads, 64)
    p_cold = min(be.imad_probe(blocks, threads, iters) for _ in range(3))

    def step():
        ring.op(ADD, da, da, db, BATCH, 0)          # a += b
        ring.op(MUL, da, da, db, BATCH, 0)          # a *= b
        ring.ntt(da, BATCH, True, 0)                # a.SwitchFormat(): EVALUATION -> COEFFICIENT
        ring.ntt(da, BATCH, False, 0)               # a.SwitchFormat(): COEFFICIENT -> EVALUATION

    for _ in range(warmup):
        step()
    be.sync()
    barrier(dist, local)
    l0 = lib.fhe_launch_count()
    be.timer_start(0)
    for _ in range(args.steps):
        step()
    ms = be.timer_stop(0)
    launches = lib.fhe_launch_count() - l0
    barrier(dist, local)
    ms = max_over_ranks(dist, local, ms)
    value = world * OPS_PER_SLOT * coeffs * args.steps / (ms * 1e-3)

    # ---- rooflines: every kernel family of the step timed alone on its launching stream ----
    def timed(f, reps):
        for _ in range(2):
            f()
        be.sync()
        be.timer_start(0)
        for _ in range(reps):
            f()
        return be.timer_stop(0) / reps

    reps = 5
    fwd_ms = timed(lambda: ring.ntt(da, BATCH, False, 0), reps)
    inv_ms = timed(lambda: ring.ntt(da, BATCH, True, 0), reps)
    mul_ms = timed(lambda: ring.op(MUL, da, da, db, BATCH, 0), reps)
    add_ms = timed(lambda: ring.op(ADD, da, da, db, BATCH, 0), reps)
    be.sync()
    time.sleep(0.2)                                  # let the power governor settle after the HBM-bound loops
    p_hot = min(be.imad_probe(blocks, threads, iters) for _ in range(5))
    p_ms = min(p_cold, p_hot)
    shoup_peak = blocks * threads * iters * 8 / (p_ms * 1e-3)
    peaks, peak_src = measured_peaks()
    traffic = {}
    tpath = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tpath):
        with open(tpath) as f:
            traffic = json.load(f)
    logn = N_RING.bit_length() - 1
    bf = coeffs * logn / 2                            # (N/2) log2 N butterflies per tower
    moved = int(traffic.get("ntt_moved_bytes_per_coeff", 32))
    ew_bytes = 24 * coeffs
    ew_ach = ew_bytes / (mul_ms * 1e-3) / 1e9
    roofline = {
        "bound": "integer pipe", "kernel": "forward NTT (fhe_ntt_forward: all its launches)",
        "achieved": bf / (fwd_ms * 1e-3), "peak": shoup_peak, "unit": "butterflies/s vs Shoup products/s (mulhi64 + 2 mullo64)",
        "frac": bf / (fwd_ms * 1e-3) / shoup_peak,
        "traffic": traffic.get("ntt_fwd_dram_bytes_per_transform_b1024"),
        "algorithmic_bytes_per_launch": 16 * coeffs, "launch_ms": fwd_ms,
        "peak_source": "fhe_imad_probe, measured live in this run (148x8 CTAs x 256 threads x 8 chains), best of a probe "
                       "before the timed loops and one after them",
        "probe_ms": {"before": p_cold, "after": p_hot},
        "coeff_per_s": coeffs / (fwd_ms * 1e-3),
        "inverse": {"launch_ms": inv_ms, "achieved": bf / (inv_ms * 1e-3), "frac": bf / (inv_ms * 1e-3) / shoup_peak,
                    "coeff_per_s": coeffs / (inv_ms * 1e-3)},
        "hbm": {"algorithmic_bytes_per_coeff": 16, "moved_bytes_per_coeff": moved,
                "achieved_gbs_moved": moved * coeffs / (fwd_ms * 1e-3) / 1e9,
                "frac_of_measured_peak": moved * coeffs / (fwd_ms * 1e-3) / 1e9 / peaks["hbm_gbs"]},
        "elementwise": {"bound": "hbm", "kernel": "ew_ring_kernel<MUL> (fhe_dcrt_op)", "achieved": ew_ach,
                        "peak": peaks["hbm_gbs"], "unit": "GB/s", "frac": ew_ach / peaks["hbm_gbs"],
                        "traffic": traffic.get("ew_mul_dram_bytes_per_launch_b1024"),
                        "algorithmic_bytes_per_launch": ew_bytes, "launch_ms": mul_ms,
                        "peak_source": f"{peak_src} copy bandwidth (burst: kernel timed alone); above 1.0 = faster than the "
                                       "driver's plain copy, part of the result is still in L2 when the kernel ends",
                        "frac_of_nominal_8TBs": ew_ach / 8000.0,
                        "modadd": {"launch_ms": add_ms, "achieved": ew_bytes / (add_ms * 1e-3) / 1e9,
                                   "frac": ew_bytes / (add_ms * 1e-3) / 1e9 / peaks["hbm_gbs"]}},
    }

    # ---- end to end: the same step on HOST operands through the C ABI ----
    # pinned a and b; batch = the largest power-of-two share of the 1024 whose two pinned operands fit in 40 % of the
    # host memory this rank may take (all 1024 on a box with >= 80 GB free per rank)
    OPS = [ADD, MUL, pkg.OP_INTT, pkg.OP_NTT]
    item_bytes = TOWERS * N_RING * 8
    avail = host_mem_available()
    local_world = int(os.environ.get("LOCAL_WORLD_SIZE", world))
    e2e_batch = BATCH
    while e2e_batch > 32 and avail is not None and 2 * e2e_batch * item_bytes * local_world > 0.4 * avail:
        e2e_batch //= 2
    ha = hb = None
    while ha is None:
        try:
            ha, pa = be.host_alloc(e2e_batch * TOWERS * N_RING)
            hb, pb = be.host_alloc(e2e_batch * TOWERS * N_RING)
        except pkg.FheError:
            if ha is not None:
                be.host_free(pa)
                ha = None
            if e2e_batch <= 32:
                raise
            e2e_batch //= 2
    e2e_words = e2e_batch * TOWERS * N_RING
    ring.fill_uniform(da, BATCH, 20260002 + shard.b0 * 7919)
    be.d2h(da, e2e_words, out=ha)
    be.d2h(db, e2e_words, out=hb)
    e2e_steps = max(2, min(args.steps, 3))
    ring.chain_host(OPS, ha, hb, e2e_batch)                     # warm-up (also first touch of the pipeline's buffers)
    barrier(dist, local)
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        ring.chain_host(OPS, ha, hb, e2e_batch)
    be.sync()
    e2e_s = max_over_ranks(dist, local, time.perf_counter() - t0)
    e2e = {"value": world * OPS_PER_SLOT * e2e_words * e2e_steps / e2e_s, "unit": UNIT,
           "h2d_bytes_per_step": 2 * e2e_words * 8, "d2h_bytes_per_step": e2e_words * 8, "steps": e2e_steps,
           "ms_per_step": 1e3 * e2e_s / e2e_steps, "batch_per_gpu": e2e_batch,
           "pcie": {"h2d_gbs": 2 * e2e_words * 8 / (e2e_s / e2e_steps) / 1e9,
                    "d2h_gbs": e2e_words * 8 / (e2e_s / e2e_steps) / 1e9,
                    "note": "PCIe-bound: 16 B up and 8 B down per coefficient slot for 4 coefficient-ops"},
           "api": "fhe_dcrt_chain_host(ADD, MUL, INTT, NTT): a += b; a *= b; a.SwitchFormat() x2 on pinned host operands "
                  "(the reference arm's step), chunked H2D / kernels / D2H over 4 streams",
           "host_mem_available_gb": None if avail is None else avail / 1e9}
    be.host_free(pa)
    be.host_free(pb)

    # ---- optional result gather (north_star: "NCCL over NVLink only to gather results"): every rank receives a
    # 64-polynomial block (1 GB) of every rank's result, HBM to HBM; device-timed on torch's stream, max over ranks ----
    gather = None
    if dist is not None:
        g_batch = 64
        g_words = g_batch * TOWERS * N_RING
        be.sync()
        ids = [pkg.Comm.unique_id() if rank == 0 else None]
        dist.broadcast_object_list(ids, src=0)                       # torch carries the 128-byte id, nothing else
        comm = pkg.Comm(be, world, rank, ids[0])                     # ncclCommInitRank inside libfhe_b200.so
        gbuf = be.alloc(world * g_words * 8)
        comm.all_gather(gbuf, da, g_words, 0)                        # warm-up: NCCL channels
        be.sync()
        own = be.d2h(gbuf + rank * g_words * 8, 4096)
        assert (own == be.d2h(da, 4096)).all(), "gathered block differs from this rank's result"
        for _ in range(2):                                            # NCCL sets its channels up lazily
            comm.all_gather(gbuf, da, g_words, 0)
        be.sync()
        reps_g, rounds = 5, []
        for _ in range(3):                                            # best of three rounds, each max over ranks
            barrier(dist, local)
            be.timer_start(0)
            for _ in range(reps_g):
                comm.all_gather(gbuf, da, g_words, 0)
            rounds.append(max_over_ranks(dist, local, be.timer_stop(0) / reps_g))
        be.free(gbuf)
        comm.close()
        g_ms = min(rounds)
        gather = {"what": f"fhe_comm_all_gather (ncclAllGather on the library's own stream, NVLink/NVSwitch) of a {g_batch}-polynomial "
                          "block of each rank's result",
                  "bytes_per_rank": g_words * 8, "ms": g_ms, "rounds_ms": rounds,
                  "recv_gbs_per_rank": (world - 1) * g_words * 8 / (g_ms * 1e-3) / 1e9,
                  "note": "results stay sharded (a full gather is 16.1 GB per rank); value excludes the gather"}
    roofline["gather"] = gather

    sampler.stop_flag = True
    sampler.join()

    be.free(da)
    be.free(db)
    extra = {}
    if not args.no_extra and rank == 0:
        extra = ntt_extra(pkg, be)

    cpu = None
    if world == 1 and rank == 0:
        try:
            from oracle import cpu_baseline
            r = cpu_baseline.time_step(N_RING, TOWERS, CPU_SAMPLE_BATCH, NBITS, steps=5, warmup=1, threads=host_threads)
            med = statistics.median(r["step_seconds"])
            cpu = {"value": r["coeff_ops_per_step"] / med, "unit": UNIT, "cores": r["cores"], "kind": r["kind"],
                   "host_threads": r["host_threads"], "omp_num_threads_env_was": omp_was,
                   "sample": cpu_sample_note(CPU_SAMPLE_BATCH, 5, "omp-for over the batch on every host thread, median")}
            if r["kind"] == "reference":      # beside it: the reference's OWN parallelism (OpenMP over the towers of one poly)
                rf = cpu_baseline.time_step(N_RING, TOWERS, 8, NBITS, steps=3, warmup=1, threads=host_threads, faithful=True)
                mf = statistics.median(rf["step_seconds"])
                cpu["reference_faithful"] = {
                    "value": rf["coeff_ops_per_step"] / mf, "unit": UNIT, "cores": rf["cores"],
                    "sample": cpu_sample_note(8, 3, "serial over the batch, OpenFHE's own omp-for over <= T towers per call "
                                                    "(dcrtpoly-impl.h:412, utils/parallel.h:112-118), median")}
        except Exception as ex:      # the checker is optional for the product, never the other way round
            cpu = {"value": None, "unit": UNIT, "cores": 0, "kind": "unavailable", "sample": str(ex)[:200]}

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": warmup, "ms_per_step": ms / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "u64", "data": "synthetic",
            "config": CONFIG, "parallelism": f"batch-sharded x{world}, no data-path collective",
            "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": int(launches),
            "clocks": sampler.summary(), "extra": extra,
        }
        emit(line)
    ring.close()
    be.close()
    if dist is not None:
        dist.destroy_process_group()


def ntt_extra(pkg, be):
    """Secondary figures (rank 0): BASELINE configs[1] (add + mult, N=2^14, 4 and 8 towers, batch 256), the NTT at the
    other ring sizes against the integer roofline (rate of lazy Shoup products measured live by fhe_imad_probe;
    achieved = butterflies/s of the transform), the ciphertext pipelines of configs[3] / configs[4] and the
    neighbouring DCRTPoly methods."""
    out = {}
    try:
        peaks, _ = measured_peaks()
        for towers in (4, 8):                          # BASELINE configs[1]: "4 and 8 towers, batch 256"
            n, batch = 1 << 14, 256
            q, psi = pkg.gen_params(60, n, towers)
            ring = pkg.Ring(be, n, q, psi)
            words = ring.words(batch)
            a, b, c = (be.alloc(words * 8) for _ in range(3))
            ring.fill_uniform(a, batch, 20260002); ring.fill_uniform(b, batch, 20261002)
            res = {}
            for name, op in (("add", pkg.OP_ADD), ("mul", pkg.OP_MUL)):
                for _ in range(3):
                    ring.op(op, c, a, b, batch, 0)
                be.sync()
                reps = 50
                be.timer_start(0)
                for _ in range(reps):
                    ring.op(op, c, a, b, batch, 0)
                ms = be.timer_stop(0) / reps
                res[name] = {"launch_ms": ms, "coeff_ops_per_s": words / (ms * 1e-3), "hbm_gbs": 24 * words / (ms * 1e-3) / 1e9,
                             "frac_of_measured_hbm_peak": 24 * words / (ms * 1e-3) / 1e9 / peaks["hbm_gbs"]}
            res["add_plus_mul_coeff_ops_per_s"] = 2 * words / ((res["add"]["launch_ms"] + res["mul"]["launch_ms"]) * 1e-3)
            res["note"] = ("operands %d MB each" % (words * 8 >> 20)) + (": the three buffers fit the 126 MB L2 only partly" if towers == 4 else "")
            out[f"cfg1_addmul_n16384_t{towers}_b256"] = res
            for p_ in (a, b, c):
                be.free(p_)
            ring.close()
        blocks, threads, iters = 148 * 8, 256, 4096
        p_ms = min(be.imad_probe(blocks, threads, iters) for _ in range(3))
        shoup_peak = blocks * threads * iters * 8 / (p_ms * 1e-3)
        out["int_probe"] = {"shoup_products_per_s": shoup_peak, "ms": p_ms,
                            "note": "8 independent chains of mulhi64 + 2 mullo64 per thread, 148x8 CTAs x 256 threads"}
        for logn, towers, batch in ((15, 16, 128), (14, 8, 256)):
            n = 1 << logn
            q, psi = pkg.gen_params(60, n, towers)
            ring = pkg.Ring(be, n, q, psi)
            words = ring.words(batch)
            x = be.alloc(words * 8)
            ring.fill_uniform(x, batch, 20260003)
            for _ in range(3):
                ring.ntt(x, batch, False); ring.ntt(x, batch, True)
            be.sync()
            reps = 10
            be.timer_start(0)
            for _ in range(reps):
                ring.ntt(x, batch, False)
            f_ms = be.timer_stop(0) / reps
            be.timer_start(0)
            for _ in range(reps):
                ring.ntt(x, batch, True)
            i_ms = be.timer_stop(0) / reps
            bf = words * logn / 2                       # (N/2) log2 N butterflies per tower
            out[f"ntt_n{n}_t{towers}_b{batch}"] = {
                "fwd_ms": f_ms, "inv_ms": i_ms,
                "fwd_coeff_per_s": words / (f_ms * 1e-3), "inv_coeff_per_s": words / (i_ms * 1e-3),
                "roofline": {"bound": "integer pipe (fmaheavy)", "achieved": bf / (f_ms * 1e-3), "peak": shoup_peak,
                             "unit": "butterflies/s vs Shoup products/s", "frac": bf / (f_ms * 1e-3) / shoup_peak,
                             "frac_inverse": bf / (i_ms * 1e-3) / shoup_peak},
                "hbm": {"algorithmic_bytes_per_coeff": 16, "moved_bytes_per_coeff": 32,
                        "achieved_gbs_moved": 32 * words / (f_ms * 1e-3) / 1e9,
                        "frac_of_measured_peak": 32 * words / (f_ms * 1e-3) / 1e9 / peaks["hbm_gbs"]}}
            be.free(x)
            ring.close()
        # ciphertext multiplication pipelines (BASELINE configs[3] / configs[4] shapes, batch cut to fit a short run)
        for name, logn, towers, batch, transform in (("ct_tensor_n65536_t40_b32", 16, 40, 32, 0),
                                                     ("ct_mul_pipeline_cfg4_n65536_t40_b32", 16, 40, 32, 2),
                                                     ("ct_mul_pipeline_cfg5_n32768_t16_b128", 15, 16, 128, 1)):
            n = 1 << logn
            q, psi = pkg.gen_params(60, n, towers)
            ring = pkg.Ring(be, n, q, psi)
            words = ring.words(batch)
            bufs = [be.alloc(words * 8) for _ in range(7)]
            for i in range(4):
                ring.fill_uniform(bufs[i], batch, 20260010 + i)
            args = (bufs[4:7], bufs[0:2], bufs[2:4], batch, transform)
            for _ in range(2):
                ring.ct_multiply(*args)
            be.sync()
            reps = 5
            be.timer_start(0)
            for _ in range(reps):
                ring.ct_multiply(*args)
            ms = be.timer_stop(0) / reps
            ops_per_slot = 5 + (7 if transform else 0)          # 4 mul + 1 add (+ 4 + 3 transforms)
            out[name] = {"ms": ms, "ciphertext_products_per_s": batch / (ms * 1e-3),
                         "coeff_ops_per_s": ops_per_slot * words / (ms * 1e-3), "coeff_ops_per_slot": ops_per_slot,
                         "hbm_gbs_tensor_only": (56 * words / (ms * 1e-3) / 1e9) if not transform else None}
            for b_ in bufs:
                be.free(b_)
            ring.close()
        # automorphism (rotation) in both formats: 16 B of HBM traffic per coefficient
        n, towers, batch = 1 << 16, 30, 64
        q, psi = pkg.gen_params(60, n, towers)
        ring = pkg.Ring(be, n, q, psi)
        words = ring.words(batch)
        src, dst = be.alloc(words * 8), be.alloc(words * 8)
        ring.fill_uniform(src, batch, 20260020)
        lib = pkg._native.lib
        for fmt, name in ((0, "automorphism_eval_n65536_t30_b64"), (1, "automorphism_coef_n65536_t30_b64")):
            for _ in range(2):
                pkg._native.check(lib.fhe_automorphism(ring.h, dst, src, batch, 5, fmt, 0))
            be.sync()
            reps = 10
            be.timer_start(0)
            for _ in range(reps):
                pkg._native.check(lib.fhe_automorphism(ring.h, dst, src, batch, 5, fmt, 0))
            ms = be.timer_stop(0) / reps
            out[name] = {"ms": ms, "coeff_per_s": words / (ms * 1e-3), "hbm_gbs": 16 * words / (ms * 1e-3) / 1e9}
        # the two-argument form (PrecomputeAutoMap table resident in HBM, poly-impl.h:368-380)
        import ctypes as _ct
        import numpy as _np
        amap = _np.empty(n, _np.uint32)
        pkg._native.check(lib.fhe_precompute_automap(n, 5, amap.ctypes.data_as(_ct.c_void_p)))
        dmap = be.alloc(n * 4)
        pkg._native.check(lib.fhe_h2d(be.h, dmap, amap.ctypes.data_as(_ct.c_void_p), n * 4, 0))
        be.sync()
        for _ in range(2):
            pkg._native.check(lib.fhe_automorphism_map(ring.h, dst, src, batch, dmap, 0))
        be.sync()
        be.timer_start(0)
        for _ in range(reps):
            pkg._native.check(lib.fhe_automorphism_map(ring.h, dst, src, batch, dmap, 0))
        ms = be.timer_stop(0) / reps
        out["automorphism_map_n65536_t30_b64"] = {"ms": ms, "coeff_per_s": words / (ms * 1e-3), "hbm_gbs": 16 * words / (ms * 1e-3) / 1e9}
        be.free(dmap)
        be.free(src); be.free(dst)
        ring.close()
        # RNS basis conversion (hybrid key-switching shapes): ApproxSwitchCRTBasis Q -> P, ApproxModUp, ApproxModDown, rescale
        n, sq, sp, batch = 1 << 16, 30, 10, 16
        q, psi = pkg.gen_params(60, n, sq + sp)
        ring = pkg.Ring(be, n, q, psi)
        rq, rp = pkg.RingView(ring, 0, sq), pkg.RingView(ring, sq, sp)
        q2p, p2q, plan = pkg.Basis(rq, rp), pkg.Basis(rp, rq), pkg.Rescale(ring)
        wq, wp = batch * sq * n, batch * sp * n
        dq, dp, dqp, dr = be.alloc(wq * 8), be.alloc(wp * 8), be.alloc((wq + wp) * 8), be.alloc((wq + wp) * 8)
        rq.fill_uniform(dq, batch, 20260030)
        ring.fill_uniform(dqp, batch, 20260031)

        def timed(f, reps=5):
            for _ in range(2):
                f()
            be.sync()
            be.timer_start(0)
            for _ in range(reps):
                f()
            return be.timer_stop(0) / reps

        ms = timed(lambda: q2p.approx_switch(dp, dq, batch))
        out["approx_switch_crt_basis_n65536_q30_p10_b16"] = {
            "ms": ms, "mac128_per_s": batch * n * sq * sp / (ms * 1e-3), "polys_per_s": batch / (ms * 1e-3),
            "note": "one 64x64->128 multiply-accumulate per (source tower, target tower, coefficient)"}
        ms = timed(lambda: p2q.approx_switch(dq, dp, batch))
        out["approx_switch_crt_basis_n65536_q10_p30_b16"] = {
            "ms": ms, "mac128_per_s": batch * n * sq * sp / (ms * 1e-3), "polys_per_s": batch / (ms * 1e-3),
            "note": "the key-switching shape: few source towers, many target towers"}
        rq.fill_uniform(dq, batch, 20260030)
        ms = timed(lambda: q2p.approx_mod_up(dr, dq, batch, 0))
        out["approx_mod_up_n65536_q30_p10_b16"] = {"ms": ms, "polys_per_s": batch / (ms * 1e-3)}
        ms = timed(lambda: p2q.approx_mod_down(dq, dqp, batch, 0))
        out["approx_mod_down_n65536_q30_p10_b16"] = {"ms": ms, "polys_per_s": batch / (ms * 1e-3)}
        ms = timed(lambda: plan.drop_last_element_and_scale(dr, dqp, batch, 0))
        out["drop_last_element_and_scale_n65536_t40_b16"] = {"ms": ms, "polys_per_s": batch / (ms * 1e-3)}
        # hybrid key switching (CKKS relinearisation / rotation shape): 30 towers of Ql, 10 of P, 3 digits of 10 towers
        ks = pkg.KeySwitch(ring, sq, 10)
        kb, ka = be.alloc(ks.digits * (sq + sp) * n * 8), be.alloc(ks.digits * (sq + sp) * n * 8)
        ring.fill_uniform(kb, ks.digits, 20260033); ring.fill_uniform(ka, ks.digits, 20260034)
        o0, o1 = be.alloc(wq * 8), be.alloc(wq * 8)
        rq.fill_uniform(dq, batch, 20260030)
        ms = timed(lambda: ks(o0, o1, dq, kb, ka, batch, 0))
        out["keyswitch_hybrid_n65536_ql30_p10_alpha10_b16"] = {
            "ms": ms, "key_switches_per_s": batch / (ms * 1e-3),
            "note": "KeySwitchHYBRID::KeySwitchCore: 3 digits x (INTT, ApproxSwitchCRTBasis to the complement basis, NTT), fused "
                    "inner product with the key over 40 towers, 2 x ApproxModDown; device-resident"}
        ks.close()
        for b_ in (kb, ka, o0, o1):
            be.free(b_)
        # HPSPOVERQ tails (tower_scale_kernel): HBM-bound, 8 B in + 8 B out per coefficient per tower (+ the extra tower)
        rq1 = pkg.RingView(ring, 0, sq + sp - 1)
        kq = [int(v) // 3 for v in q[:sq + sp - 1]]
        ts = pkg.TowerScale(rq1, kq)
        ms = timed(lambda: ts.scale_and_round_p_over_q(dr, dqp, batch))
        out["scale_and_round_p_over_q_n65536_t39_b16"] = {
            "ms": ms, "polys_per_s": batch / (ms * 1e-3), "hbm_gbs": batch * n * (2 * (sq + sp) - 1) * 8 / (ms * 1e-3) / 1e9}
        ms = timed(lambda: ts.expand_crt_basis_ql_hat(dr, dq, sq, batch))
        out["expand_crt_basis_ql_hat_n65536_l30_t39_b16"] = {
            "ms": ms, "polys_per_s": batch / (ms * 1e-3), "hbm_gbs": batch * n * (sq + sq + sp - 1) * 8 / (ms * 1e-3) / 1e9}
        ts.close()
        rq1.close()
        for o in (q2p, p2q, plan, rq, rp):
            o.close()
        for b_ in (dq, dp, dqp, dr):
            be.free(b_)
        ring.close()
        # BFVrns EvalMult core with the HPS technique (BASELINE configs[4] as a real EvalMult): Q = 16 towers, R = 17,
        # N = 2^15; per ciphertext pair: 4 ExpandCRTBasis, tensor product over QR, 3 x (INTT, ScaleAndRound, SwitchCRTBasis)
        n, sq, sr, batch = 1 << 15, 16, 17, 32
        q, psi = pkg.gen_params(60, n, sq + sr)
        ring = pkg.Ring(be, n, q, psi)
        frac, modr = pkg.hps_scale_tables(q[:sq], q[sq:], 65537)
        mult = pkg.BfvHpsMult(ring, sq, frac, modr)
        wq = batch * sq * n
        ins = [be.alloc(wq * 8) for _ in range(4)]
        outs = [be.alloc(wq * 8) for _ in range(3)]
        for i, p_ in enumerate(ins):
            mult.rq.fill_uniform(p_, batch, 20260040 + i)
        ms = timed(lambda: mult(outs, ins[0:2], ins[2:4], batch), reps=3)
        out["bfv_hps_evalmult_core_n32768_q16_r17_b32"] = {
            "ms": ms, "ciphertext_products_per_s": batch / (ms * 1e-3),
            "note": "ExpandCRTBasis x4 -> tensor product over QR -> INTT -> ScaleAndRound -> SwitchCRTBasis, device-resident"}
        mult.close()
        ring.close()
        # the same product with the HPSPOVERQ technique (bfvrns-leveledshe.cpp:208-235, 357-366): R = 16 towers;
        # a: 2 ExpandCRTBasis, b: 2 x (INTT, FastExpandCRTBasisPloverQ, NTT), tensor product, 3 x (INTT, ScaleAndRound to Q)
        q, psi = pkg.gen_params(60, n, 2 * sq)
        ring = pkg.Ring(be, n, q, psi)
        neg, qinv = pkg.hps_p_over_q_tables(q[:sq], q[sq:])
        frac, modq = pkg.hps_scale_tables(q[sq:], q[:sq], 65537)
        mult = pkg.BfvHpsPOverQMult(ring, sq, neg, qinv, frac, modq)
        ms = timed(lambda: mult(outs, ins[0:2], ins[2:4], batch), reps=3)
        out["bfv_hpspoverq_evalmult_core_n32768_q16_r16_b32"] = {
            "ms": ms, "ciphertext_products_per_s": batch / (ms * 1e-3),
            "note": "ExpandCRTBasis x2 + (INTT, FastExpandCRTBasisPloverQ, NTT) x2 -> tensor product over QR -> INTT -> "
                    "ScaleAndRound to Q, device-resident"}
        mult.close()
        ring.close()
        # the same product with the BEHZ technique (bfvrns-leveledshe.cpp:275-297, 367-388): Bsk = 16 towers of B and msk;
        # 4 FastBaseConvqToBskMontgomery, tensor product over QBsk, 3 x (INTT, FastRNSFloorq, FastBaseConvSK)
        q, psi = pkg.gen_params(60, n, 2 * sq + 1)
        ring = pkg.Ring(be, n, q, psi)
        mult = pkg.BfvBehzMult(ring, sq, 65537)
        ms = timed(lambda: mult(outs, ins[0:2], ins[2:4], batch), reps=3)
        out["bfv_behz_evalmult_core_n32768_q16_bsk17_b32"] = {
            "ms": ms, "ciphertext_products_per_s": batch / (ms * 1e-3),
            "note": "FastBaseConvqToBskMontgomery x4 -> tensor product over QBsk -> INTT -> FastRNSFloorq -> FastBaseConvSK, "
                    "device-resident"}
        mult.close()
        for p_ in ins + outs:
            be.free(p_)
        ring.close()
    except Exception as ex:
        out["error"] = str(ex)[:300]
    return out


if __name__ == "__main__":
    main()
