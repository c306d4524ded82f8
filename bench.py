#!/usr/bin/env python
"""bench.py — RNS coefficient-ops/s of the B200 backend on BASELINE.json's headline configuration.

Workload (configs[1]): DCRTPoly add + mult, N = 2^14, 8 towers of 60-bit primes, batch 256
polynomials per GPU.  One step = one modadd launch + one modmul launch over the whole batch
(2 * 256 * 8 * 16384 = 67.1 M coefficient-ops per GPU).  Operands are 268 MB each (805 MB touched
per launch), several times the 126 MB L2, so no L2 flush is needed between iterations.

  value    : device-resident throughput, CUDA events on the launching stream, max over ranks
  e2e      : same metric through the host-operand C-ABI call (pinned host buffers; H2D of both
             operands and D2H of both results inside the timed region)
  roofline : modmul kernel alone, 24 B algorithmic traffic per coefficient, vs the measured HBM peak
  cpu_baseline : the reference's own CPU path (oracle/_ref) on the host cores, rank 0, N=1 only

`--impl reference` times only the CPU reference on the same config/metric.
Multi-GPU: one process per GPU under torchrun; the global batch (256 x N polynomials) is sharded
by upmem--openfhe_b200/sharding.py (weak scaling: 256 polynomials per GPU), no data-path collective;
NCCL carries the barrier, the max-over-ranks reduction and a checksum exchange of the results.
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

N_RING, TOWERS, BATCH, NBITS = 1 << 14, 8, 256, 60
WORKLOAD = "dcrtpoly add+mult N=2^14 T=8 batch=256 per GPU, 60-bit moduli (BASELINE configs[1])"
METRIC = "rns_coeff_ops_per_sec"
UNIT = "coeff-ops/s"


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


def run_reference(args):
    world, rank, local, _ = int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("RANK", "0")), 0, None
    if rank != 0:
        return
    from oracle import cpu_baseline
    r = cpu_baseline.time_addmul(N_RING, TOWERS, BATCH, NBITS, steps=args.steps, warmup=args.warmup)
    total = sum(r["step_seconds"])
    value = r["coeff_ops_per_step"] * len(r["step_seconds"]) / total
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * total / len(r["step_seconds"]), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u64", "data": "synthetic", "impl": "reference",
        "config": {"workload": WORKLOAD, "N": N_RING, "towers": TOWERS, "batch_per_gpu": BATCH,
                   "note": "CPU reference runs ONE GPU's share (batch 256) on the host cores"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": r["cores"], "kind": r["kind"],
                         "sample": f"full batch {BATCH}, in-place += and *= of the reference's DCRTPoly, "
                                   f"omp-for over the batch, {args.steps} steps"},
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


def main():
    args = parse()
    claim_stdout()
    if args.impl == "reference":
        return run_reference(args)
    world, rank, local, dist = dist_setup(args.gpus)
    pkg = importlib.import_module("upmem--openfhe_b200")
    be = pkg.Backend(local, "streams=4,arena_mb=1024")
    q, psi = pkg.gen_params(NBITS, N_RING, TOWERS)
    ring = pkg.Ring(be, N_RING, q, psi)
    sharding = importlib.import_module("upmem--openfhe_b200.sharding")
    shard = sharding.plan(BATCH * world, TOWERS, world)[rank]      # this rank's block of the global batch
    assert shard.batch == BATCH and shard.towers == TOWERS
    words = ring.words(BATCH)
    coeffs = words                                   # coefficient-ops per launch
    lib, ADD, MUL = pkg._native.lib, pkg.OP_ADD, pkg.OP_MUL

    da, db, dc, dd = (be.alloc(words * 8) for _ in range(4))
    ring.fill_uniform(da, BATCH, 20260002 + shard.b0 * 7919)
    ring.fill_uniform(db, BATCH, 20261002 + shard.b0 * 7919)
    be.sync()

    sampler = ClockSampler(local)
    sampler.start()

    def step():
        ring.op(ADD, dc, da, db, BATCH, 0)
        ring.op(MUL, dd, da, db, BATCH, 0)

    for _ in range(max(args.warmup, 3)):
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
    value = world * 2 * coeffs * args.steps / (ms * 1e-3)

    # ---- roofline of the dominant kernel (modmul), timed alone on its launching stream ----
    reps = max(20, min(args.steps, 200))
    for _ in range(3):
        ring.op(MUL, dd, da, db, BATCH, 0)
    be.sync()
    be.timer_start(0)
    for _ in range(reps):
        ring.op(MUL, dd, da, db, BATCH, 0)
    mul_ms = be.timer_stop(0) / reps
    be.timer_start(0)
    for _ in range(reps):
        ring.op(ADD, dc, da, db, BATCH, 0)
    add_ms = be.timer_stop(0) / reps
    peaks, peak_src = measured_peaks()
    alg_bytes = 24 * coeffs
    achieved = alg_bytes / (mul_ms * 1e-3) / 1e9
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tpath):
        with open(tpath) as f:
            traffic = json.load(f).get("ew_mul_dram_bytes_per_launch")
    roofline = {"bound": "hbm", "kernel": "ew_ring_kernel<MUL> (fhe_dcrt_op)", "achieved": achieved,
                "peak": peaks["hbm_gbs"], "unit": "GB/s", "frac": achieved / peaks["hbm_gbs"],
                "traffic": traffic, "algorithmic_bytes_per_launch": alg_bytes, "launch_ms": mul_ms,
                "peak_source": f"{peak_src} copy bandwidth (burst: kernel timed alone)",
                "frac_of_nominal_8TBs": achieved / 8000.0,
                "modadd": {"launch_ms": add_ms, "achieved": alg_bytes / (add_ms * 1e-3) / 1e9,
                           "frac": alg_bytes / (add_ms * 1e-3) / 1e9 / peaks["hbm_gbs"]}}

    # ---- end to end: host operands through the C ABI ----
    ha, pa = be.host_alloc(words)
    hb, pb = be.host_alloc(words)
    hc, pc = be.host_alloc(words)
    hd, pd = be.host_alloc(words)
    be.d2h(da, words, out=ha)
    be.d2h(db, words, out=hb)
    e2e_steps = max(3, min(args.steps, 10))
    # (i) the reference arm's own step, on host operands: a += b; a *= b in place (one result comes back)
    for _ in range(2):
        ring.chain_host([ADD, MUL], ha, hb, BATCH)
    barrier(dist, local)
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        ring.chain_host([ADD, MUL], ha, hb, BATCH)
    be.sync()
    e2e_s = max_over_ranks(dist, local, time.perf_counter() - t0)
    # (ii) the out-of-place form (a + b and a * b both returned): twice the D2H bytes
    be.d2h(da, words, out=ha)
    ring.multi_op_host([ADD, MUL], [hc, hd], ha, hb, BATCH)
    barrier(dist, local)
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        ring.multi_op_host([ADD, MUL], [hc, hd], ha, hb, BATCH)
    be.sync()
    e2e2_s = max_over_ranks(dist, local, time.perf_counter() - t0)
    # PCIe ceiling for this access pattern: both directions busy at once, pinned memory, same sizes
    N_ = pkg._native
    be.sync()
    t0 = time.perf_counter()
    N_.check(N_.lib.fhe_h2d(be.h, da, pa, words * 8, 0))
    N_.check(N_.lib.fhe_d2h(be.h, pc, dc, words * 8, 1))
    be.sync()
    pcie_both = words * 8 / (time.perf_counter() - t0) / 1e9
    ring.fill_uniform(da, BATCH, 20260002 + shard.b0 * 7919)
    be.sync()
    e2e = {"value": world * 2 * coeffs * e2e_steps / e2e_s, "unit": UNIT,
           "pcie": {"h2d_gbs": 2 * words * 8 / (e2e_s / e2e_steps) / 1e9,
                    "d2h_gbs": words * 8 / (e2e_s / e2e_steps) / 1e9,
                    "ceiling_gbs_each_way_bidirectional": pcie_both,
                    "note": "the host-operand path is PCIe-bound: 16 B up and 8 B down per coefficient slot "
                            "for 2 coefficient-ops"},
           "h2d_bytes_per_step": 2 * words * 8, "d2h_bytes_per_step": words * 8, "steps": e2e_steps,
           "ms_per_step": 1e3 * e2e_s / e2e_steps,
           "api": "fhe_dcrt_chain_host(ADD, MUL): a += b; a *= b on pinned host operands (the reference arm's "
                  "step), chunked H2D/kernels/D2H over 4 streams",
           "out_of_place": {"value": world * 2 * coeffs * e2e_steps / e2e2_s, "ms_per_step": 1e3 * e2e2_s / e2e_steps,
                            "h2d_bytes_per_step": 2 * words * 8, "d2h_bytes_per_step": 2 * words * 8,
                            "api": "fhe_dcrt_multi_op_host(ADD, MUL): a + b and a * b both returned"}}

    # ---- optional result gather (north_star: "NCCL over NVLink only to gather results"): every rank receives every
    # rank's product of one step, HBM to HBM; device-timed on torch's stream, max over ranks ----
    gather = None
    if dist is not None:
        import torch
        step()
        be.sync()
        dev = torch.device("cuda", local)
        g = sharding.gather_device(dd, words, dist, dev)             # warm-up: NCCL channels, output allocation
        torch.cuda.synchronize()
        own = g[rank, :4096].cpu().numpy().view("uint64")
        assert (own == be.d2h(dd, 4096)).all(), "gathered block differs from this rank's result"
        for _ in range(2):                                            # NCCL sets its channels up lazily
            g = sharding.gather_device(dd, words, dist, dev, out=g)
        torch.cuda.synchronize()
        reps_g, rounds = 5, []
        for _ in range(3):                                            # best of three rounds, each max over ranks
            barrier(dist, local)
            ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            ev0.record()
            for _ in range(reps_g):
                g = sharding.gather_device(dd, words, dist, dev, out=g)       # the output buffer is reused
            ev1.record()
            torch.cuda.synchronize()
            rounds.append(max_over_ranks(dist, local, ev0.elapsed_time(ev1) / reps_g))
        del g
        g_ms = min(rounds)
        step_ms = ms / args.steps
        gather = {"what": "all_gather_into_tensor of each rank's modmul result of one step (NCCL over NVLink/NVSwitch)",
                  "bytes_per_rank": words * 8, "ms": g_ms, "rounds_ms": rounds, "recv_gbs_per_rank": (world - 1) * words * 8 / (g_ms * 1e-3) / 1e9,
                  "value_with_gather": world * 2 * coeffs / ((step_ms + g_ms) * 1e-3),
                  "note": "value (without the gather) is the headline; a step moves 1.6 GB through HBM in "
                          f"{step_ms:.3f} ms, its 268 MB product takes {g_ms:.2f} ms to reach the other {world - 1} GPU(s)"}

    sampler.stop_flag = True
    sampler.join()

    extra = {}
    if not args.no_extra and rank == 0:
        extra = ntt_extra(pkg, be)

    cpu = None
    if world == 1 and rank == 0:
        try:
            from oracle import cpu_baseline
            r = cpu_baseline.time_addmul(N_RING, TOWERS, BATCH, NBITS, steps=5, warmup=1)
            med = statistics.median(r["step_seconds"])
            cpu = {"value": r["coeff_ops_per_step"] / med, "unit": UNIT, "cores": r["cores"], "kind": r["kind"],
                   "sample": f"full batch {BATCH} (same workload), in-place += and *=, omp-for over the batch, median of 5"}
            if not args.no_extra:     # the reference's own SwitchFormat on the host cores, beside the NTT figure
                rn = cpu_baseline.time_ntt(1 << 16, 30, 16, NBITS, steps=3, warmup=1)
                mn = statistics.median(rn["step_seconds"])
                if "ntt_n65536_t30_b64" in extra:
                    extra["ntt_n65536_t30_b64"]["cpu_reference"] = {
                        "coeff_per_s": rn["coeff_ops_per_step"] / mn, "cores": rn["cores"], "kind": rn["kind"],
                        "sample": "batch 16 of the same ring, forward + inverse SwitchFormat, omp-for over the batch, median of 3"}
        except Exception as ex:      # the checker is optional for the product, never the other way round
            cpu = {"value": None, "unit": UNIT, "cores": 0, "kind": "unavailable", "sample": str(ex)[:200]}

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "u64", "data": "synthetic",
            "config": {"workload": WORKLOAD, "N": N_RING, "towers": TOWERS, "batch_per_gpu": BATCH,
                       "l2": "operands 268 MB each > 126 MB L2: no flush needed", "parallelism": f"batch-sharded x{world}"},
            "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": int(launches), "gather": gather,
            "clocks": sampler.summary(), "extra": extra,
        }
        emit(line)
    ring.close()
    be.close()
    if dist is not None:
        dist.destroy_process_group()


def ntt_extra(pkg, be):
    """Secondary figures (rank 0): batched forward / inverse NTT at BASELINE configs[2]'s shape
    (N=2^16, 30 towers; batch cut to 64 so the run stays short, 1 GB in place > L2) against the
    integer roofline.  Integer roofline = rate of lazy Shoup products (mulhi64 + 2 mullo64, the multiply
    work of one butterfly) measured live by fhe_imad_probe; achieved = butterflies/s of the transform."""
    out = {}
    try:
        blocks, threads, iters = 148 * 8, 256, 4096
        p_ms = min(be.imad_probe(blocks, threads, iters) for _ in range(3))
        shoup_peak = blocks * threads * iters * 8 / (p_ms * 1e-3)
        out["int_probe"] = {"shoup_products_per_s": shoup_peak, "ms": p_ms,
                            "note": "8 independent chains of mulhi64 + 2 mullo64 per thread, 148x8 CTAs x 256 threads"}
        peaks, _ = measured_peaks()
        for logn, towers, batch in ((16, 30, 64), (14, 8, 256)):
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
        for p_ in ins + outs:
            be.free(p_)
        ring.close()
    except Exception as ex:
        out["error"] = str(ex)[:300]
    return out


if __name__ == "__main__":
    main()
