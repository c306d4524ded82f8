"""BASELINE.json configs[3] and configs[4] at their full batch, sharded over the GPUs of one node.

  cfg3: CKKS ciphertext tensor product (INTT -> pointwise products -> NTT), N = 2^16, 40 towers, 4096 ciphertext pairs
  cfg4: BFVrns EvalMult, N = 2^15, 16 towers: (a) the primitive pipeline NTT -> tensor product -> INTT and (b) the real
        RNS core with the HPS technique (ExpandCRTBasis x4, tensor product over QR, INTT, ScaleAndRound, SwitchCRTBasis);
        batch 4096 pairs; rank 0 also times the reference's own DCRTPoly on the host cores for (b) on a small sample.

One process per GPU (torchrun, like bench.py); the batch is split into contiguous blocks (sharding.plan), no data-path
collective; NCCL carries the barrier and the max-over-ranks of the device time.  A rank streams its share through
chunk-sized buffers filled on the device (cfg3's operand set is 172 GB: it cannot be resident on one GPU), so every
pair is computed, on the values of its chunk slot.

  python tools/config_sweep.py [--gpus N] [--batch 4096] [--chunk 32] [--cpu-sample 8]
"""
import argparse
import ctypes
import importlib
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def cpu_bfv(n, sq, sr, pairs, t):
    """The reference's own DCRTPoly running the HPS EvalMult core on all host threads (hooked build, no handle attached)."""
    from tests import oracle_py
    so = os.path.join(os.path.dirname(oracle_py.REF_SO), "libopenfhe_hooked.so")
    if not os.path.exists(so):
        return None
    vp, sz, ci, u32, u64 = ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int, ctypes.c_uint32, ctypes.c_uint64
    H = ctypes.CDLL(so)
    H.ref_ctx_create.restype = vp; H.ref_ctx_create.argtypes = [u32, u32, vp, vp]
    H.ref_polys_load.restype = vp; H.ref_polys_load.argtypes = [vp, vp, sz, ci]
    H.ref_polys_free.argtypes = [vp]
    H.ref_max_threads.restype = ci
    H.ref_bfv_hps_mult_timed.argtypes = [vp, vp, vp, vp, vp, vp, u64, ctypes.POINTER(vp), ci, ctypes.POINTER(ctypes.c_double)]
    O = oracle_py.load()
    p = lambda a: a.ctypes.data_as(vp)
    q, psi = O.gen_params(60, n, sq + sr)
    mk = lambda a, b: H.ref_ctx_create(n, b - a, p(q[a:b].copy()), p(psi[a:b].copy()))
    cQ, cR, cQR = mk(0, sq), mk(sq, sq + sr), mk(0, sq + sr)
    hs = [H.ref_polys_load(cQ, p(O.fill_ring(pairs, n, q[:sq].copy(), 700 + k)), pairs, 0) for k in range(4)]
    threads = max(1, len(os.sched_getaffinity(0)))      # not omp_get_max_threads(): torchrun exports OMP_NUM_THREADS=1
    best = None
    for _ in range(2):
        outs, sec = (vp * 3)(), ctypes.c_double()
        assert H.ref_bfv_hps_mult_timed(hs[0], hs[1], hs[2], hs[3], cR, cQR, t, outs, threads, ctypes.byref(sec)) == 0
        best = sec.value if best is None else min(best, sec.value)
        for k in range(3):
            H.ref_polys_free(outs[k])
    return {"ciphertext_products_per_s": pairs / best, "seconds": best, "pairs": pairs, "threads": threads,
            "kind": "reference (its DCRTPoly: ExpandCRTBasis, *, SetFormat, ScaleAndRound, SwitchCRTBasis), omp-for over pairs"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--batch", type=int, default=4096)
    ap.add_argument("--chunk", type=int, default=32)
    ap.add_argument("--host-stream", type=int, default=0, help="1 GPU: also run cfg3 with host operands for this many chunks")
    ap.add_argument("--cpu-sample", type=int, default=8)
    args = ap.parse_args()
    world, rank, local = int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0"))
    out_fd = os.dup(1)
    os.dup2(2, 1)
    dist = None
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if dist is not None:
            import torch
            dist.barrier(device_ids=[local])
            torch.cuda.synchronize()

    def max_ranks(x):
        if dist is None:
            return x
        import torch
        t = torch.tensor([x], dtype=torch.float64, device=f"cuda:{local}")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    pkg = importlib.import_module("upmem--openfhe_b200")
    sharding = importlib.import_module("upmem--openfhe_b200.sharding")
    be = pkg.Backend(local, "streams=2,arena_mb=2048")
    result = {"n_gpus": world, "batch": args.batch, "chunk": args.chunk, "scaling": "strong"}

    def run(name, n, towers_total, make_call, pairs_words):
        share = sharding.plan(args.batch, 1, world)[rank].batch
        chunk = min(args.chunk, share)
        call, cleanup = make_call(chunk)
        call(chunk)
        call(chunk)
        be.sync()
        barrier()
        be.timer_start(0)
        done = 0
        while done < share:
            c = min(chunk, share - done)
            call(c)
            done += c
        ms = max_ranks(be.timer_stop(0))
        cleanup()
        result[name] = {"ms": ms, "ciphertext_products_per_s": args.batch / (ms * 1e-3), "pairs_per_gpu": share,
                        "N": n, "towers": towers_total}

    # ---- cfg3
    n, T = 1 << 16, 40
    q, psi = pkg.gen_params(60, n, T)
    ring = pkg.Ring(be, n, q, psi)

    def mk_cfg3(chunk):
        w = ring.words(chunk) * 8
        bufs = [be.alloc(w) for _ in range(7)]
        for i in range(4):
            ring.fill_uniform(bufs[i], chunk, 900 + i + 17 * rank)
        return (lambda c: ring.ct_multiply(bufs[4:7], bufs[0:2], bufs[2:4], c, 2)), (lambda: [be.free(b) for b in bufs])

    run("cfg3_ckks_tensor_intt_mul_ntt_n65536_t40", n, T, mk_cfg3, None)
    if args.host_stream > 0 and world == 1:
        # The same pipeline with HOST operands: every chunk's four input polynomials cross PCIe from pinned memory and its
        # three outputs go back, double-buffered over two streams (H2D / kernels / D2H of neighbouring chunks overlap).  The
        # 172 GB operand set of the full batch is represented by two pinned slots that are re-sent for every chunk.
        import time
        N_ = pkg._native
        chunk, w = args.chunk, ring.words(args.chunk)
        dev = [[be.alloc(w * 8) for _ in range(7)] for _ in range(2)]
        hin = [[be.host_alloc(w) for _ in range(4)] for _ in range(2)]
        hout = [[be.host_alloc(w) for _ in range(3)] for _ in range(2)]
        for sl in range(2):
            for i in range(4):
                ring.fill_uniform(dev[sl][i], chunk, 1900 + i + 7 * sl)
                be.d2h(dev[sl][i], w, out=hin[sl][i][0])

        def one(sl):
            for i in range(4):
                N_.check(N_.lib.fhe_h2d(be.h, dev[sl][i], ctypes.c_void_p(hin[sl][i][1]), w * 8, sl))
            ring.ct_multiply(dev[sl][4:7], dev[sl][0:2], dev[sl][2:4], chunk, 2, sl)
            for j in range(3):
                N_.check(N_.lib.fhe_d2h(be.h, ctypes.c_void_p(hout[sl][j][1]), dev[sl][4 + j], w * 8, sl))

        one(0); one(1); be.sync()
        t0 = time.perf_counter()
        for k in range(args.host_stream):
            be.sync(k % 2)
            one(k % 2)
        be.sync()
        sec = time.perf_counter() - t0
        pairs = args.host_stream * chunk
        result["cfg3_host_streamed"] = {
            "pairs": pairs, "seconds": sec, "ciphertext_products_per_s": pairs / sec,
            "h2d_gbs": pairs * 4 * T * n * 8 / sec / 1e9, "d2h_gbs": pairs * 3 * T * n * 8 / sec / 1e9,
            "full_batch_seconds_at_this_rate": args.batch / (pairs / sec),
            "note": "operands in pinned host memory, every chunk crosses PCIe both ways; two pinned slots stand for the 172 GB set"}
        for sl in range(2):
            for b in dev[sl]:
                be.free(b)
            for _, hp in hin[sl] + hout[sl]:
                be.host_free(hp)
    result["operands"] = ("device-generated: each rank fills chunk-sized buffers once on the device and reuses them for every chunk "
                          "of its share (device throughput; nothing crosses PCIe); cfg3_host_streamed, when present, is the "
                          "host-operand form")
    ring.close()

    # ---- cfg4
    n, sq, sr, t = 1 << 15, 16, 17, 65537
    q, psi = pkg.gen_params(60, n, sq + sr)
    ring = pkg.Ring(be, n, q, psi)
    rq = pkg.RingView(ring, 0, sq)

    def mk_cfg4a(chunk):
        w = rq.words(chunk) * 8
        bufs = [be.alloc(w) for _ in range(7)]
        for i in range(4):
            rq.fill_uniform(bufs[i], chunk, 950 + i + 17 * rank)
        return (lambda c: rq.ct_multiply(bufs[4:7], bufs[0:2], bufs[2:4], c, 1)), (lambda: [be.free(b) for b in bufs])

    run("cfg4_bfv_primitive_ntt_mul_intt_n32768_t16", n, sq, mk_cfg4a, None)
    frac, modr = pkg.hps_scale_tables(q[:sq], q[sq:], t)
    mult = pkg.BfvHpsMult(ring, sq, frac, modr)

    def mk_cfg4b(chunk):
        w = rq.words(chunk) * 8
        bufs = [be.alloc(w) for _ in range(7)]
        for i in range(4):
            rq.fill_uniform(bufs[i], chunk, 970 + i + 17 * rank)
        return (lambda c: mult(bufs[4:7], bufs[0:2], bufs[2:4], c)), (lambda: [be.free(b) for b in bufs])

    run("cfg4_bfv_hps_evalmult_core_n32768_q16_r17", n, sq + sr, mk_cfg4b, None)
    mult.close()
    rq.close()
    ring.close()
    # the same product with the reference's default BFV technique, HPSPOVERQ (R = 16 towers)
    q2, psi2 = pkg.gen_params(60, n, 2 * sq)
    ring = pkg.Ring(be, n, q2, psi2)
    rq = pkg.RingView(ring, 0, sq)
    neg, qinv = pkg.hps_p_over_q_tables(q2[:sq], q2[sq:])
    frac2, modq2 = pkg.hps_scale_tables(q2[sq:], q2[:sq], t)
    mult = pkg.BfvHpsPOverQMult(ring, sq, neg, qinv, frac2, modq2)

    def mk_cfg4c(chunk):
        w = rq.words(chunk) * 8
        bufs = [be.alloc(w) for _ in range(7)]
        for i in range(4):
            rq.fill_uniform(bufs[i], chunk, 990 + i + 17 * rank)
        return (lambda c: mult(bufs[4:7], bufs[0:2], bufs[2:4], c)), (lambda: [be.free(b) for b in bufs])

    run("cfg4_bfv_hpspoverq_evalmult_core_n32768_q16_r16", n, 2 * sq, mk_cfg4c, None)
    mult.close()
    rq.close()
    ring.close()
    be.close()
    if rank == 0:
        if args.cpu_sample > 0:
            try:
                result["cfg4_cpu_reference"] = cpu_bfv(n, sq, sr, args.cpu_sample, t)
            except Exception as ex:                 # noqa: BLE001
                result["cfg4_cpu_reference"] = {"error": str(ex)[:200]}
        os.write(out_fd, (json.dumps(result) + "\n").encode())
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
