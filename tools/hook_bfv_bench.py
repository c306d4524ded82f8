"""BASELINE configs[4] shape through the LIVE intercept: the RNS core of BFVrns EvalMult (HPS) written against the
reference's own DCRTPoly (oracle/ref_driver.cpp bfv_hps_mult_impl: ExpandCRTBasis x4, Times / +=, SetFormat,
ScaleAndRound, SwitchCRTBasis), N = 2^15, 16 + 17 towers.
  cpu   the hooked build's CPU path, one ciphertext pair per host thread (all threads)
  b200  the same code with the operands attached (SetPim): one launch sequence per pair, operands uploaded once,
        the three result polys downloaded (ref_polys_store) inside the timed region
Both modes run one ciphertext pair per host thread; in the b200 mode every host thread enqueues on its own stream
(pim/PimHook.h hook_stream).  gpu_threads = 1 gives the single-thread figure of round 1.
Usage: python tools/hook_bfv_bench.py [pairs] [hps|hpspoverq] [gpu_threads]   ->  one JSON line
hpspoverq: the reference's default BFV technique (bfv_hps_p_over_q_mult_impl: ExpandCRTBasis x2, SetFormat +
FastExpandCRTBasisPloverQ + SetFormat x2, Times / +=, SetFormat, ScaleAndRound to Q), 16 + 16 towers."""
import ctypes
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tests import oracle_py  # noqa: E402

pairs = int(sys.argv[1]) if len(sys.argv) > 1 else 16
tech = sys.argv[2] if len(sys.argv) > 2 else "hps"
host_threads = max(1, len(os.sched_getaffinity(0)))
gpu_threads = int(sys.argv[3]) if len(sys.argv) > 3 else min(16, host_threads)
n, sq, sr, t = 1 << 15, 16, (17 if tech == "hps" else 16), 65537
vp, sz, ci, u32, u64 = ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int, ctypes.c_uint32, ctypes.c_uint64
H = ctypes.CDLL(os.path.join(os.path.dirname(oracle_py.REF_SO), "libopenfhe_hooked.so"))
H.ref_ctx_create.restype = vp; H.ref_ctx_create.argtypes = [u32, u32, vp, vp]
H.ref_polys_load.restype = vp; H.ref_polys_load.argtypes = [vp, vp, sz, ci]
H.ref_polys_store_mt.argtypes = [vp, vp, ci]; H.ref_polys_free.argtypes = [vp]
H.ref_polys_set_pim.argtypes = [vp, ci]
H.ref_max_threads.restype = ci
H.ref_last_error.restype = ctypes.c_char_p
mult_timed = H.ref_bfv_hps_mult_timed if tech == "hps" else H.ref_bfv_hps_p_over_q_mult_timed
mult_timed.argtypes = [vp, vp, vp, vp, vp, vp, u64, ctypes.POINTER(vp), ci, ctypes.POINTER(ctypes.c_double)]
p = lambda a: a.ctypes.data_as(vp)
O = oracle_py.load()
q, psi = O.gen_params(60, n, sq + sr)
mk = lambda a, b: H.ref_ctx_create(n, b - a, p(q[a:b].copy()), p(psi[a:b].copy()))
cQ, cR, cQR = mk(0, sq), mk(sq, sq + sr), mk(0, sq + sr)
data = [O.fill_ring(pairs, n, q[:sq].copy(), 700 + k) for k in range(4)]
out, results = {"technique": tech, "N": n, "towers_q": sq, "towers_r": sr, "pairs": pairs}, {}
for mode in ("cpu", "b200"):
    threads = host_threads if mode == "cpu" else gpu_threads
    best = None
    for trial in range(3):
        hs = [H.ref_polys_load(cQ, p(d), pairs, 0) for d in data]
        if mode == "b200":
            for h in hs:
                assert H.ref_polys_set_pim(h, 1) == 0, H.ref_last_error()
        outs, sec = (vp * 3)(), ctypes.c_double()
        res = [np.empty((pairs, sq, n), np.uint64) for _ in range(3)]
        t0 = time.perf_counter()
        assert mult_timed(hs[0], hs[1], hs[2], hs[3], cR, cQR, t, outs, threads, ctypes.byref(sec)) == 0, H.ref_last_error()
        for k in range(3):
            H.ref_polys_store_mt(outs[k], p(res[k]), threads)          # download (b200) / copy out (cpu)
        dt = time.perf_counter() - t0
        if best is None or dt < best:
            best, loop = dt, sec.value
        for k in range(3):
            H.ref_polys_free(outs[k])
        for h in hs:
            H.ref_polys_free(h)
    results[mode] = res
    out[mode] = {"seconds": best, "ciphertext_products_per_s": pairs / best, "threads": threads,
                 "loop_seconds": loop, "note": "seconds = table construction (host BigInteger, both modes) + products + results on the host"}
out["bit_exact"] = all(np.array_equal(results["cpu"][k], results["b200"][k]) for k in range(3))
out["speedup"] = out["cpu"]["seconds"] / out["b200"]["seconds"]
print(json.dumps(out))
