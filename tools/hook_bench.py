"""The reference's own DCRTPoly (hooked build, oracle/_ref/libopenfhe_hooked.so) on a chain of operators: CPU path vs
B200 through the live intercept in its resident form.  One step = for every polynomial of the batch:
    a += b; a *= b; a.SwitchFormat(); a.SwitchFormat(); a -= b        (5 operators, 2 of them transforms)
The GPU figure includes the one upload of a and b and the one download of a (ref_polys_store_mt).
Both modes run the reference's loop over the batch on every host thread (its callers are OpenMP loops); in the GPU
mode every host thread enqueues on its own stream (pim/PimHook.h hook_stream).  `gpu_threads` = 1 gives the
single-thread figure of round 1.
Usage: python tools/hook_bench.py [logN] [towers] [batch] [chain_reps] [gpu_threads]"""
import ctypes, json, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tests import oracle_py
logn = int(sys.argv[1]) if len(sys.argv) > 1 else 14
towers = int(sys.argv[2]) if len(sys.argv) > 2 else 8
batch = int(sys.argv[3]) if len(sys.argv) > 3 else 64
reps = int(sys.argv[4]) if len(sys.argv) > 4 else 4
host_threads = max(1, len(os.sched_getaffinity(0)))
gpu_threads = int(sys.argv[5]) if len(sys.argv) > 5 else min(16, host_threads)
n = 1 << logn
vp, sz, ci, u32 = ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int, ctypes.c_uint32
H = ctypes.CDLL(os.path.join(os.path.dirname(oracle_py.REF_SO), "libopenfhe_hooked.so"))
H.ref_ctx_create.restype = vp; H.ref_ctx_create.argtypes = [u32, u32, vp, vp]
H.ref_polys_load.restype = vp; H.ref_polys_load.argtypes = [vp, vp, sz, ci]
H.ref_polys_store_mt.argtypes = [vp, vp, ci]; H.ref_polys_free.argtypes = [vp]
H.ref_polys_op.restype = ctypes.c_double; H.ref_polys_op.argtypes = [vp, vp, ci, ci]
H.ref_polys_switch_format.restype = ctypes.c_double; H.ref_polys_switch_format.argtypes = [vp, ci]
H.ref_polys_set_pim.argtypes = [vp, ci]
H.ref_max_threads.restype = ci
H.ref_last_error.restype = ctypes.c_char_p
p = lambda a: a.ctypes.data_as(vp)
O = oracle_py.load()
q, psi = O.gen_params(60, n, towers)
a, b = O.fill_ring(batch, n, q, 1), O.fill_ring(batch, n, q, 2)
ctx = H.ref_ctx_create(n, towers, p(q), p(psi))
res = {}
out = {}
for mode in ("cpu_1thread_per_poly_all_cores", "b200_resident"):
    threads = host_threads if mode.startswith("cpu") else gpu_threads
    best = None
    for trial in range(3):
        ha, hb = H.ref_polys_load(ctx, p(a), batch, 0), H.ref_polys_load(ctx, p(b), batch, 0)
        if mode == "b200_resident":
            assert H.ref_polys_set_pim(ha, 1) == 0, H.ref_last_error()
        r = np.empty_like(a)
        t0 = time.perf_counter()
        for _ in range(reps):
            assert H.ref_polys_op(ha, hb, 0, threads) >= 0, H.ref_last_error()
            H.ref_polys_op(ha, hb, 2, threads)
            H.ref_polys_switch_format(ha, threads)
            H.ref_polys_switch_format(ha, threads)
            H.ref_polys_op(ha, hb, 1, threads)
        H.ref_polys_store_mt(ha, p(r), threads)
        dt = time.perf_counter() - t0
        best = dt if best is None else min(best, dt)
        H.ref_polys_free(ha); H.ref_polys_free(hb)
    res[mode] = r
    out[mode] = {"seconds": best, "coeff_ops_per_s": 5 * reps * a.size / best, "threads": threads}
assert np.array_equal(res["cpu_1thread_per_poly_all_cores"], res["b200_resident"]), "GPU result differs from the CPU path"
out["speedup"] = out["cpu_1thread_per_poly_all_cores"]["seconds"] / out["b200_resident"]["seconds"]
out["config"] = {"N": n, "towers": towers, "batch": batch, "chain_reps": reps, "ops_per_rep": 5}
print(json.dumps(out))
