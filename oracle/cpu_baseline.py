"""CPU timing of the reference's path (TEST / MEASUREMENT INFRASTRUCTURE ONLY).

Used by bench.py's `cpu_baseline` leg and by `bench.py --impl reference`, nowhere else.
kind "reference": oracle/_ref/libopenfhe_ref.so — the reference's own DCRTPoly operator+= / *=
                  (dcrtpoly-impl.h:410-416, dcrtpoly.h:142-148) compiled from its sources.
kind "port"     : oracle/libfhe_oracle.so — our C restatement, when _ref was not built.
Threads: an OpenMP loop over the batch with every host thread (the reference's own parallelism is
OpenMP over the towers of ONE polynomial, <= T threads; the batch loop is the CPU's best case).
The thread count is taken from the CPUs this process may run on and passed explicitly
(`num_threads(...)` in the driver), never from OMP_NUM_THREADS: torchrun exports OMP_NUM_THREADS=1.
"""
import ctypes
import os
import subprocess
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF_SO = os.path.join(HERE, "_ref", "libopenfhe_ref.so")
PORT_SO = os.path.join(HERE, "libfhe_oracle.so")
vp, sz, ci, u32, u64 = ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int, ctypes.c_uint32, ctypes.c_uint64


def _p(a):
    return a.ctypes.data_as(vp)


def host_threads():
    """CPUs this process may run on (what `nproc` prints)."""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return max(1, os.cpu_count() or 1)


def claim_host_threads():
    """Must run before libgomp is loaded: the reference sizes its own thread pool from omp_get_max_threads() in a
    static constructor (utils/parallel.h:52, lib/utils/parallel.cpp:40), i.e. from OMP_NUM_THREADS.  Returns
    (threads, the value OMP_NUM_THREADS had)."""
    was = os.environ.get("OMP_NUM_THREADS")
    n = host_threads()
    os.environ["OMP_NUM_THREADS"] = str(n)
    return n, was


def _port():
    if not os.path.exists(PORT_SO):
        subprocess.check_call(["make", "-C", HERE, "oracle"])
    L = ctypes.CDLL(PORT_SO)
    L.orc_ctx_create.restype = vp
    L.orc_ctx_create.argtypes = [u32, u32, vp, vp]
    L.orc_dcrt_elementwise.argtypes = [vp, ci, vp, vp, vp, sz, ci]
    L.orc_dcrt_ntt.argtypes = [vp, ci, vp, sz, ci]
    L.orc_gen_dcrt_params.argtypes = [u32, u32, u32, vp, vp]
    L.orc_fill_uniform.argtypes = [vp, sz, u64, u64, u64]
    L.orc_max_threads.restype = ci
    return L


def _ref():
    R = ctypes.CDLL(REF_SO)
    R.ref_ctx_create.restype = vp
    R.ref_ctx_create.argtypes = [u32, u32, vp, vp]
    R.ref_polys_load.restype = vp
    R.ref_polys_load.argtypes = [vp, vp, sz, ci]
    R.ref_polys_free.argtypes = [vp]
    R.ref_polys_op.restype = ctypes.c_double
    R.ref_polys_op.argtypes = [vp, vp, ci, ci]
    R.ref_polys_switch_format.restype = ctypes.c_double
    R.ref_polys_switch_format.argtypes = [vp, ci]
    R.ref_max_threads.restype = ci
    R.ref_set_threads.argtypes = [ci]
    return R


def _inputs(L, n, towers, batch, nbits, seed):
    q = np.zeros(towers, np.uint64)
    psi = np.zeros(towers, np.uint64)
    L.orc_gen_dcrt_params(nbits, n, towers, _p(q), _p(psi))
    a = np.empty((batch, towers, n), np.uint64)
    b = np.empty_like(a)
    for bi in range(batch):
        for t in range(towers):
            first = (bi * towers + t) * n
            L.orc_fill_uniform(_p(a[bi, t]), n, int(q[t]), seed, first)
            L.orc_fill_uniform(_p(b[bi, t]), n, int(q[t]), seed + 1000, first)
    return q, psi, a, b


def time_addmul(n, towers, batch, nbits=60, steps=5, warmup=1, seed=20260002, threads=0):
    """One step = a += b then a *= b over the whole batch (2*batch*towers*n coefficient-ops).
    Returns dict(kind, cores, step_seconds (list), coeff_ops_per_step)."""
    P = _port()
    q, psi, a, b = _inputs(P, n, towers, batch, nbits, seed)
    ops = 2 * batch * towers * n
    times = []
    if os.path.exists(REF_SO):
        R = _ref()
        cores = threads or host_threads()
        ctx = R.ref_ctx_create(n, towers, _p(q), _p(psi))
        ha, hb = R.ref_polys_load(ctx, _p(a), batch, 0), R.ref_polys_load(ctx, _p(b), batch, 0)
        for i in range(warmup + steps):
            t = R.ref_polys_op(ha, hb, 0, cores) + R.ref_polys_op(ha, hb, 2, cores)
            if i >= warmup:
                times.append(t)
        R.ref_polys_free(ha); R.ref_polys_free(hb)
        kind = "reference"
    else:
        cores = threads or host_threads()
        ctx = P.orc_ctx_create(n, towers, _p(q), _p(psi))
        for i in range(warmup + steps):
            t0 = time.perf_counter()
            P.orc_dcrt_elementwise(ctx, 0, _p(a), _p(a), _p(b), batch, cores)
            P.orc_dcrt_elementwise(ctx, 2, _p(a), _p(a), _p(b), batch, cores)
            if i >= warmup:
                times.append(time.perf_counter() - t0)
        kind = "port"
    return {"kind": kind, "cores": int(cores), "step_seconds": times, "coeff_ops_per_step": ops}


def time_ntt(n, towers, batch, nbits=60, steps=3, warmup=1, seed=20260003, threads=0):
    """One step = forward then inverse transform of the whole batch (2*batch*towers*n coefficient-ops)."""
    P = _port()
    q, psi, a, _ = _inputs(P, n, towers, batch, nbits, seed)
    ops = 2 * batch * towers * n
    times = []
    if os.path.exists(REF_SO):
        R = _ref()
        cores = threads or host_threads()
        ctx = R.ref_ctx_create(n, towers, _p(q), _p(psi))
        ha = R.ref_polys_load(ctx, _p(a), batch, 1)
        for i in range(warmup + steps):
            t = R.ref_polys_switch_format(ha, cores) + R.ref_polys_switch_format(ha, cores)
            if i >= warmup:
                times.append(t)
        R.ref_polys_free(ha)
        kind = "reference"
    else:
        cores = threads or host_threads()
        ctx = P.orc_ctx_create(n, towers, _p(q), _p(psi))
        for i in range(warmup + steps):
            t0 = time.perf_counter()
            P.orc_dcrt_ntt(ctx, 0, _p(a), batch, cores)
            P.orc_dcrt_ntt(ctx, 1, _p(a), batch, cores)
            if i >= warmup:
                times.append(time.perf_counter() - t0)
        kind = "port"
    return {"kind": kind, "cores": int(cores), "step_seconds": times, "coeff_ops_per_step": ops}


def time_step(n, towers, batch, nbits=60, steps=5, warmup=1, seed=20260002, threads=0, faithful=False):
    """One step = the bench step on `batch` polynomials: a += b; a *= b; a.SwitchFormat(); a.SwitchFormat()
    (EVALUATION -> COEFFICIENT -> EVALUATION), in place, 4*batch*towers*n coefficient-ops.
    faithful=False: omp-for over the batch on `threads` (default: all) host threads, the CPU's best case.
    faithful=True : the reference's own parallelism only — serial over the batch, OpenMP over the towers of one
                    polynomial inside each call (dcrtpoly-impl.h:412, utils/parallel.h:112-118), <= towers threads."""
    P = _port()
    q, psi, a, b = _inputs(P, n, towers, batch, nbits, seed)
    ops = 4 * batch * towers * n
    times = []
    cores = threads or host_threads()
    if os.path.exists(REF_SO):
        R = _ref()
        R.ref_set_threads(cores)
        ctx = R.ref_ctx_create(n, towers, _p(q), _p(psi))
        ha, hb = R.ref_polys_load(ctx, _p(a), batch, 0), R.ref_polys_load(ctx, _p(b), batch, 0)
        outer = 1 if faithful else cores
        for i in range(warmup + steps):
            t = (R.ref_polys_op(ha, hb, 0, outer) + R.ref_polys_op(ha, hb, 2, outer) +
                 R.ref_polys_switch_format(ha, outer) + R.ref_polys_switch_format(ha, outer))
            if i >= warmup:
                times.append(t)
        R.ref_polys_free(ha); R.ref_polys_free(hb)
        kind = "reference"
        used = min(cores, towers, R.ref_max_threads()) if faithful else cores
    else:
        if faithful:
            raise RuntimeError("the reference-faithful figure needs oracle/_ref (the compiled reference)")
        ctx = P.orc_ctx_create(n, towers, _p(q), _p(psi))
        for i in range(warmup + steps):
            t0 = time.perf_counter()
            P.orc_dcrt_elementwise(ctx, 0, _p(a), _p(a), _p(b), batch, cores)
            P.orc_dcrt_elementwise(ctx, 2, _p(a), _p(a), _p(b), batch, cores)
            P.orc_dcrt_ntt(ctx, 1, _p(a), batch, cores)
            P.orc_dcrt_ntt(ctx, 0, _p(a), batch, cores)
            if i >= warmup:
                times.append(time.perf_counter() - t0)
        kind, used = "port", cores
    return {"kind": kind, "cores": int(used), "host_threads": host_threads(), "step_seconds": times,
            "coeff_ops_per_step": ops, "batch": batch}
