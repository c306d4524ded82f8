"""End-to-end (host operands) sweep of the chunk size / stream count of the host pipelines."""
import importlib, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pkg = importlib.import_module("upmem--openfhe_b200")
n, towers, batch = 1 << 14, 8, 256
for opts in ["streams=4,chunk_mb=32", "streams=4,chunk_mb=16", "streams=4,chunk_mb=8", "streams=4,chunk_mb=4", "streams=8,chunk_mb=8", "streams=8,chunk_mb=16", "streams=3,chunk_mb=16", "streams=2,chunk_mb=32"]:
    with pkg.Backend(0, opts + ",arena_mb=512") as be:
        q, psi = pkg.gen_params(60, n, towers)
        ring = pkg.Ring(be, n, q, psi)
        words = ring.words(batch)
        bufs = [be.host_alloc(words)[0] for _ in range(4)]
        d = be.alloc(words * 8)
        ring.fill_uniform(d, batch, 1); be.d2h(d, words, out=bufs[0])
        ring.fill_uniform(d, batch, 2); be.d2h(d, words, out=bufs[1])
        for _ in range(2):
            ring.multi_op_host([0, 2], [bufs[2], bufs[3]], bufs[0], bufs[1], batch)
        t0 = time.perf_counter()
        reps = 8
        for _ in range(reps):
            ring.multi_op_host([0, 2], [bufs[2], bufs[3]], bufs[0], bufs[1], batch)
        dt = (time.perf_counter() - t0) / reps
        print(f"{opts:28s} out-of-place {dt*1e3:7.2f} ms/step  {2*words/dt/1e9:6.2f} G coeff-ops/s  PCIe {words*8*2/dt/1e9:5.1f} GB/s each way")
        for _ in range(2):
            ring.chain_host([0, 2], bufs[0], bufs[1], batch)
        t0 = time.perf_counter()
        for _ in range(reps):
            ring.chain_host([0, 2], bufs[0], bufs[1], batch)
        dt = (time.perf_counter() - t0) / reps
        print(f"{opts:28s} in-place     {dt*1e3:7.2f} ms/step  {2*words/dt/1e9:6.2f} G coeff-ops/s  H2D {words*8*2/dt/1e9:5.1f} GB/s")
        ring.close()
