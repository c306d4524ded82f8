"""Short NTT workload for timing / ncu captures.
Usage: python tools/prof_ntt.py [logN] [towers] [batch] [reps] [algo]   (algo 0 persistent, 2 legacy two-pass)"""
import importlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pkg = importlib.import_module("upmem--openfhe_b200")
logn = int(sys.argv[1]) if len(sys.argv) > 1 else 16
towers = int(sys.argv[2]) if len(sys.argv) > 2 else 30
batch = int(sys.argv[3]) if len(sys.argv) > 3 else 64
reps = int(sys.argv[4]) if len(sys.argv) > 4 else 3
algo = int(sys.argv[5]) if len(sys.argv) > 5 else 0
n = 1 << logn
with pkg.Backend(0, "arena_mb=1024") as be:
    q, psi = pkg.gen_params(60, n, towers)
    ring = pkg.Ring(be, n, q, psi)
    words = ring.words(batch)
    x = be.alloc(words * 8)
    ring.fill_uniform(x, batch, 1)
    for _ in range(2):
        ring.ntt(x, batch, False, algo=algo); ring.ntt(x, batch, True, algo=algo)
    be.sync()
    be.timer_start(0)
    for _ in range(reps):
        ring.ntt(x, batch, False, algo=algo)
    f = be.timer_stop(0) / reps
    be.timer_start(0)
    for _ in range(reps):
        ring.ntt(x, batch, True, algo=algo)
    i = be.timer_stop(0) / reps
    print(f"N=2^{logn} T={towers} B={batch} algo={algo}: fwd {f:.3f} ms ({words/f/1e6:.1f} Gcoeff/s)  inv {i:.3f} ms ({words/i/1e6:.1f} Gcoeff/s)")
    ring.close()
