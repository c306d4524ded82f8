"""PCIe ceiling for the end-to-end path: H2D alone, D2H alone, both at once (pinned memory, C-ABI copies)."""
import importlib, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pkg = importlib.import_module("upmem--openfhe_b200")
N = pkg._native
words = 64 << 20          # 512 MB
with pkg.Backend(0, "streams=4,arena_mb=1024") as be:
    ha, pa = be.host_alloc(words)
    hb, pb = be.host_alloc(words)
    ha[:] = 1; hb[:] = 2
    da, db = be.alloc(words * 8), be.alloc(words * 8)
    def run(h2d, d2h, reps=5, pieces=1):
        be.sync()
        t0 = time.perf_counter()
        for _ in range(reps):
            step = words // pieces
            for k in range(pieces):
                if h2d: N.check(N.lib.fhe_h2d(be.h, da + k * step * 8, pa + k * step * 8, step * 8, 0))
                if d2h: N.check(N.lib.fhe_d2h(be.h, pb + k * step * 8, db + k * step * 8, step * 8, 1))
        be.sync()
        return words * 8 * reps / (time.perf_counter() - t0) / 1e9
    run(True, True, 1)
    print(f"H2D alone {run(True, False):.1f} GB/s; D2H alone {run(False, True):.1f} GB/s; both at once {run(True, True):.1f} GB/s each way")
    print(f"both at once in 16 pieces {run(True, True, 5, 16):.1f} GB/s each way; 64 pieces {run(True, True, 5, 64):.1f}")
