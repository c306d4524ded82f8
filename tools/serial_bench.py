"""Wire-format throughput of device-resident polys: fhe_serialize / fhe_deserialize beside a plain d2h / h2d of the same
towers, pageable and pinned host memory.  N = 2^16, 40 towers (21 MB per poly), batch 16.

  python tools/serial_bench.py  ->  one JSON line
"""
import ctypes
import importlib
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    pkg = importlib.import_module("upmem--openfhe_b200")
    N = pkg._native
    be = pkg.Backend(0, "streams=2,arena_mb=1024")
    n, towers, batch = 1 << 16, 40, 16
    q, psi = pkg.gen_params(60, n, towers)
    ring = pkg.Ring(be, n, q, psi)
    words = ring.words(batch)
    d = be.alloc(words * 8)
    ring.fill_uniform(d, batch, 1)
    be.sync()
    size = ring.serial_bytes() * batch
    pinned_words = (size + 7) // 8
    harr, hptr = be.host_alloc(pinned_words)
    pageable = np.empty(size, np.uint8)
    plain = np.empty(words, np.uint64)
    fmt, used = ctypes.c_int(), ctypes.c_size_t()

    def timed(f, reps=5):
        f()
        be.sync()
        best = None
        for _ in range(reps):
            t0 = time.perf_counter()
            f()
            be.sync()
            dt = time.perf_counter() - t0
            best = dt if best is None else min(best, dt)
        return best

    res = {"N": n, "towers": towers, "batch": batch, "stream_bytes": size, "payload_bytes": words * 8}
    cases = {     # order matters: the pinned block must hold streams when deserialize_pinned runs
        "d2h_pageable": lambda: be.d2h(d, words, out=plain),
        "d2h_pinned": lambda: N.check(N.lib.fhe_d2h(be.h, hptr, d, words * 8, 0)),
        "h2d_pageable": lambda: be.h2d(d, plain),
        "h2d_pinned": lambda: N.check(N.lib.fhe_h2d(be.h, d, hptr, words * 8, 0)),
        "serialize_pageable": lambda: N.check(N.lib.fhe_serialize(ring.h, pageable.ctypes.data, d, batch, 0, 0)),
        "serialize_pinned": lambda: N.check(N.lib.fhe_serialize(ring.h, hptr, d, batch, 0, 0)),
        "deserialize_pageable": lambda: N.check(N.lib.fhe_deserialize(ring.h, d, pageable.ctypes.data, size, batch, ctypes.byref(fmt), ctypes.byref(used), 0)),
        "deserialize_pinned": lambda: N.check(N.lib.fhe_deserialize(ring.h, d, hptr, size, batch, ctypes.byref(fmt), ctypes.byref(used), 0)),
    }
    for name, f in cases.items():
        s = timed(f)
        res[name] = {"ms": s * 1e3, "payload_gbs": words * 8 / s / 1e9}
    print(json.dumps(res))
    ring.close()
    be.close()


if __name__ == "__main__":
    main()
