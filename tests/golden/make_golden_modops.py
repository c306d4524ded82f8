"""Generates tests/golden/modops_vectors.json from the REFERENCE (oracle/_ref/libopenfhe_ref.so):
NativeVector SwitchModulus / ModEq / ModByTwoEq / ModExpEq / ModInverseEq / MultiplyAndRoundEq /
DivideAndRoundEq (mubintvecnat.cpp:111-195, 370-487) on seeded inputs.  Build container only:
    make -C oracle ref && python tests/golden/make_golden_modops.py
"""
import ctypes
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from tests import oracle_py  # noqa: E402

MODS = [73, 8353, (1 << 40) + 15, 1152921504606748673]


def cases():
    for mod in MODS:
        yield from [(0, mod, mod + 1000, 0), (0, mod, max(3, mod // 3), 0), (1, mod, 2, 0), (2, mod, max(5, mod // 7), 0),
                    (3, mod, 0, 0), (5, mod, 65537, 0), (6, mod, mod + 5, 0), (7, mod, 0, 0),
                    (9, mod, 12345, max(7, mod // 5)), (10, mod, mod - 1, mod), (11, mod, 7, 0),
                    (12, mod, max(2, mod // 1000), 0)]


def inputs(O, kind, mod, n=64):
    x = O.fill_uniform(n, mod, 424242 + kind)
    if kind in (7, 8):
        x[x == 0] = 1
    x[0], x[1], x[2], x[3] = mod - 1, mod >> 1, (mod >> 1) + 1, 1
    return x


def main():
    O = oracle_py.load()
    R = ctypes.CDLL(oracle_py.REF_SO)
    R.ref_vec_modop.argtypes = [ctypes.c_int, ctypes.c_void_p, ctypes.c_size_t] + [ctypes.c_uint64] * 3
    out = []
    for kind, mod, p, q in cases():
        x = inputs(O, kind, mod)
        y = x.copy()
        rc = R.ref_vec_modop(kind, y.ctypes.data, y.size, mod, p, q)
        assert rc == 0
        out.append({"kind": kind, "mod": mod, "p": p, "q": q, "out": [int(v) for v in y]})
    with open(os.path.join(ROOT, "tests", "golden", "modops_vectors.json"), "w") as f:
        json.dump({"generator": "tests/golden/make_golden_modops.py over oracle/_ref (reference commit e70da11)",
                   "cases": out}, f)
    print("wrote", len(out), "cases")


if __name__ == "__main__":
    main()
