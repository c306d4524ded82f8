"""Golden vectors of the reference's binary wire format of a DCRTPoly, written by the REFERENCE itself
(oracle/_ref/libopenfhe_ref.so: Serial::Serialize(poly, ss, SerType::BINARY), utils/serial.h:99-103).

  python tests/golden/make_golden_serial.py      -> tests/golden/serial_vectors.json

Each case names (N, towers, format, seed); the tower payload is oracle fill_ring(1, N, q, seed).  Small cases carry
the whole stream (base64), large ones its SHA-256.  `detached` cases were serialised after giving every tower a
private copy of its params, so ILDCRTParams carries the tower params inline (reader coverage).
"""
import base64
import ctypes
import hashlib
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from tests import oracle_py  # noqa: E402

vp, sz, ci, u32 = ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int, ctypes.c_uint32

CASES = [  # n, towers, format, seed, detached, full bytes?
    (2, 1, 1, 11, False, True),
    (4, 2, 0, 12, False, True),
    (8, 3, 1, 13, False, True),
    (8, 3, 0, 14, True, True),
    (16, 5, 0, 15, False, True),
    (1024, 12, 1, 16, False, False),
    (1024, 12, 0, 17, True, False),
    (4096, 40, 0, 18, False, False),
]


def bind(R):
    R.ref_ctx_create.restype = vp
    R.ref_ctx_create.argtypes = [u32, u32, vp, vp]
    R.ref_polys_load.restype = vp
    R.ref_polys_load.argtypes = [vp, vp, sz, ci]
    R.ref_polys_free.argtypes = [vp]
    for f in (R.ref_polys_serialize, R.ref_polys_serialize_detached):
        f.restype = sz
        f.argtypes = [vp, sz, vp, sz]
    return R


def ref_stream(R, ctx, data, fmt, detached):
    h = R.ref_polys_load(ctx, data.ctypes.data, 1, fmt)
    f = R.ref_polys_serialize_detached if detached else R.ref_polys_serialize
    n = f(h, 0, None, 0)
    buf = np.zeros(n, np.uint8)
    f(h, 0, buf.ctypes.data, n)
    R.ref_polys_free(h)
    return buf


def main():
    O = oracle_py.load()
    R = bind(ctypes.CDLL(oracle_py.REF_SO))
    out = []
    for n, towers, fmt, seed, detached, full in CASES:
        q, psi = O.gen_params(60, n, towers)
        ctx = R.ref_ctx_create(n, towers, q.ctypes.data, psi.ctypes.data)
        data = O.fill_ring(1, n, q, seed)
        blob = ref_stream(R, ctx, data, fmt, detached)
        case = {"n": n, "towers": towers, "format": fmt, "seed": seed, "detached": detached, "bytes": int(blob.size),
                "q": [int(x) for x in q], "psi": [int(x) for x in psi], "sha256": hashlib.sha256(blob.tobytes()).hexdigest()}
        if full:
            case["stream_b64"] = base64.b64encode(blob.tobytes()).decode()
        out.append(case)
    path = os.path.join(ROOT, "tests", "golden", "serial_vectors.json")
    with open(path, "w") as f:
        json.dump({"source": "oracle/_ref (reference Serial::Serialize, SerType::BINARY, MATHBACKEND 4)", "cases": out}, f, indent=1)
    print(path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
