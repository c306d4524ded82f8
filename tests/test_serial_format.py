"""Wire format of a DCRTPoly (Serial::Serialize / Deserialize, SerType::BINARY; utils/serial.h:99-115): the host half
of fhe_serialize / fhe_deserialize (csrc/serial.h) against streams written by the reference itself — the committed
golden vectors (tests/golden/serial_vectors.json) and, where oracle/_ref is built, fresh ones.  No device needed."""
import base64
import ctypes
import hashlib
import importlib
import json
import os

import numpy as np
import pytest

from tests import oracle_py

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
vp, sz, ci, u32 = ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int, ctypes.c_uint32
GOLD = json.load(open(os.path.join(ROOT, "tests", "golden", "serial_vectors.json")))["cases"]


@pytest.fixture(scope="module")
def lib(pkg):
    return importlib.import_module("upmem--openfhe_b200._native").lib


def build_stream(lib, n, q, psi, fmt, data):
    """headers from the library, payload dropped into the holes it reports"""
    towers = len(q)
    off = np.zeros(towers, np.uint64)
    size = lib.fhe_serial_template(n, towers, q.ctypes.data, psi.ctypes.data, fmt, None, 0, off.ctypes.data)
    blob = np.zeros(size, np.uint8)
    assert lib.fhe_serial_template(n, towers, q.ctypes.data, psi.ctypes.data, fmt, blob.ctypes.data, size, off.ctypes.data) == size
    for t in range(towers):
        blob[int(off[t]):int(off[t]) + 8 * n] = data.reshape(towers, n)[t].view(np.uint8)
    return blob, off


def parse(lib, n, q, psi, blob):
    off = np.zeros(len(q), np.uint64)
    fmt, used = ci(-1), sz(0)
    rc = lib.fhe_serial_parse(n, len(q), q.ctypes.data, psi.ctypes.data if psi is not None else None, blob.ctypes.data, blob.size,
                              off.ctypes.data, ctypes.byref(fmt), ctypes.byref(used))
    return rc, off, fmt.value, used.value


@pytest.mark.parametrize("case", GOLD, ids=lambda c: f"n{c['n']}_t{c['towers']}_f{c['format']}{'_detached' if c['detached'] else ''}")
def test_golden_streams(lib, oracle, case):
    n, towers, fmt = case["n"], case["towers"], case["format"]
    q, psi = np.array(case["q"], np.uint64), np.array(case["psi"], np.uint64)
    data = oracle.fill_ring(1, n, q, case["seed"])
    if not case["detached"]:                                   # writer: byte-identical to the reference's stream
        blob, _ = build_stream(lib, n, q, psi, fmt, data)
        assert blob.size == case["bytes"]
        assert hashlib.sha256(blob.tobytes()).hexdigest() == case["sha256"]
        if "stream_b64" in case:
            assert blob.tobytes() == base64.b64decode(case["stream_b64"])
    if "stream_b64" in case:                                   # reader: finds the payloads of the reference's stream
        ref = np.frombuffer(base64.b64decode(case["stream_b64"]), np.uint8).copy()
        rc, off, f, used = parse(lib, n, q, psi, ref)
        assert rc == 0 and f == fmt and used == ref.size
        for t in range(towers):
            got = ref[int(off[t]):int(off[t]) + 8 * n].view(np.uint64)
            assert np.array_equal(got, data.reshape(towers, n)[t])


def test_reader_rejects_what_the_reference_rejects_or_another_ring(lib, oracle):
    n, towers = 8, 3
    q, psi = oracle.gen_params(60, n, towers)
    data = oracle.fill_ring(1, n, q, 3)
    blob, off = build_stream(lib, n, q, psi, 0, data)
    assert parse(lib, n, q, psi, blob)[0] == 0
    assert parse(lib, n, q, None, blob)[0] == 0                # a ring without roots does not check them
    bad = blob.copy(); bad[0] = 0                              # big-endian flag
    assert parse(lib, n, q, psi, bad)[0] != 0
    bad = blob.copy(); bad[1] = 2                              # DCRTPolyImpl from a later library (dcrtpoly.h:358-361 throws)
    assert parse(lib, n, q, psi, bad)[0] != 0
    for cut in (1, 20, int(off[0]) + 5, blob.size - 1):        # truncated
        assert parse(lib, n, q, psi, blob[:cut].copy())[0] != 0
    q2 = q.copy(); q2[1] = q[2]; q2[2] = q[1]                  # same moduli, other order
    assert parse(lib, n, q2, psi, blob)[0] != 0
    psi2 = psi.copy(); psi2[0] ^= 1
    assert parse(lib, n, q, psi2, blob)[0] != 0
    assert parse(lib, 2 * n, q, psi, blob)[0] != 0
    assert parse(lib, n, q[:2].copy(), psi[:2].copy(), blob)[0] != 0
    bad = blob.copy()                                          # one tower claims the other format
    f_at = int(off[0]) + 8 * n + 4 + 8                         # tower 0: NativeInteger version, modulus, then "f"
    bad[f_at] = 1
    assert parse(lib, n, q, psi, bad)[0] != 0


@pytest.mark.skipif(not os.path.exists(oracle_py.REF_SO), reason="oracle/_ref not built here")
@pytest.mark.parametrize("n,towers,fmt", [(4, 1, 0), (32, 2, 1), (256, 7, 0), (2048, 30, 1), (64, 45, 0)])
def test_fresh_streams_against_the_reference(lib, oracle, n, towers, fmt):
    R = ctypes.CDLL(oracle_py.REF_SO)
    R.ref_ctx_create.restype = vp; R.ref_ctx_create.argtypes = [u32, u32, vp, vp]
    R.ref_polys_load.restype = vp; R.ref_polys_load.argtypes = [vp, vp, sz, ci]
    R.ref_polys_store.argtypes = [vp, vp]
    R.ref_polys_free.argtypes = [vp]
    R.ref_polys_format.argtypes = [vp]
    R.ref_polys_deserialize.restype = vp; R.ref_polys_deserialize.argtypes = [vp, vp, sz]
    for f in (R.ref_polys_serialize, R.ref_polys_serialize_detached):
        f.restype = sz; f.argtypes = [vp, sz, vp, sz]
    q, psi = oracle.gen_params(60, n, towers)
    ctx = R.ref_ctx_create(n, towers, q.ctypes.data, psi.ctypes.data)
    data = oracle.fill_ring(1, n, q, 1000 + n)
    h = R.ref_polys_load(ctx, data.ctypes.data, 1, fmt)
    size = R.ref_polys_serialize(h, 0, None, 0)
    ref = np.zeros(size, np.uint8)
    R.ref_polys_serialize(h, 0, ref.ctypes.data, size)
    mine, _ = build_stream(lib, n, q, psi, fmt, data)
    assert mine.tobytes() == ref.tobytes()
    # the reference reads the stream this library wrote
    back = R.ref_polys_deserialize(ctx, mine.ctypes.data, mine.size)
    assert back, "reference Serial::Deserialize rejected the stream"
    got = np.empty_like(data)
    R.ref_polys_store(back, got.ctypes.data)
    assert np.array_equal(got, data) and R.ref_polys_format(back) == fmt
    # and this library reads a stream whose tower params are written inline
    size = R.ref_polys_serialize_detached(h, 0, None, 0)
    det = np.zeros(size, np.uint8)
    R.ref_polys_serialize_detached(h, 0, det.ctypes.data, size)
    assert det.size > ref.size
    rc, off, f, used = parse(lib, n, q, psi, det)
    assert rc == 0 and f == fmt and used == det.size
    for t in range(towers):
        assert np.array_equal(det[int(off[t]):int(off[t]) + 8 * n].view(np.uint64), data.reshape(towers, n)[t])
    R.ref_polys_free(h); R.ref_polys_free(back)
