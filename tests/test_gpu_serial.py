"""Device-resident polys <-> the reference's binary wire format (fhe_serialize / fhe_deserialize through the C ABI):
streams byte-identical to what the reference's Serial::Serialize writes (golden vectors from oracle/_ref), the
reference's streams loaded into HBM, round trips at full size."""
import base64
import hashlib
import json
import os

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = json.load(open(os.path.join(ROOT, "tests", "golden", "serial_vectors.json")))["cases"]
pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("case", GOLD, ids=lambda c: f"n{c['n']}_t{c['towers']}_f{c['format']}{'_detached' if c['detached'] else ''}")
def test_golden_streams_through_the_device(pkg, backend, oracle, case):
    n, towers, fmt = case["n"], case["towers"], case["format"]
    q, psi = np.array(case["q"], np.uint64), np.array(case["psi"], np.uint64)
    ring = pkg.Ring(backend, n, q, psi)
    data = oracle.fill_ring(1, n, q, case["seed"])
    poly = pkg.DCRTPolyBatch(ring, 1, fmt, data)
    if not case["detached"]:
        blob = poly.Serialize()
        assert blob.size == case["bytes"] == ring.serial_bytes()
        assert hashlib.sha256(blob.tobytes()).hexdigest() == case["sha256"]
    if "stream_b64" in case:
        ref = base64.b64decode(case["stream_b64"])
        back = pkg.DCRTPolyBatch.Deserialize(ring, ref, 1)
        assert back.format == fmt
        assert np.array_equal(back.to_host().reshape(-1), data.reshape(-1))
    ring.close()


@pytest.mark.parametrize("n,towers,batch", [(1, 1, 2), (2, 2, 3), (64, 3, 5), (1 << 12, 9, 4), (1 << 16, 40, 3)])
def test_batch_round_trip(pkg, backend, n, towers, batch):
    q, psi = pkg.gen_params(60 if n > 1 else 30, max(n, 2), towers)
    ring = pkg.Ring(backend, n, q, psi if n > 1 else None)
    a = pkg.DCRTPolyBatch(ring, batch, pkg.COEFFICIENT)
    ring.fill_uniform(a.dptr, batch, 77)
    host = a.to_host()
    blob = a.Serialize()
    size = ring.serial_bytes()
    assert blob.size == size * batch
    # every stream holds its poly's towers at the offsets the host-only template reports
    nat = __import__("importlib").import_module("upmem--openfhe_b200._native").lib
    off = np.zeros(towers, np.uint64)
    tmpl = np.zeros(size, np.uint8)
    psi_ptr = psi.ctypes.data if n > 1 else None
    assert nat.fhe_serial_template(n, towers, q.ctypes.data, psi_ptr, 1, tmpl.ctypes.data, size, off.ctypes.data) == size
    for b in range(batch):
        one = blob[b * size:(b + 1) * size].copy()
        for t in range(towers):
            lo = int(off[t])
            assert np.array_equal(one[lo:lo + 8 * n].copy().view(np.uint64), host[b, t])
            one[lo:lo + 8 * n] = 0
        assert np.array_equal(one, tmpl)
    back = pkg.DCRTPolyBatch.Deserialize(ring, blob, batch)
    assert back.format == pkg.COEFFICIENT
    assert backend.checksum(back.dptr, back.nwords) == backend.checksum(a.dptr, a.nwords)
    assert np.array_equal(back.to_host(), host)
    ring.close()


def test_rejected_stream_leaves_the_destination_untouched(pkg, backend):
    n, towers = 256, 4
    q, psi = pkg.gen_params(60, n, towers)
    ring = pkg.Ring(backend, n, q, psi)
    other = pkg.Ring(backend, n, q[::-1].copy(), psi[::-1].copy())
    a = pkg.DCRTPolyBatch(ring, 2)
    ring.fill_uniform(a.dptr, 2, 5)
    blob = a.Serialize()
    dst = pkg.DCRTPolyBatch(other, 2)
    other.fill_uniform(dst.dptr, 2, 6)
    before = dst.to_host().copy()
    with pytest.raises(Exception, match="modulus"):
        other.deserialize(dst.dptr, blob, 2)
    bad = blob.copy()
    bad[ring.serial_bytes() + 1] = 9                             # second stream: class version from a later library
    with pytest.raises(Exception, match="later version"):
        ring.deserialize(dst.dptr, bad, 2)
    with pytest.raises(Exception):
        ring.deserialize(dst.dptr, blob[:-3].copy(), 2)          # truncated
    assert np.array_equal(dst.to_host(), before)
    fmt, used = ring.deserialize(dst.dptr, blob, 1)              # a prefix of the batch is fine
    assert (fmt, used) == (pkg.EVALUATION, ring.serial_bytes())
    ring.close(); other.close()


def test_serialize_after_device_ops(pkg, backend, oracle):
    """the stream of a computed poly equals the stream of the oracle's result (headers + payload)"""
    n, towers = 1 << 10, 5
    q, psi = pkg.gen_params(60, n, towers)
    ring = pkg.Ring(backend, n, q, psi)
    x, y = oracle.fill_ring(2, n, q, 1), oracle.fill_ring(2, n, q, 2)
    a, b = pkg.DCRTPolyBatch(ring, 2, pkg.EVALUATION, x), pkg.DCRTPolyBatch(ring, 2, pkg.EVALUATION, y)
    a *= b
    a.SwitchFormat()
    octx = oracle.ctx(n, q, psi)
    exp = octx.ntt(octx.elementwise(2, x.reshape(2, towers, n), y.reshape(2, towers, n)), inverse=True)
    want = pkg.DCRTPolyBatch(ring, 2, pkg.COEFFICIENT, exp).Serialize()
    assert a.Serialize().tobytes() == want.tobytes()
    ring.close()
