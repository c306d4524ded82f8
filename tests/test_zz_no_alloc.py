"""fhe_fast_rns_floor_q_noalloc (static device scratch, no allocation) against fhe_fast_rns_floor_q through the C ABI.
Collected last on purpose (the name)."""
import ctypes

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_noalloc_floor_matches_the_allocating_entry_point(pkg, backend, oracle):
    L = pkg._native.lib
    n, nq, nB, t, batch = 1 << 12, 3, 3, 65537, 4
    nb, T = nB + 1, nq + nB + 1
    q, psi = oracle.gen_params(55, n, T)
    ring = pkg.Ring(backend, n, q, psi)
    z = pkg.Behz(ring, nq, t)
    words = batch * T * n
    da, db = backend.alloc(words * 8), backend.alloc(words * 8)
    reserved = (ctypes.c_size_t(), ctypes.c_size_t())
    for call, seed in enumerate((901, 902)):                       # two calls in a row: the array keeps the last call's contents
        y = oracle.fill_ring(batch, n, q, seed)
        y[0, :, 0] = q - np.uint64(1)
        backend.h2d(da, y); backend.h2d(db, y)
        pkg._native.check(L.fhe_fast_rns_floor_q(z.h, da, batch, 1, 0))
        pkg._native.check(L.fhe_arena_stats(backend.h, ctypes.byref(reserved[0]), None))
        pkg._native.check(L.fhe_fast_rns_floor_q_noalloc(z.h, db, batch, 1, 0))
        pkg._native.check(L.fhe_arena_stats(backend.h, ctypes.byref(reserved[1]), None))
        backend.sync()
        want, got = backend.d2h(da, words), backend.d2h(db, words)
        assert np.array_equal(got, want), f"call {call}"           # integer work: bit exact
        assert reserved[0].value == reserved[1].value, "the arena grew during the no-alloc call"
        if call == 0:
            assert np.array_equal(got.reshape(batch, T, n), oracle.behz(n, q, psi, nq, t).fast_rns_floor_q(y))
    # beyond the static array: 2^20 words = batch * nb * n  ->  batch 64 is the limit at this shape
    limit = (1 << 20) // (nb * n)
    big = backend.alloc((limit + 1) * T * n * 8)
    assert L.fhe_fast_rns_floor_q_noalloc(z.h, big, limit + 1, 0, 0) != 0
    assert b"exceeds the static scratch" in L.fhe_last_error()
    y = oracle.fill_ring(limit, n, q, 903)
    backend.h2d(big, y)
    pkg._native.check(L.fhe_fast_rns_floor_q_noalloc(z.h, big, limit, 0, 0))      # the largest batch that fits
    got = backend.d2h(big, limit * T * n).reshape(limit, T, n)
    backend.h2d(big, y)
    pkg._native.check(L.fhe_fast_rns_floor_q(z.h, big, limit, 0, 0))
    assert np.array_equal(got, backend.d2h(big, limit * T * n).reshape(limit, T, n))
    for p in (da, db, big):
        backend.free(p)
    z.close()
    ring.close()
