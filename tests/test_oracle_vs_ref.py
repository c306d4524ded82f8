"""Oracle restatement vs the reference's own compiled sources (oracle/_ref), on fresh random inputs.
Only where oracle/_ref was built (the container that has /root/reference); elsewhere the committed
golden vectors (tests/test_oracle_golden.py) carry the same evidence."""
import ctypes
import os

import numpy as np
import pytest

from tests import oracle_py

pytestmark = pytest.mark.skipif(not os.path.exists(oracle_py.REF_SO), reason="oracle/_ref not built here")
vp, sz, ci, u32, u64 = ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int, ctypes.c_uint32, ctypes.c_uint64


@pytest.fixture(scope="module")
def ref():
    R = ctypes.CDLL(oracle_py.REF_SO)
    R.ref_gen_params.argtypes = [u32, u32, u32, vp, vp]
    R.ref_root_of_unity.restype = u64
    R.ref_root_of_unity.argtypes = [u32, u64]
    R.ref_ctx_create.restype = vp
    R.ref_ctx_create.argtypes = [u32, u32, vp, vp]
    R.ref_ctx_destroy.argtypes = [vp]
    R.ref_polys_load.restype = vp
    R.ref_polys_load.argtypes = [vp, vp, sz, ci]
    R.ref_polys_store.argtypes = [vp, vp]
    R.ref_polys_free.argtypes = [vp]
    R.ref_polys_op.restype = ctypes.c_double
    R.ref_polys_op.argtypes = [vp, vp, ci, ci]
    R.ref_polys_switch_format.restype = ctypes.c_double
    R.ref_polys_switch_format.argtypes = [vp, ci]
    R.ref_polys_automorphism.restype = vp
    R.ref_polys_automorphism.argtypes = [vp, u32, ci]
    R.ref_precompute_automap.argtypes = [u32, u32, vp]
    R.ref_last_error.restype = ctypes.c_char_p
    return R


def p(a):
    return a.ctypes.data_as(vp)


@pytest.mark.parametrize("nbits,n,towers", [(20, 16, 2), (40, 512, 3), (60, 1 << 13, 2), (60, 1 << 14, 3), (55, 1 << 16, 2)])
def test_params_and_ops_match_reference(oracle, ref, nbits, n, towers):
    q, psi = oracle.gen_params(nbits, n, towers)
    qr, pr = np.zeros(towers, np.uint64), np.zeros(towers, np.uint64)
    ref.ref_gen_params(nbits, n, towers, p(qr), p(pr))
    assert np.array_equal(q, qr) and np.array_equal(psi, pr)
    rng = np.random.default_rng(n + nbits)
    batch = 2
    a = np.stack([rng.integers(0, int(m), size=(batch, n), dtype=np.uint64) for m in q], axis=1)
    b = np.stack([rng.integers(0, int(m), size=(batch, n), dtype=np.uint64) for m in q], axis=1)
    a[0, :, 0], b[0, :, 0] = q - 1, q - 1
    octx = oracle.ctx(n, q, psi)
    ctx = ref.ref_ctx_create(n, towers, p(q), p(psi))
    res = np.empty_like(a)
    for op in range(3):
        ha, hb = ref.ref_polys_load(ctx, p(a), batch, 0), ref.ref_polys_load(ctx, p(b), batch, 0)
        # in-place forms only: the fork's out-of-place Plus/Times copy NativeVectors, whose PimData
        # member is copied shallowly (pim/PimData.h:36-44) and double-frees under the dpu.h stub
        ref.ref_polys_op(ha, hb, op, 1)                      # += / -= / *=
        ref.ref_polys_store(ha, p(res))
        assert np.array_equal(res, octx.elementwise(op, a, b)), f"in-place op {op}"
        for h in (ha, hb):
            ref.ref_polys_free(h)
    ha = ref.ref_polys_load(ctx, p(a), batch, 1)
    ref.ref_polys_switch_format(ha, 1)
    ref.ref_polys_store(ha, p(res))
    assert np.array_equal(res, octx.ntt(a, inverse=False))
    ref.ref_polys_switch_format(ha, 1)
    ref.ref_polys_store(ha, p(res))
    assert np.array_equal(res, a)
    ref.ref_polys_free(ha)
    hb = ref.ref_polys_load(ctx, p(b), batch, 0)
    ref.ref_polys_switch_format(hb, 1)
    ref.ref_polys_store(hb, p(res))
    assert np.array_equal(res, octx.ntt(b, inverse=True))
    ref.ref_polys_free(hb)
    ref.ref_ctx_destroy(ctx)


@pytest.mark.parametrize("n,towers", [(8, 2), (1 << 10, 3), (1 << 14, 2)])
def test_automorphism_matches_reference(oracle, ref, n, towers):
    """DCRTPoly::AutomorphismTransform(k) and (k, precomp), both formats."""
    q, psi = oracle.gen_params(60, n, towers)
    rng = np.random.default_rng(n)
    batch = 2
    a = np.stack([rng.integers(0, int(m), size=(batch, n), dtype=np.uint64) for m in q], axis=1)
    a[0, :, :3] = 0                                     # negated zeros: the reference stores q
    octx = oracle.ctx(n, q, psi)
    ctx = ref.ref_ctx_create(n, towers, p(q), p(psi))
    res = np.empty_like(a)
    for k in (3, 5, 2 * n - 1, n + 1, 25 % (2 * n) | 1):
        mr = np.empty(n, np.uint32)
        ref.ref_precompute_automap(n, k, p(mr))
        assert np.array_equal(mr, oracle.automap(n, k))
        for fmt in (0, 1):
            ha = ref.ref_polys_load(ctx, p(a), batch, fmt)
            hr = ref.ref_polys_automorphism(ha, k, 0)
            assert hr, ref.ref_last_error()
            ref.ref_polys_store(hr, p(res))
            assert np.array_equal(res, octx.automorphism(a, k, fmt)), (k, fmt)
            ref.ref_polys_free(hr)
            if fmt == 0:
                hr = ref.ref_polys_automorphism(ha, k, 1)
                assert hr, ref.ref_last_error()
                ref.ref_polys_store(hr, p(res))
                assert np.array_equal(res, octx.automorphism(a, k, 0)), (k, "precomp")
                ref.ref_polys_free(hr)
            ref.ref_polys_free(ha)
    ha = ref.ref_polys_load(ctx, p(a), batch, 0)
    assert not ref.ref_polys_automorphism(ha, 4, 0) and b"not odd" in ref.ref_last_error()
    ref.ref_polys_free(ha)
    ref.ref_ctx_destroy(ctx)
