"""the intercept LIVE inside the reference.  oracle/_ref/libopenfhe_hooked.so is the
reference's own DCRTPolyImpl / PolyImpl / NativeVector compiled against include/pim/*.h of this repository
with the four intercept calls patched in (oracle/hook_patch.py).  With a handle attached through the
reference's own SetPim, DCRTPoly += / -= / *= and SwitchFormat run on the B200; results must equal the
CPU path bit for bit.  The library is built where /root/reference exists and travels to the GPU box."""
import ctypes
import os

import numpy as np
import pytest

from tests import oracle_py

HOOKED_SO = os.path.join(os.path.dirname(oracle_py.REF_SO), "libopenfhe_hooked.so")
pytestmark = pytest.mark.skipif(not os.path.exists(HOOKED_SO), reason="oracle/_ref/libopenfhe_hooked.so not built")
vp, sz, ci, u32, u64 = ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int, ctypes.c_uint32, ctypes.c_uint64


def load_hooked():
    H = ctypes.CDLL(HOOKED_SO)
    H.ref_ctx_create.restype = vp
    H.ref_ctx_create.argtypes = [u32, u32, vp, vp]
    H.ref_ctx_destroy.argtypes = [vp]
    H.ref_polys_load.restype = vp
    H.ref_polys_load.argtypes = [vp, vp, sz, ci]
    H.ref_polys_store.argtypes = [vp, vp]
    H.ref_polys_free.argtypes = [vp]
    H.ref_polys_op.restype = ctypes.c_double
    H.ref_polys_op.argtypes = [vp, vp, ci, ci]
    H.ref_polys_switch_format.restype = ctypes.c_double
    H.ref_polys_switch_format.argtypes = [vp, ci]
    H.ref_polys_set_pim.argtypes = [vp, ci]
    H.ref_polys_format.argtypes = [vp]
    H.ref_last_error.restype = ctypes.c_char_p
    H.ref_hook_counters.argtypes = [vp, vp]
    H.ref_polys_automorphism.restype = vp
    H.ref_polys_automorphism.argtypes = [vp, u32, ci]
    H.ref_polys_rescale.argtypes = [vp, vp, u64]
    H.ref_polys_binary.restype = vp
    H.ref_polys_binary.argtypes = [vp, vp, ci]
    H.ref_polys_approx_switch.restype = vp
    H.ref_polys_approx_switch.argtypes = [vp, vp]
    H.ref_set_table_tweak.argtypes = [ctypes.c_int]
    H.ref_polys_crt_decompose.restype = vp
    H.ref_polys_crt_decompose.argtypes = [vp, u32]
    H.ref_keyswitch_bv.restype = vp
    H.ref_keyswitch_bv.argtypes = [vp, vp, vp]
    H.ref_behz_create.restype = vp
    H.ref_behz_create.argtypes = [vp, u32, u64]
    H.ref_behz_destroy.argtypes = [vp]
    H.ref_polys_behz_q_to_bsk.argtypes = [vp, vp]
    H.ref_polys_behz_floor.argtypes = [vp, vp]
    H.ref_polys_behz_conv_sk.argtypes = [vp, vp, vp]
    H.ref_polys_approx_mod_down.restype = vp
    H.ref_polys_approx_mod_down.argtypes = [vp, vp, vp, u64]
    H.ref_polys_approx_mod_up.argtypes = [vp, vp, vp]
    H.ref_bfv_hps_mult.argtypes = [vp, vp, vp, vp, vp, vp, u64, ctypes.POINTER(vp)]
    H.ref_polys_serialize.restype = sz
    H.ref_polys_serialize.argtypes = [vp, sz, vp, sz]
    H.ref_polys_deserialize.restype = vp
    H.ref_polys_deserialize.argtypes = [vp, vp, sz]
    return H


@pytest.fixture(scope="module")
def hooked():
    return load_hooked()


def counters(H):
    up, down = ctypes.c_long(), ctypes.c_long()
    H.ref_hook_counters(ctypes.byref(up), ctypes.byref(down))
    return up.value, down.value


def p(a):
    return a.ctypes.data_as(vp)


def run_ops(H, oracle, n, towers, batch, pim_on):
    q, psi = oracle.gen_params(60, n, towers)
    octx = oracle.ctx(n, q, psi)
    a = oracle.fill_ring(batch, n, q, 31)
    b = oracle.fill_ring(batch, n, q, 32)
    ctx = H.ref_ctx_create(n, towers, p(q), p(psi))
    res = np.empty_like(a)
    try:
        for op in range(3):
            ha, hb = H.ref_polys_load(ctx, p(a), batch, 0), H.ref_polys_load(ctx, p(b), batch, 0)
            if pim_on:
                assert H.ref_polys_set_pim(ha, 1) == 0, H.ref_last_error()
            assert H.ref_polys_op(ha, hb, op, 1) >= 0, H.ref_last_error()
            H.ref_polys_store(ha, p(res))
            assert np.array_equal(res, octx.elementwise(op, a, b)), f"op {op}"
            H.ref_polys_free(ha); H.ref_polys_free(hb)
        ha = H.ref_polys_load(ctx, p(a), batch, 1)                 # COEFFICIENT
        if pim_on:
            assert H.ref_polys_set_pim(ha, 1) == 0, H.ref_last_error()
        assert H.ref_polys_switch_format(ha, 1) >= 0, H.ref_last_error()
        assert H.ref_polys_format(ha) == 0                         # now EVALUATION
        H.ref_polys_store(ha, p(res))
        assert np.array_equal(res, octx.ntt(a, inverse=False))
        assert H.ref_polys_switch_format(ha, 1) >= 0, H.ref_last_error()
        assert H.ref_polys_format(ha) == 1
        H.ref_polys_store(ha, p(res))
        assert np.array_equal(res, a)
        # a transform on the GPU followed by a multiply on the CPU path (handle detached): formats stay coherent
        if pim_on:
            H.ref_polys_switch_format(ha, 1)
            assert H.ref_polys_set_pim(ha, 0) == 0
            hb = H.ref_polys_load(ctx, p(b), batch, 0)
            assert H.ref_polys_op(ha, hb, 2, 1) >= 0
            H.ref_polys_store(ha, p(res))
            assert np.array_equal(res, octx.elementwise(2, octx.ntt(a, inverse=False), b))
            H.ref_polys_free(hb)
        H.ref_polys_free(ha)
    finally:
        H.ref_ctx_destroy(ctx)


def test_hooked_build_without_handle_is_the_cpu_path(hooked, oracle):
    run_ops(hooked, oracle, 1 << 10, 2, 2, pim_on=False)


def test_hooked_build_fails_loudly_without_gpu(hooked, oracle):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    q, psi = oracle.gen_params(60, 1 << 10, 2)
    a = oracle.fill_ring(1, 1 << 10, q, 1)
    ctx = hooked.ref_ctx_create(1 << 10, 2, p(q), p(psi))
    ha = hooked.ref_polys_load(ctx, p(a), 1, 0)
    assert hooked.ref_polys_set_pim(ha, 1) == 1
    assert b"no CPU fallback" in hooked.ref_last_error()
    hooked.ref_polys_free(ha)
    hooked.ref_ctx_destroy(ctx)


@pytest.mark.gpu
@pytest.mark.parametrize("n,towers,batch", [(1 << 14, 2, 2), (1 << 12, 5, 1), (1 << 16, 3, 1), (8, 2, 1)])
def test_live_intercept_matches_cpu_path(hooked, oracle, pkg, n, towers, batch):
    before = pkg._native.lib.fhe_launch_count()
    run_ops(hooked, oracle, n, towers, batch, pim_on=True)
    assert pkg._native.lib.fhe_launch_count() > before, "the intercept did not launch any kernel"


@pytest.mark.gpu
def test_resident_form_keeps_towers_on_the_device(hooked, oracle, pkg):
    """Resident form of the intercept (pim/PimHook.h, oracle/hook_patch.py step 4): a chain of hooked operators on the
    reference's own DCRTPoly uploads each operand once and downloads the result once, when something else reads it;
    CPU-only methods called in between see the current values."""
    H = hooked
    n, towers, batch = 1 << 12, 4, 3
    q, psi = oracle.gen_params(60, n, towers)
    octx = oracle.ctx(n, q, psi)
    a = oracle.fill_ring(batch, n, q, 41)
    b = oracle.fill_ring(batch, n, q, 42)
    ctx = H.ref_ctx_create(n, towers, p(q), p(psi))
    ha, hb = H.ref_polys_load(ctx, p(a), batch, 0), H.ref_polys_load(ctx, p(b), batch, 0)
    assert H.ref_polys_set_pim(ha, 1) == 0, H.ref_last_error()
    up0, down0 = counters(H)
    launches0 = pkg._native.lib.fhe_launch_count()
    assert H.ref_polys_op(ha, hb, 0, 1) >= 0, H.ref_last_error()        # a += b
    assert H.ref_polys_op(ha, hb, 2, 1) >= 0                            # a *= b
    assert H.ref_polys_switch_format(ha, 1) >= 0                        # -> COEFFICIENT
    assert H.ref_polys_switch_format(ha, 1) >= 0                        # -> EVALUATION
    assert H.ref_polys_op(ha, hb, 1, 1) >= 0                            # a -= b
    up1, down1 = counters(H)
    assert (up1 - up0, down1 - down0) == (2 * batch, 0), "operands must be uploaded once and nothing downloaded yet"
    assert pkg._native.lib.fhe_launch_count() - launches0 >= 5 * batch
    want = octx.elementwise(1, octx.elementwise(2, octx.elementwise(0, a, b), b), b)
    res = np.empty_like(a)
    H.ref_polys_store(ha, p(res))                                       # const access: lazy download
    assert np.array_equal(res, want)
    up2, down2 = counters(H)
    assert (up2 - up1, down2 - down1) == (0, batch)
    H.ref_polys_store(ha, p(res))                                       # host copy is current now
    assert counters(H) == (up2, down2)
    assert H.ref_polys_op(ha, hb, 0, 1) >= 0                            # device copy was still valid: no new upload
    assert counters(H)[0] == up2
    want = octx.elementwise(0, want, b)
    # a CPU-only const method of the reference (AutomorphismTransform) sees the device result
    hr = H.ref_polys_automorphism(ha, 5, 0)
    assert hr, H.ref_last_error()
    res2 = np.empty_like(a)
    H.ref_polys_store(hr, p(res2))
    assert np.array_equal(res2, octx.automorphism(want, 5, 0))
    H.ref_polys_free(hr)
    # a CPU-only mutating method (DropLastElementAndScale) takes the current values and invalidates the device copy
    rest = H.ref_ctx_create(n, towers - 1, p(q[:-1].copy()), p(psi[:-1].copy()))
    assert H.ref_polys_rescale(ha, rest, 0) == 0, H.ref_last_error()
    res3 = np.empty((batch, towers - 1, n), np.uint64)
    H.ref_polys_store(ha, p(res3))
    assert np.array_equal(res3, oracle.rescale(n, q, psi, want, 0, 0))
    H.ref_polys_free(ha); H.ref_polys_free(hb)
    H.ref_ctx_destroy(rest); H.ref_ctx_destroy(ctx)


@pytest.mark.gpu
def test_resident_out_of_place_operators(hooked, oracle, pkg):
    """Plus / Minus / Times of the reference's DCRTPoly (dcrtpoly.h:164-200, dcrtpoly-impl.h:398-407) with a handle
    attached: the result is born on the device and keeps the handle, so (a + b) * b - b never touches the host until
    it is read."""
    H = hooked
    n, towers, batch = 1 << 13, 3, 2
    q, psi = oracle.gen_params(60, n, towers)
    octx = oracle.ctx(n, q, psi)
    a, b = oracle.fill_ring(batch, n, q, 51), oracle.fill_ring(batch, n, q, 52)
    ctx = H.ref_ctx_create(n, towers, p(q), p(psi))
    ha, hb = H.ref_polys_load(ctx, p(a), batch, 0), H.ref_polys_load(ctx, p(b), batch, 0)
    assert H.ref_polys_set_pim(ha, 1) == 0, H.ref_last_error()
    up0, down0 = counters(H)
    h1 = H.ref_polys_binary(ha, hb, 0)            # a + b
    h2 = H.ref_polys_binary(h1, hb, 2)            # (a + b) * b   (the handle travelled with the result)
    h3 = H.ref_polys_binary(h2, hb, 1)            # ... - b
    assert counters(H) == (up0 + 2 * batch, down0)
    res = np.empty_like(a)
    H.ref_polys_store(h3, p(res))
    want = octx.elementwise(1, octx.elementwise(2, octx.elementwise(0, a, b), b), b)
    assert np.array_equal(res, want)
    assert counters(H) == (up0 + 2 * batch, down0 + batch)
    H.ref_polys_store(ha, p(res))                 # the operands were not modified
    assert np.array_equal(res, a)
    for h in (ha, hb, h1, h2, h3):
        H.ref_polys_free(h)
    H.ref_ctx_destroy(ctx)


@pytest.mark.gpu
def test_resident_neighbouring_methods(hooked, oracle, pkg):
    """AutomorphismTransform, DropLastElementAndScale, ApproxSwitchCRTBasis and ApproxModDown of the reference's own
    DCRTPoly with the handle attached (SURVEY §8(f) ranks 3, 4 through the live intercept): a CKKS-like chain
    a *= b; sigma_5; ModDown; rescale stays on the device until the result is read, and equals the CPU path."""
    H = hooked
    n, sq, sp, batch = 1 << 12, 4, 2, 2
    q, psi = oracle.gen_params(60, n, sq + sp)
    ob = oracle.basis(n, q, psi, sq)
    oQP = oracle.ctx(n, q, psi)
    mk = lambda a, b: H.ref_ctx_create(n, b - a, p(q[a:b].copy()), p(psi[a:b].copy()))
    cQ, cP, cQP, cRest = mk(0, sq), mk(sq, sq + sp), mk(0, sq + sp), mk(0, sq - 1)
    a, b = oracle.fill_ring(batch, n, q, 61), oracle.fill_ring(batch, n, q, 62)
    ha, hb = H.ref_polys_load(cQP, p(a), batch, 0), H.ref_polys_load(cQP, p(b), batch, 0)
    assert H.ref_polys_set_pim(ha, 1) == 0, H.ref_last_error()
    up0, down0 = counters(H)
    launches0 = pkg._native.lib.fhe_launch_count()
    assert H.ref_polys_op(ha, hb, 2, 1) >= 0                               # a *= b over QP
    h1 = H.ref_polys_automorphism(ha, 5, 0)                                # sigma_5 (result carries the handle)
    assert h1, H.ref_last_error()
    h2 = H.ref_polys_approx_mod_down(h1, cQ, cP, 0)                        # QP -> Q
    assert h2, H.ref_last_error()
    assert H.ref_polys_rescale(h2, cRest, 0) == 0, H.ref_last_error()      # drop the last tower of Q and scale
    assert counters(H) == (up0 + 2 * batch, down0), "nothing may be downloaded before the result is read"
    assert pkg._native.lib.fhe_launch_count() - launches0 >= 4 * batch
    want = oracle.rescale(n, q[:sq].copy(), psi[:sq].copy(), ob.mod_down(oQP.automorphism(oQP.elementwise(2, a, b), 5, 0), 0), 0, 0)
    res = np.empty((batch, sq - 1, n), np.uint64)
    H.ref_polys_store(h2, p(res))
    assert np.array_equal(res, want)
    assert counters(H) == (up0 + 2 * batch, down0 + batch)
    # ApproxSwitchCRTBasis Q -> P on a resident operand in COEFFICIENT format
    x = oracle.fill_ring(batch, n, q[:sq].copy(), 63)
    hx = H.ref_polys_load(cQ, p(x), batch, 1)
    assert H.ref_polys_set_pim(hx, 1) == 0
    hs = H.ref_polys_approx_switch(hx, cP)
    res2 = np.empty((batch, sp, n), np.uint64)
    H.ref_polys_store(hs, p(res2))
    assert np.array_equal(res2, ob.approx_switch(x))
    # an even automorphism index still raises the reference's own error
    assert not H.ref_polys_automorphism(ha, 4, 0) and b"not odd" in H.ref_last_error()
    for h in (ha, hb, h1, h2, hx, hs):
        H.ref_polys_free(h)
    for c in (cQ, cP, cQP, cRest):
        H.ref_ctx_destroy(c)


@pytest.mark.gpu
@pytest.mark.parametrize("fmt", [0, 1])
def test_resident_behz_evalmult_core(hooked, oracle, pkg, fmt):
    """BFV EvalMult with the BEHZ technique on the reference's own DCRTPoly with the handle attached
    (bfvrns-leveledshe.cpp:275-297, 322-334, 367-388): FastBaseConvqToBskMontgomery of both operands, Times over QBsk,
    SetFormat(COEFFICIENT), FastRNSFloorq, FastBaseConvSK.  Two uploads, nothing downloaded until the result is read,
    and the result equals the oracle's (pinned against the CPU path of the same methods)."""
    H = hooked
    n, nq, t, batch = 1 << 11, 3, 65537, 2
    T = 2 * nq + 1
    q, psi = oracle.gen_params(55, n, T)
    ob, octx = oracle.behz(n, q, psi, nq, t), oracle.ctx(n, q, psi)
    cQBsk = H.ref_ctx_create(n, T, p(q), p(psi))
    cQ = H.ref_ctx_create(n, nq, p(q[:nq].copy()), p(psi[:nq].copy()))
    z = H.ref_behz_create(cQBsk, nq, t)
    a, b = oracle.fill_ring(batch, n, q[:nq].copy(), 91), oracle.fill_ring(batch, n, q[:nq].copy(), 92)
    ha, hb = H.ref_polys_load(cQ, p(a), batch, fmt), H.ref_polys_load(cQ, p(b), batch, fmt)
    assert H.ref_polys_set_pim(ha, 1) == 0 and H.ref_polys_set_pim(hb, 1) == 0, H.ref_last_error()
    up0, down0 = counters(H)
    launches0 = pkg._native.lib.fhe_launch_count()
    assert H.ref_polys_behz_q_to_bsk(ha, z) == 0, H.ref_last_error()
    assert H.ref_polys_behz_q_to_bsk(hb, z) == 0, H.ref_last_error()
    assert H.ref_polys_format(ha) == 0
    assert H.ref_polys_op(ha, hb, 2, 1) >= 0                               # a *= b over QBsk
    H.ref_polys_switch_format(ha, 1)                                       # -> COEFFICIENT
    assert H.ref_polys_behz_floor(ha, z) == 0, H.ref_last_error()
    assert H.ref_polys_behz_conv_sk(ha, z, cQ) == 0, H.ref_last_error()
    assert counters(H) == (up0 + 2 * batch, down0), "nothing may be downloaded before the result is read"
    assert pkg._native.lib.fhe_launch_count() - launches0 >= 8 * batch
    res = np.empty((batch, nq, n), np.uint64)
    H.ref_polys_store(ha, p(res))
    prod = octx.elementwise(2, ob.q_to_bsk_montgomery(a, fmt), ob.q_to_bsk_montgomery(b, fmt))
    want = ob.fast_base_conv_sk(ob.fast_rns_floor_q(octx.ntt(prod, inverse=True)))
    assert np.array_equal(res, want)
    assert counters(H) == (up0 + 2 * batch, down0 + batch)
    # a caller with another plaintext modulus' tables gets what ITS tables say (a second plan for the same ring)
    z2 = H.ref_behz_create(cQBsk, nq, 786433)
    y = oracle.fill_ring(batch, n, q, 93)
    hy = H.ref_polys_load(cQBsk, p(y), batch, 1)
    assert H.ref_polys_set_pim(hy, 1) == 0
    assert H.ref_polys_behz_floor(hy, z2) == 0, H.ref_last_error()
    got = np.empty((batch, T, n), np.uint64)
    H.ref_polys_store(hy, p(got))
    assert np.array_equal(got, oracle.behz(n, q, psi, nq, 786433).fast_rns_floor_q(y))
    for h in (ha, hb, hy):
        H.ref_polys_free(h)
    H.ref_behz_destroy(z); H.ref_behz_destroy(z2)
    H.ref_ctx_destroy(cQBsk); H.ref_ctx_destroy(cQ)


@pytest.mark.gpu
@pytest.mark.parametrize("fmt", [0, 1])
def test_resident_crt_decompose_and_bv_key_switch(hooked, oracle, pkg, fmt):
    """DCRTPolyImpl::CRTDecompose(0) with the handle attached returns device-resident digits (views of one block), and the
    BV key-switching core built on it (keyswitch-bv.cpp:314-338: sum of digit x key products) stays in HBM."""
    H = hooked
    n, T, batch = 1 << 11, 4, 2
    q, psi = oracle.gen_params(55, n, T)
    octx = oracle.ctx(n, q, psi)
    c = H.ref_ctx_create(n, T, p(q), p(psi))
    x = oracle.fill_ring(batch, n, q, 95)
    hx = H.ref_polys_load(c, p(x), batch, fmt)
    assert H.ref_polys_set_pim(hx, 1) == 0, H.ref_last_error()
    up0, down0 = counters(H)
    want = octx.crt_decompose(x, fmt)
    handles = [H.ref_polys_crt_decompose(hx, i) for i in range(T)]
    assert counters(H) == (up0 + batch, down0)
    for i, hd in enumerate(handles):
        got = np.empty((batch, T, n), np.uint64)
        H.ref_polys_store(hd, p(got))
        assert np.array_equal(got, want[i]), f"digit {i}"
        H.ref_polys_free(hd)
    if fmt == 0:
        kb, ka = oracle.fill_ring(T, n, q, 96), oracle.fill_ring(T, n, q, 97)
        x1 = np.ascontiguousarray(x[:1])
        h1 = H.ref_polys_load(c, p(x1), 1, 0)
        hb, ha = H.ref_polys_load(c, p(kb), T, 0), H.ref_polys_load(c, p(ka), T, 0)
        assert H.ref_polys_set_pim(h1, 1) == 0 and H.ref_polys_set_pim(hb, 1) == 0 and H.ref_polys_set_pim(ha, 1) == 0
        launches0 = pkg._native.lib.fhe_launch_count()
        hr = H.ref_keyswitch_bv(h1, hb, ha)
        assert pkg._native.lib.fhe_launch_count() - launches0 >= 2 * T
        got = np.empty((2, T, n), np.uint64)
        H.ref_polys_store(hr, p(got))
        dig = octx.crt_decompose(x1, 0)
        ct0 = np.zeros((1, T, n), np.uint64)
        ct1 = np.zeros((1, T, n), np.uint64)
        for i in range(T):
            ct0 = octx.elementwise(0, ct0, octx.elementwise(2, kb[i:i + 1], dig[i]))
            ct1 = octx.elementwise(0, ct1, octx.elementwise(2, ka[i:i + 1], dig[i]))
        assert np.array_equal(got[0:1], ct0) and np.array_equal(got[1:2], ct1)
        for h in (h1, hb, ha, hr):
            H.ref_polys_free(h)
    H.ref_polys_free(hx)
    H.ref_ctx_destroy(c)


@pytest.mark.gpu
def test_foreign_tables_stay_on_the_cpu_path(hooked, oracle, pkg):
    """The device plans derive their constants from the moduli.  A caller that passes OTHER constants (here
    [(Q/q_0)^-1]_{q_0} + 1) must get what the reference computes from ITS tables: ApproxSwitchCRTBasis, ApproxModUp
    and ApproxModDown notice the difference and leave the call to the CPU statements (no launch), with or without
    the handle the result is the same."""
    H = hooked
    n, sq, sp, batch = 1 << 10, 3, 2, 2
    q, psi = oracle.gen_params(60, n, sq + sp)
    mk = lambda a, b: H.ref_ctx_create(n, b - a, p(q[a:b].copy()), p(psi[a:b].copy()))
    cQ, cP, cQP = mk(0, sq), mk(sq, sq + sp), mk(0, sq + sp)
    x = oracle.fill_ring(batch, n, q[:sq].copy(), 81)
    y = oracle.fill_ring(batch, n, q, 82)
    ob = oracle.basis(n, q, psi, sq)

    def run(attach):
        out = []
        hx = H.ref_polys_load(cQ, p(x), batch, 1)
        if attach:
            assert H.ref_polys_set_pim(hx, 1) == 0, H.ref_last_error()
        hs = H.ref_polys_approx_switch(hx, cP)
        r = np.empty((batch, sp, n), np.uint64)
        H.ref_polys_store(hs, p(r))
        out.append(r)
        assert H.ref_polys_approx_mod_up(hx, cP, cQP) == 0, H.ref_last_error()
        r = np.empty((batch, sq + sp, n), np.uint64)
        H.ref_polys_store(hx, p(r))
        out.append(r)
        hy = H.ref_polys_load(cQP, p(y), batch, 0)
        if attach:
            assert H.ref_polys_set_pim(hy, 1) == 0
        hd = H.ref_polys_approx_mod_down(hy, cQ, cP, 0)
        r = np.empty((batch, sq, n), np.uint64)
        H.ref_polys_store(hd, p(r))
        out.append(r)
        for h in (hx, hs, hy, hd):
            H.ref_polys_free(h)
        return out

    H.ref_set_table_tweak(1)
    try:
        cpu = run(False)
        launches0 = pkg._native.lib.fhe_launch_count()
        dev = run(True)
        launched = pkg._native.lib.fhe_launch_count() - launches0
    finally:
        H.ref_set_table_tweak(0)
    for c, d in zip(cpu, dev):
        assert np.array_equal(c, d)
    assert not np.array_equal(cpu[0], ob.approx_switch(x)), "the tweak must change the result, else the test proves nothing"
    # with the moduli's own tables the same calls do run on the device
    launches1 = pkg._native.lib.fhe_launch_count()
    good = run(True)
    assert pkg._native.lib.fhe_launch_count() - launches1 > launched
    assert np.array_equal(good[0], ob.approx_switch(x))
    for c in (cQ, cP, cQP):
        H.ref_ctx_destroy(c)


@pytest.mark.gpu
@pytest.mark.parametrize("fmt", [0, 1])
def test_resident_mod_up_and_mod_reduce(hooked, oracle, pkg, fmt):
    """ApproxModUp and ModReduce through the live intercept: the hybrid key-switching core
    ModUp(x) * k -> ModDown over the reference's own DCRTPoly stays in HBM (two uploads, one download), and BGV's
    ModReduce keeps the format it was given.  Results equal the oracle's (which is pinned to the CPU path)."""
    H = hooked
    n, sq, sp, batch = 1 << 11, 3, 2, 2
    q, psi = oracle.gen_params(60, n, sq + sp)
    ob = oracle.basis(n, q, psi, sq)
    oQP = oracle.ctx(n, q, psi)
    mk = lambda a, b: H.ref_ctx_create(n, b - a, p(q[a:b].copy()), p(psi[a:b].copy()))
    cQ, cP, cQP, cRest = mk(0, sq), mk(sq, sq + sp), mk(0, sq + sp), mk(0, sq - 1)
    x = oracle.fill_ring(batch, n, q[:sq].copy(), 71)
    k = oracle.fill_ring(batch, n, q, 72)
    hx, hk = H.ref_polys_load(cQ, p(x), batch, fmt), H.ref_polys_load(cQP, p(k), batch, 0)
    assert H.ref_polys_set_pim(hx, 1) == 0, H.ref_last_error()
    up0, down0 = counters(H)
    assert H.ref_polys_approx_mod_up(hx, cP, cQP) == 0, H.ref_last_error()      # Q -> QP, EVALUATION
    assert H.ref_polys_format(hx) == 0
    assert H.ref_polys_op(hx, hk, 2, 1) >= 0                                    # * key over QP
    hd = H.ref_polys_approx_mod_down(hx, cQ, cP, 0)                             # QP -> Q
    assert hd, H.ref_last_error()
    assert counters(H) == (up0 + 2 * batch, down0)
    want = ob.mod_down(oQP.elementwise(2, ob.mod_up(x, fmt), k), 0)
    res = np.empty((batch, sq, n), np.uint64)
    H.ref_polys_store(hd, p(res))
    assert np.array_equal(res, want)
    assert counters(H) == (up0 + 2 * batch, down0 + batch)
    # ModReduce (t = 65537) on a resident operand in either format
    y = oracle.fill_ring(batch, n, q[:sq].copy(), 73)
    hy = H.ref_polys_load(cQ, p(y), batch, fmt)
    assert H.ref_polys_set_pim(hy, 1) == 0
    assert H.ref_polys_rescale(hy, cRest, 65537) == 0, H.ref_last_error()
    assert H.ref_polys_format(hy) == fmt
    res2 = np.empty((batch, sq - 1, n), np.uint64)
    H.ref_polys_store(hy, p(res2))
    assert np.array_equal(res2, oracle.rescale(n, q[:sq].copy(), psi[:sq].copy(), y, fmt, 65537))
    for h in (hx, hk, hd, hy):
        H.ref_polys_free(h)
    for c in (cQ, cP, cQP, cRest):
        H.ref_ctx_destroy(c)


@pytest.mark.gpu
def test_resident_bfv_evalmult_core(hooked, oracle, pkg):
    """The RNS core of BFVrns EvalMult (HPS; bfvrns-leveledshe.cpp:193-207, 322-356) written against the reference's own
    DCRTPoly — ExpandCRTBasis x4, Times / +=, SetFormat, ScaleAndRound, SwitchCRTBasis — with the operands attached:
    every step runs on the device (4 uploads and 3 downloads per ciphertext pair) and the three result polys equal the
    same code on the CPU path bit for bit."""
    H = hooked
    n, sq, sr, batch, t = 1 << 11, 3, 4, 2, 65537
    q, psi = oracle.gen_params(60, n, sq + sr)
    mk = lambda a, b: H.ref_ctx_create(n, b - a, p(q[a:b].copy()), p(psi[a:b].copy()))
    cQ, cR, cQR = mk(0, sq), mk(sq, sq + sr), mk(0, sq + sr)
    data = [oracle.fill_ring(batch, n, q[:sq].copy(), 80 + k) for k in range(4)]

    def run(attach):
        hs = [H.ref_polys_load(cQ, p(d), batch, 0) for d in data]
        if attach:
            for h in hs:
                assert H.ref_polys_set_pim(h, 1) == 0, H.ref_last_error()
        outs = (vp * 3)()
        c0 = counters(H)
        assert H.ref_bfv_hps_mult(hs[0], hs[1], hs[2], hs[3], cR, cQR, t, outs) == 0, H.ref_last_error()
        c1 = counters(H)
        res = []
        for k in range(3):
            r = np.empty((batch, sq, n), np.uint64)
            H.ref_polys_store(outs[k], p(r))
            res.append(r)
            H.ref_polys_free(outs[k])
        c2 = counters(H)
        for h in hs:
            H.ref_polys_free(h)
        return res, (c1[0] - c0[0], c1[1] - c0[1]), (c2[0] - c1[0], c2[1] - c1[1])

    launches0 = pkg._native.lib.fhe_launch_count()
    got, during, after = run(True)
    assert pkg._native.lib.fhe_launch_count() - launches0 >= 20 * batch
    assert during == (4 * batch, 0), during                 # operands go up once, nothing comes down during the product
    assert after == (0, 3 * batch), after                   # the three result polys come down when they are read
    want, cpu_during, _ = run(False)
    assert cpu_during == (0, 0)
    for k in range(3):
        assert np.array_equal(got[k], want[k]), f"c{k}"
    for c in (cQ, cR, cQR):
        H.ref_ctx_destroy(c)


@pytest.mark.gpu
def test_resident_bfv_hps_p_over_q_evalmult_core(hooked, oracle, pkg):
    """The same with the HPSPOVERQ technique (bfvrns-leveledshe.cpp:208-235, 322-334, 357-366; the default of the reference's
    BFV): ExpandCRTBasis x2, SetFormat + FastExpandCRTBasisPloverQ + SetFormat x2, Times / +=, SetFormat, ScaleAndRound to Q
    on the reference's own DCRTPoly with the operands attached — all on the device, results equal to the CPU path."""
    H = hooked
    H.ref_bfv_hps_p_over_q_mult.argtypes = [vp, vp, vp, vp, vp, vp, u64, ctypes.POINTER(vp)]
    n, sq, batch, t = 1 << 11, 3, 2, 65537
    q, psi = oracle.gen_params(60, n, 2 * sq)
    mk = lambda a, b: H.ref_ctx_create(n, b - a, p(q[a:b].copy()), p(psi[a:b].copy()))
    cQ, cR, cQR = mk(0, sq), mk(sq, 2 * sq), mk(0, 2 * sq)
    data = [oracle.fill_ring(batch, n, q[:sq].copy(), 180 + k) for k in range(4)]

    def run(attach):
        hs = [H.ref_polys_load(cQ, p(d), batch, 0) for d in data]
        if attach:
            for h in hs:
                assert H.ref_polys_set_pim(h, 1) == 0, H.ref_last_error()
        outs = (vp * 3)()
        c0 = counters(H)
        assert H.ref_bfv_hps_p_over_q_mult(hs[0], hs[1], hs[2], hs[3], cR, cQR, t, outs) == 0, H.ref_last_error()
        c1 = counters(H)
        res = []
        for k in range(3):
            r = np.empty((batch, sq, n), np.uint64)
            H.ref_polys_store(outs[k], p(r))
            res.append(r)
            H.ref_polys_free(outs[k])
        c2 = counters(H)
        for h in hs:
            H.ref_polys_free(h)
        return res, (c1[0] - c0[0], c1[1] - c0[1]), (c2[0] - c1[0], c2[1] - c1[1])

    launches0 = pkg._native.lib.fhe_launch_count()
    got, during, after = run(True)
    assert pkg._native.lib.fhe_launch_count() - launches0 >= 20 * batch
    assert during == (4 * batch, 0), during                 # operands go up once, nothing comes down during the product
    assert after == (0, 3 * batch), after
    want, cpu_during, _ = run(False)
    assert cpu_during == (0, 0)
    for k in range(3):
        assert np.array_equal(got[k], want[k]), f"c{k}"
    ob = oracle.basis(n, q, psi, sq)
    frac, modq = pkg.hps_scale_tables(q[sq:], q[:sq], t)
    orc = ob.bfv_hps_p_over_q_mult(data[0], data[1], data[2], data[3], frac, modq)
    for k in range(3):
        assert np.array_equal(got[k], orc[k]), f"oracle c{k}"
    for c in (cQ, cR, cQR):
        H.ref_ctx_destroy(c)


@pytest.mark.gpu
def test_resident_p_over_q_tails_and_reverse_expand(hooked, oracle, pkg):
    """ExpandCRTBasisReverseOrder, ExpandCRTBasisQlHat and ScaleAndRoundPOverQ (dcrtpoly-impl.h:1361-1410, 1514-1534,
    2052-2066) of the reference's own DCRTPoly with the handle attached: one upload, no download until the result is read,
    values equal to the CPU path and the oracle."""
    H = hooked
    H.ref_polys_expand_crt_basis_reverse.argtypes = [vp, vp, vp, ci]
    H.ref_polys_expand_ql_hat.argtypes = [vp, vp, vp]
    H.ref_polys_scale_round_p_over_q.argtypes = [vp, vp, vp]
    n, sq, sp, batch = 1 << 11, 3, 2, 2
    q, psi = oracle.gen_params(60, n, sq + sp)
    mk = lambda qq, pp: H.ref_ctx_create(n, len(qq), p(np.ascontiguousarray(qq, np.uint64)), p(np.ascontiguousarray(pp, np.uint64)))
    cQ, cP, cQP = mk(q[:sq], psi[:sq]), mk(q[sq:], psi[sq:]), mk(q, psi)
    cPQ = mk(np.concatenate([q[sq:], q[:sq]]), np.concatenate([psi[sq:], psi[:sq]]))
    cQ1 = mk(q[:sq + 1], psi[:sq + 1])
    x = oracle.fill_ring(batch, n, q[:sq].copy(), 301)
    z = oracle.fill_ring(batch, n, q[:sq + 1].copy(), 302)
    k = oracle.fill_ring(1, 1, q.copy(), 303).reshape(-1).copy()
    ob = oracle.basis(n, q, psi, sq)

    def run(attach, src, ctx, fmt, call, towers):
        h = H.ref_polys_load(ctx, p(src), batch, fmt)
        if attach:
            assert H.ref_polys_set_pim(h, 1) == 0, H.ref_last_error()
        c0 = counters(H)
        assert call(h) == 0, H.ref_last_error()
        c1 = counters(H)
        res = np.empty((batch, towers, n), np.uint64)
        H.ref_polys_store(h, p(res))
        H.ref_polys_free(h)
        return res, (c1[0] - c0[0], c1[1] - c0[1])

    cases = [
        ("reverse f0->r0", x, cQ, 0, lambda h: H.ref_polys_expand_crt_basis_reverse(h, cP, cPQ, 0), sq + sp, lambda: ob.expand_reverse(x, 0, 0)),
        ("reverse f1->r0", x, cQ, 1, lambda h: H.ref_polys_expand_crt_basis_reverse(h, cP, cPQ, 0), sq + sp, lambda: ob.expand_reverse(x, 1, 0)),
        ("reverse f0->r1", x, cQ, 0, lambda h: H.ref_polys_expand_crt_basis_reverse(h, cP, cPQ, 1), sq + sp, lambda: ob.expand_reverse(x, 0, 1)),
        ("ql hat", x, cQ, 0, lambda h: H.ref_polys_expand_ql_hat(h, cQP, p(k)), sq + sp, lambda: oracle.expand_ql_hat(x, n, q, sq, k[:sq])),
        ("p over q", z, cQ1, 1, lambda h: H.ref_polys_scale_round_p_over_q(h, cQ, p(k)), sq,
         lambda: oracle.scale_round_p_over_q(z, n, q[:sq].copy(), k[:sq])),
    ]
    for name, src, ctx, fmt, call, towers, want in cases:
        got, moved = run(True, src, ctx, fmt, call, towers)
        assert moved == (batch, 0), (name, moved)
        cpu, cpu_moved = run(False, src, ctx, fmt, call, towers)
        assert cpu_moved == (0, 0), name
        assert np.array_equal(got, cpu), name
        assert np.array_equal(got, want()), name + " vs oracle"
    # Times(const std::vector<NativeInteger>&): device result born on the device; constants above q_i are reduced first
    H.ref_polys_times_towers.restype = vp
    H.ref_polys_times_towers.argtypes = [vp, vp]
    big = (k[:sq] + q[:sq]).astype(np.uint64)                # unreduced, as NativeVectorT::ModMulEq accepts them
    for attach in (True, False):
        h = H.ref_polys_load(cQ, p(x), batch, 0)
        if attach:
            assert H.ref_polys_set_pim(h, 1) == 0, H.ref_last_error()
        c0 = counters(H)
        hr = H.ref_polys_times_towers(h, p(big))
        c1 = counters(H)
        assert (c1[0] - c0[0], c1[1] - c0[1]) == ((batch, 0) if attach else (0, 0))
        res = np.empty((batch, sq, n), np.uint64)
        H.ref_polys_store(hr, p(res))
        assert np.array_equal(res, oracle.expand_ql_hat(x, n, q[:sq].copy(), sq, k[:sq])), attach
        H.ref_polys_free(hr); H.ref_polys_free(h)
    for c in (cQ, cP, cQP, cPQ, cQ1):
        H.ref_ctx_destroy(c)


@pytest.mark.gpu
def test_resident_poly_serialises_through_the_reference(hooked, oracle, pkg):
    """Serial::Serialize of the reference (DCRTPolyImpl::save, dcrtpoly.h:349-354) reads the towers through the const
    accessor, so a device-resident result is downloaded once and written in the wire format; the stream equals the one
    fhe_serialize writes for the same values, and the reference's Deserialize gives a poly the hook can use again."""
    H = hooked
    n, towers = 1 << 11, 3
    q, psi = oracle.gen_params(60, n, towers)
    octx = oracle.ctx(n, q, psi)
    a, b = oracle.fill_ring(1, n, q, 51), oracle.fill_ring(1, n, q, 52)
    ctx = H.ref_ctx_create(n, towers, p(q), p(psi))
    ha, hb = H.ref_polys_load(ctx, p(a), 1, 0), H.ref_polys_load(ctx, p(b), 1, 0)
    assert H.ref_polys_set_pim(ha, 1) == 0, H.ref_last_error()
    assert H.ref_polys_op(ha, hb, 2, 1) >= 0                            # a *= b on the device
    _, down0 = counters(H)
    size = H.ref_polys_serialize(ha, 0, None, 0)
    blob = np.zeros(size, np.uint8)
    assert H.ref_polys_serialize(ha, 0, p(blob), size) == size
    assert counters(H)[1] - down0 == 1                                  # one download, the second call found the host copy current
    want = octx.elementwise(2, a, b)
    be = pkg.Backend(0)
    ring = pkg.Ring(be, n, q, psi)
    mine = pkg.DCRTPolyBatch(ring, 1, pkg.EVALUATION, want).Serialize()
    assert mine.tobytes() == blob.tobytes()
    hc = H.ref_polys_deserialize(ctx, p(blob), size)
    assert hc, H.ref_last_error()
    assert H.ref_polys_set_pim(hc, 1) == 0, H.ref_last_error()
    assert H.ref_polys_op(hc, hb, 0, 1) >= 0                            # the loaded poly goes back through the hook
    res = np.empty_like(a)
    H.ref_polys_store(hc, p(res))
    assert np.array_equal(res, octx.elementwise(0, want, b))
    ring.close(); be.close()
    for h in (ha, hb, hc):
        H.ref_polys_free(h)
    H.ref_ctx_destroy(ctx)


def _attach_child():
    """Runs in a fresh process with FHE_B200_ATTACH=1: no SetPim anywhere, yet the operators run on the device."""
    from tests import oracle_py as op
    H = load_hooked()
    O = op.load()
    up0, _ = counters(H)
    run_ops(H, O, 1 << 12, 3, 2, pim_on=False)
    # copies of an attached poly are device copies: a = polys[i]; a.SwitchFormat() as the reference's benchmark does
    n, towers = 1 << 12, 3
    q, psi = O.gen_params(60, n, towers)
    x = O.fill_ring(1, n, q, 9)
    ctx = H.ref_ctx_create(n, towers, p(q), p(psi))
    hx = H.ref_polys_load(ctx, p(x), 1, 1)
    up1, down1 = counters(H)
    for _ in range(3):
        hc = H.ref_polys_binary(hx, hx, 3 + _ % 2)    # op 3 / 4: copy construction / copy assignment of the first operand
        assert hc, H.ref_last_error()
        assert H.ref_polys_switch_format(hc, 1) >= 0
        res = np.empty_like(x)
        H.ref_polys_store(hc, p(res))
        assert np.array_equal(res, O.ctx(n, q, psi).ntt(x, inverse=False))
        H.ref_polys_free(hc)
    up2, down2 = counters(H)
    assert (up2 - up1, down2 - down1) == (1, 3), (up2 - up1, down2 - down1)   # the source went up once; each result came down once
    H.ref_polys_store(hx, p(res))
    assert np.array_equal(res, x)                     # the source itself is untouched
    assert up2 > up0
    print("attach-child ok")


@pytest.mark.gpu
def test_default_attach_runs_unmodified_callers_on_the_device():
    """FHE_B200_ATTACH=1 (pim/PimHook.h default_handle): every DCRTPoly is born with the handle, so code that never calls
    SetPim — the reference's own poly-benchmark-16k.cpp, tools/polybench.sh — runs its operators on the GPU."""
    import subprocess
    import sys
    env = dict(os.environ, FHE_B200_ATTACH="1")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, "-c", "from tests import test_live_intercept as t; t._attach_child()"], cwd=root, env=env,
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "attach-child ok" in r.stdout, r.stdout[-2000:] + r.stderr[-4000:]


@pytest.mark.gpu
def test_intercept_from_many_host_threads(hooked, oracle):
    """The reference's callers are OpenMP loops: with the handle attached, eight host threads run a chain of operators on
    their own polynomials at once, each enqueueing on its own stream (pim/PimHook.h hook_stream); a polynomial that one
    thread produced is then read by another (schedule differs between the operator loop and the store).  Results must
    equal the CPU path."""
    H = hooked
    H.ref_polys_store_mt.argtypes = [vp, vp, ci]
    n, towers, batch, threads = 1 << 13, 5, 48, 8
    q, psi = oracle.gen_params(60, n, towers)
    octx = oracle.ctx(n, q, psi)
    a, b = oracle.fill_ring(batch, n, q, 71), oracle.fill_ring(batch, n, q, 72)
    want = a
    for _ in range(3):
        want = octx.elementwise(0, want, b)
        want = octx.elementwise(2, want, b)
        want = octx.ntt(octx.ntt(want, inverse=True), inverse=False)
        want = octx.elementwise(1, want, b)
    ctx = H.ref_ctx_create(n, towers, p(q), p(psi))
    try:
        for trial in range(3):
            ha, hb = H.ref_polys_load(ctx, p(a), batch, 0), H.ref_polys_load(ctx, p(b), batch, 0)
            assert H.ref_polys_set_pim(ha, 1) == 0, H.ref_last_error()
            for _ in range(3):
                assert H.ref_polys_op(ha, hb, 0, threads) >= 0, H.ref_last_error()
                H.ref_polys_op(ha, hb, 2, threads)
                H.ref_polys_switch_format(ha, threads)
                H.ref_polys_switch_format(ha, threads)
                H.ref_polys_op(ha, hb, 1, threads)
            res = np.empty_like(a)
            H.ref_polys_store_mt(ha, p(res), 3 if trial else threads)      # another thread count: streams are handed over
            assert np.array_equal(res, want), f"trial {trial}"
            H.ref_polys_free(ha); H.ref_polys_free(hb)
    finally:
        H.ref_ctx_destroy(ctx)
