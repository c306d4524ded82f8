"""ctypes wrapper of the CPU oracle (oracle/fhe_oracle.c).  TEST INFRASTRUCTURE ONLY."""
import ctypes
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_SO = os.path.join(ROOT, "oracle", "libfhe_oracle.so")
REF_SO = os.path.join(ROOT, "oracle", "_ref", "libopenfhe_ref.so")

u64, u32, vp, sz, ci = ctypes.c_uint64, ctypes.c_uint32, ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int


def _p(a):
    return a.ctypes.data_as(vp)


class Oracle:
    def __init__(self, lib):
        L = self.L = lib
        for name, res, args in [
            ("orc_msb", u32, [u64]), ("orc_compute_mu", u64, [u64]),
            ("orc_modadd_fast", u64, [u64, u64, u64]), ("orc_modsub_fast", u64, [u64, u64, u64]),
            ("orc_modmul_barrett", u64, [u64, u64, u64, u64]), ("orc_prep_mulconst", u64, [u64, u64]),
            ("orc_modmul_const", u64, [u64, u64, u64, u64]), ("orc_modexp", u64, [u64, u64, u64]),
            ("orc_modinv", u64, [u64, u64]), ("orc_is_prime", ci, [u64]),
            ("orc_first_prime", u64, [u32, u64]), ("orc_next_prime", u64, [u64, u64]),
            ("orc_previous_prime", u64, [u64, u64]), ("orc_root_of_unity", u64, [u64, u64]),
            ("orc_gen_dcrt_params", None, [u32, u32, u32, vp, vp]),
            ("orc_vec_modadd", None, [vp, vp, vp, sz, u64]), ("orc_vec_modsub", None, [vp, vp, vp, sz, u64]),
            ("orc_vec_modmul", None, [vp, vp, vp, sz, u64]),
            ("orc_vec_modadd_scalar", None, [vp, vp, u64, sz, u64]),
            ("orc_vec_modsub_scalar", None, [vp, vp, u64, sz, u64]),
            ("orc_vec_modmul_scalar", None, [vp, vp, u64, sz, u64]),
            ("orc_precompute", None, [u64, u32, u64, vp, vp, vp, vp, vp, vp]),
            ("orc_ntt_fwd", None, [vp, u32, u64, vp, vp]),
            ("orc_ntt_inv", None, [vp, u32, u64, vp, vp, u64, u64]),
            ("orc_ctx_create", vp, [u32, u32, vp, vp]), ("orc_ctx_destroy", None, [vp]),
            ("orc_dcrt_elementwise", None, [vp, ci, vp, vp, vp, sz, ci]),
            ("orc_dcrt_ntt", None, [vp, ci, vp, sz, ci]), ("orc_max_threads", ci, []),
            ("orc_vec_modop", ci, [ci, vp, vp, sz, u64, u64, u64]),
            ("orc_automorphism", ci, [vp, vp, u32, u64, u32, ci]),
            ("orc_precompute_automap", None, [u32, u32, vp]),
            ("orc_automorphism_map", None, [vp, vp, u32, vp]),
            ("orc_dcrt_automorphism", ci, [vp, vp, vp, sz, u32, ci]),
            ("orc_switch_tables", None, [vp, u32, vp, u32, vp, vp]),
            ("orc_rescale_tables", None, [vp, u32, vp, vp]),
            ("orc_approx_switch", None, [vp, vp, sz, u32, vp, u32, vp, u32, vp, vp]),
            ("orc_approx_mod_up", None, [vp, vp, vp, vp, sz, ci]),
            ("orc_approx_mod_down", None, [vp, vp, vp, vp, sz, u64]),
            ("orc_rescale", None, [vp, vp, vp, vp, vp, sz, ci, u64]),
            ("orc_exact_tables", None, [vp, u32, vp, u32, vp, vp]),
            ("orc_switch_crt_basis", None, [vp, vp, sz, u32, vp, u32, vp, u32, vp, vp, vp, vp]),
            ("orc_expand_crt_basis", None, [vp, vp, vp, vp, sz, ci, ci]),
            ("orc_scale_and_round", None, [vp, vp, sz, u32, u32, vp, u32, ci, vp, vp]),
            ("orc_expand_crt_basis_reverse", None, [vp, vp, vp, vp, sz, ci, ci]),
            ("orc_pl_over_q_tables", None, [vp, u32, vp, u32, vp, vp]),
            ("orc_fast_expand_pl_over_q", None, [vp, vp, sz, u32, vp, u32, vp, u32, vp, u32, vp, vp]),
            ("orc_expand_ql_hat", None, [vp, vp, sz, u32, vp, u32, u32, vp]),
            ("orc_scale_round_p_over_q", None, [vp, vp, sz, u32, vp, u32, vp]),
            ("orc_uniform_at", u64, [u64, u64, u64]), ("orc_fill_uniform", None, [vp, sz, u64, u64, u64]),
            ("orc_decrypt_branch", ci, [vp, u32, u64]),
            ("orc_decrypt_tables", None, [vp, u32, u64, vp, vp, vp, vp]),
            ("orc_scale_and_round_native", None, [vp, vp, sz, u32, vp, u32, u64, vp, vp, vp, vp]),
            ("orc_behz_create", vp, [vp, u32, u64]), ("orc_behz_destroy", None, [vp]),
            ("orc_behz_table", sz, [vp, ci, vp]),
            ("orc_behz_q_to_bsk_montgomery", None, [vp, vp, vp, sz, ci]),
            ("orc_behz_fast_rns_floor_q", None, [vp, vp, sz]),
            ("orc_behz_fast_base_conv_sk", None, [vp, vp, vp, sz]),
            ("orc_behz_scale_and_round", None, [vp, vp, vp, sz]),
            ("orc_crt_decompose", None, [vp, vp, vp, sz, ci]),
            ("orc_keyswitch_digits", u32, [u32, u32]),
            ("orc_keyswitch_precompute", None, [vp, u32, u32, vp, vp, sz]),
            ("orc_keyswitch_inner", None, [vp, u32, vp, vp, vp, vp, vp, sz]),
            ("orc_keyswitch_hybrid", None, [vp, u32, u32, vp, vp, vp, vp, vp, sz, u64]),
        ]:
            f = getattr(L, name)
            f.restype, f.argtypes = res, args

    def gen_params(self, nbits, n, towers):
        q = np.zeros(towers, np.uint64)
        psi = np.zeros(towers, np.uint64)
        self.L.orc_gen_dcrt_params(nbits, n, towers, _p(q), _p(psi))
        return q, psi

    def precompute(self, psi, n, q):
        W, Wp, WI, WIp = (np.zeros(n, np.uint64) for _ in range(4))
        ni, nip = u64(), u64()
        self.L.orc_precompute(int(psi), n, int(q), _p(W), _p(Wp), _p(WI), _p(WIp), ctypes.byref(ni), ctypes.byref(nip))
        return W, Wp, WI, WIp, ni.value, nip.value

    def ntt_fwd(self, x, q, psi):
        x = np.array(x, dtype=np.uint64)
        W, Wp, _, _, _, _ = self.precompute(psi, len(x), q)
        self.L.orc_ntt_fwd(_p(x), len(x), int(q), _p(W), _p(Wp))
        return x

    def ntt_inv(self, x, q, psi):
        x = np.array(x, dtype=np.uint64)
        _, _, WI, WIp, ni, nip = self.precompute(psi, len(x), q)
        self.L.orc_ntt_inv(_p(x), len(x), int(q), _p(WI), _p(WIp), ni, nip)
        return x

    def vec(self, op, a, b, q):
        a = np.ascontiguousarray(a, np.uint64)
        b = np.ascontiguousarray(b, np.uint64)
        out = np.empty_like(a)
        f = [self.L.orc_vec_modadd, self.L.orc_vec_modsub, self.L.orc_vec_modmul][op]
        f(_p(out), _p(a), _p(b), a.size, int(q))
        return out

    def vec_scalar(self, op, a, s, q):
        a = np.ascontiguousarray(a, np.uint64)
        out = np.empty_like(a)
        f = [self.L.orc_vec_modadd_scalar, self.L.orc_vec_modsub_scalar, self.L.orc_vec_modmul_scalar][op]
        f(_p(out), _p(a), int(s), a.size, int(q))
        return out

    def modop(self, kind, x, mod, p=0, q=0):
        x = np.ascontiguousarray(x, np.uint64)
        out = np.empty_like(x)
        rc = self.L.orc_vec_modop(kind, _p(out), _p(x), x.size, int(mod), int(p), int(q))
        return rc, out

    def ctx(self, n, q, psi):
        return OracleCtx(self, n, q, psi)

    # BFV decryption: ScaleAndRound(t, ...) -> NativePoly (dcrtpoly-impl.h:1537-1816)
    def decrypt_tables(self, q, t):
        q = np.ascontiguousarray(q, np.uint64)
        T = len(q)
        modt, bmodt = np.zeros(T, np.uint64), np.zeros(T, np.uint64)
        frac, bfrac = np.zeros(T, np.float64), np.zeros(T, np.float64)
        self.L.orc_decrypt_tables(_p(q), T, int(t), _p(modt), _p(bmodt), _p(frac), _p(bfrac))
        return modt, bmodt, frac, bfrac, self.L.orc_decrypt_branch(_p(q), T, int(t))

    def scale_and_round_native(self, x, n, q, t, tables=None):
        q = np.ascontiguousarray(q, np.uint64)
        x = np.ascontiguousarray(x, np.uint64)
        modt, bmodt, frac, bfrac = (tables or self.decrypt_tables(q, t))[:4]
        B = x.size // (len(q) * n)
        out = np.empty((B, n), np.uint64)
        self.L.orc_scale_and_round_native(_p(out), _p(x), B, n, _p(q), len(q), int(t), _p(modt), _p(bmodt), _p(frac), _p(bfrac))
        return out

    # hybrid key switching (keyswitch-hybrid.cpp:324-480): qlp = OracleCtx over the sizeQl towers of Ql followed by P
    def keyswitch_precompute(self, qlp, size_ql, alpha, c):
        c = np.ascontiguousarray(c, np.uint64)
        B, T, n = c.shape[0], len(qlp.q), qlp.n
        digits = np.empty((self.L.orc_keyswitch_digits(size_ql, alpha), B, T, n), np.uint64)
        self.L.orc_keyswitch_precompute(qlp.h, size_ql, alpha, _p(digits), _p(c), B)
        return digits

    def keyswitch_inner(self, qlp, digits, evk_b, evk_a):
        digits = np.ascontiguousarray(digits, np.uint64)
        c0, c1 = np.empty(digits.shape[1:], np.uint64), np.empty(digits.shape[1:], np.uint64)
        self.L.orc_keyswitch_inner(qlp.h, digits.shape[0], _p(c0), _p(c1), _p(digits), _p(np.ascontiguousarray(evk_b, np.uint64)),
                                   _p(np.ascontiguousarray(evk_a, np.uint64)), digits.shape[1])
        return c0, c1

    def keyswitch_hybrid(self, qlp, size_ql, alpha, c, evk_b, evk_a, t=0):
        c = np.ascontiguousarray(c, np.uint64)
        B, n = c.shape[0], qlp.n
        ct0, ct1 = np.empty((B, size_ql, n), np.uint64), np.empty((B, size_ql, n), np.uint64)
        self.L.orc_keyswitch_hybrid(qlp.h, size_ql, alpha, _p(ct0), _p(ct1), _p(c), _p(np.ascontiguousarray(evk_b, np.uint64)),
                                    _p(np.ascontiguousarray(evk_a, np.uint64)), B, int(t))
        return ct0, ct1

    def fill_uniform(self, n, q, seed, first=0):
        x = np.empty(n, np.uint64)
        self.L.orc_fill_uniform(_p(x), n, int(q), int(seed), int(first))
        return x

    def fill_ring(self, batch, n, q, seed):
        """Same values as fhe_fill_uniform: flat index over [batch][tower][N]."""
        T = len(q)
        out = np.empty((batch, T, n), np.uint64)
        for b in range(batch):
            for t in range(T):
                out[b, t] = self.fill_uniform(n, q[t], seed, (b * T + t) * n)
        return out


class OracleCtx:
    def __init__(self, o, n, q, psi):
        self.o, self.n = o, n
        self.q = np.ascontiguousarray(q, np.uint64)
        self.psi = np.ascontiguousarray(psi, np.uint64)
        self.h = o.L.orc_ctx_create(n, len(self.q), _p(self.q), _p(self.psi))

    def __del__(self):
        if getattr(self, "h", None):
            self.o.L.orc_ctx_destroy(self.h)
            self.h = None

    def elementwise(self, op, a, b, threads=0):
        a = np.ascontiguousarray(a, np.uint64)
        b = np.ascontiguousarray(b, np.uint64)
        out = np.empty_like(a)
        batch = a.size // (len(self.q) * self.n)
        self.o.L.orc_dcrt_elementwise(self.h, op, _p(out), _p(a), _p(b), batch, threads)
        return out

    def ntt(self, x, inverse=False, threads=0):
        x = np.array(x, dtype=np.uint64, copy=True)
        batch = x.size // (len(self.q) * self.n)
        self.o.L.orc_dcrt_ntt(self.h, 1 if inverse else 0, _p(x), batch, threads)
        return x


def _crt_decompose(self, x, fmt=0):
    """DCRTPolyImpl::CRTDecompose(0) over [B][T][N] -> [T][B][T][N] (EVALUATION)."""
    x = np.ascontiguousarray(x, np.uint64)
    T = len(self.q)
    batch = x.size // (T * self.n)
    out = np.empty((T, batch, T, self.n), np.uint64)
    self.o.L.orc_crt_decompose(self.h, _p(out), _p(x), batch, fmt)
    return out


def _automorphism(self, x, k, fmt=0):
    """DCRTPolyImpl::AutomorphismTransform(k) over [B][T][N]; fmt 0 EVALUATION, 1 COEFFICIENT."""
    x = np.ascontiguousarray(x, np.uint64)
    out = np.empty_like(x)
    batch = x.size // (len(self.q) * self.n)
    if self.o.L.orc_dcrt_automorphism(self.h, _p(out), _p(x), batch, k, fmt):
        raise ValueError("Automorphism index not odd")
    return out


def _automap(self, n, k):
    m = np.empty(n, np.uint32)
    self.L.orc_precompute_automap(n, k, _p(m))
    return m


OracleCtx.automorphism = _automorphism
OracleCtx.crt_decompose = _crt_decompose
Oracle.automap = _automap


class OracleBehz:
    """BEHZ over QBsk = nq towers of Q followed by Bsk = B u {msk} (dcrtpoly-impl.h:2005-2050, 2069-2420)."""
    NTABLES = 17

    def __init__(self, o, n, q, psi, nq, t):
        self.o, self.n, self.nq, self.nb, self.t = o, n, nq, len(q) - nq, int(t)
        self.ctx = o.ctx(n, q, psi)
        self.h = o.L.orc_behz_create(self.ctx.h, nq, int(t))
        assert self.h, "orc_behz_create refused the shape"

    def __del__(self):
        if getattr(self, "h", None):
            self.o.L.orc_behz_destroy(self.h)
            self.h = None

    def table(self, which):
        out = np.empty(self.o.L.orc_behz_table(self.h, which, None), np.uint64)
        self.o.L.orc_behz_table(self.h, which, _p(out))
        return out

    def q_to_bsk_montgomery(self, x, fmt):
        x = np.ascontiguousarray(x, np.uint64)
        B = x.size // (self.nq * self.n)
        out = np.empty((B, self.nq + self.nb, self.n), np.uint64)
        self.o.L.orc_behz_q_to_bsk_montgomery(self.h, _p(out), _p(x), B, fmt)
        return out

    def fast_rns_floor_q(self, x):
        x = np.array(x, dtype=np.uint64, copy=True)
        B = x.size // ((self.nq + self.nb) * self.n)
        self.o.L.orc_behz_fast_rns_floor_q(self.h, _p(x), B)
        return x.reshape(B, self.nq + self.nb, self.n)

    def fast_base_conv_sk(self, x):
        x = np.ascontiguousarray(x, np.uint64)
        B = x.size // ((self.nq + self.nb) * self.n)
        out = np.empty((B, self.nq, self.n), np.uint64)
        self.o.L.orc_behz_fast_base_conv_sk(self.h, _p(out), _p(x), B)
        return out

    def scale_and_round(self, x):
        x = np.ascontiguousarray(x, np.uint64)
        B = x.size // (self.nq * self.n)
        out = np.empty((B, self.n), np.uint64)
        self.o.L.orc_behz_scale_and_round(self.h, _p(out), _p(x), B)
        return out


class OracleBasis:
    """Two-basis helper: Q = first sq towers, P = the remaining sp of (q, psi); SURVEY §8(f) rank 3."""

    def __init__(self, o, n, q, psi, sq):
        self.o, self.n, self.sq, self.sp = o, n, sq, len(q) - sq
        self.q, self.psi = np.ascontiguousarray(q, np.uint64), np.ascontiguousarray(psi, np.uint64)
        self.Q = o.ctx(n, self.q[:sq].copy(), self.psi[:sq].copy())
        self.P = o.ctx(n, self.q[sq:].copy(), self.psi[sq:].copy())

    def tables(self, reverse=False):
        a, b = (self.P, self.Q) if reverse else (self.Q, self.P)
        hi, hm = np.empty(len(a.q), np.uint64), np.empty((len(a.q), len(b.q)), np.uint64)
        self.o.L.orc_switch_tables(_p(a.q), len(a.q), _p(b.q), len(b.q), _p(hi), _p(hm))
        return hi, hm

    def approx_switch(self, x, reverse=False):
        a, b = (self.P, self.Q) if reverse else (self.Q, self.P)
        hi, hm = self.tables(reverse)
        x = np.ascontiguousarray(x, np.uint64)
        B = x.size // (len(a.q) * self.n)
        out = np.empty((B, len(b.q), self.n), np.uint64)
        self.o.L.orc_approx_switch(_p(out), _p(x), B, self.n, _p(a.q), len(a.q), _p(b.q), len(b.q), _p(hi), _p(hm))
        return out

    def mod_up(self, x, fmt):
        x = np.ascontiguousarray(x, np.uint64)
        B = x.size // (self.sq * self.n)
        out = np.empty((B, self.sq + self.sp, self.n), np.uint64)
        self.o.L.orc_approx_mod_up(self.Q.h, self.P.h, _p(out), _p(x), B, fmt)
        return out

    def mod_down(self, x, t=0):
        x = np.ascontiguousarray(x, np.uint64)
        B = x.size // ((self.sq + self.sp) * self.n)
        out = np.empty((B, self.sq, self.n), np.uint64)
        self.o.L.orc_approx_mod_down(self.Q.h, self.P.h, _p(out), _p(x), B, t)
        return out


def _exact_tables(self, reverse=False):
    a, b = (self.P, self.Q) if reverse else (self.Q, self.P)
    qinv, aq = np.empty(len(a.q), np.float64), np.empty((len(a.q) + 1, len(b.q)), np.uint64)
    self.o.L.orc_exact_tables(_p(a.q), len(a.q), _p(b.q), len(b.q), _p(qinv), _p(aq))
    return qinv, aq


def _switch_exact(self, x, reverse=False):
    a, b = (self.P, self.Q) if reverse else (self.Q, self.P)
    hi, hm = self.tables(reverse)
    qinv, aq = self.exact_tables(reverse)
    x = np.ascontiguousarray(x, np.uint64)
    B = x.size // (len(a.q) * self.n)
    out = np.empty((B, len(b.q), self.n), np.uint64)
    self.o.L.orc_switch_crt_basis(_p(out), _p(x), B, self.n, _p(a.q), len(a.q), _p(b.q), len(b.q), _p(hi), _p(hm),
                                  _p(qinv), _p(aq))
    return out


def _bfv_hps_mult(self, a0, a1, b0, b1, frac, modr):
    """RNS core of BFVrns EvalMult (HPS), bfvrns-leveledshe.cpp:193-207, 322-356: inputs over Q in EVALUATION ->
    three polynomials over Q in COEFFICIENT."""
    QP = self.o.ctx(self.n, self.q, self.psi)
    e = [self.expand(v, 0, 0) for v in (a0, a1, b0, b1)]
    c0 = QP.elementwise(2, e[0], e[2])
    c1 = QP.elementwise(0, QP.elementwise(2, e[0], e[3]), QP.elementwise(2, e[1], e[2]))
    c2 = QP.elementwise(2, e[1], e[3])
    out = []
    for c in (c0, c1, c2):
        r = self.scale_and_round(QP.ntt(c, inverse=True), frac, modr, 1)
        out.append(self.switch_exact(r, reverse=True))
    return out


def _bfv_hps_p_over_q_mult(self, a0, a1, b0, b1, frac, modq):
    """RNS core of BFVrns EvalMult (HPSPOVERQ), bfvrns-leveledshe.cpp:208-235, 322-334, 357-366: inputs over Q in
    EVALUATION -> three polynomials over Q in COEFFICIENT.  frac [sizeR], modq [sizeQ][sizeR+1]: t*Q/(QR)."""
    QP = self.o.ctx(self.n, self.q, self.psi)
    q, r = self.q[:self.sq].copy(), self.q[self.sq:].copy()
    e1 = [self.expand(v, 0, 0) for v in (a0, a1)]
    e2 = [QP.ntt(self.o.fast_expand_pl_over_q(self.Q.ntt(v, inverse=True), self.n, q, q, r)) for v in (b0, b1)]
    c0 = QP.elementwise(2, e1[0], e2[0])
    c1 = QP.elementwise(0, QP.elementwise(2, e1[0], e2[1]), QP.elementwise(2, e1[1], e2[0]))
    c2 = QP.elementwise(2, e1[1], e2[1])
    return [self.scale_and_round(QP.ntt(c, inverse=True), frac, modq, 0) for c in (c0, c1, c2)]


OracleBasis.bfv_hps_p_over_q_mult = _bfv_hps_p_over_q_mult
OracleBasis.bfv_hps_mult = _bfv_hps_mult


def _expand(self, x, fmt, result_fmt):
    x = np.ascontiguousarray(x, np.uint64)
    B = x.size // (self.sq * self.n)
    out = np.empty((B, self.sq + self.sp, self.n), np.uint64)
    self.o.L.orc_expand_crt_basis(self.Q.h, self.P.h, _p(out), _p(x), B, fmt, result_fmt)
    return out


def _scale_and_round(self, x, frac, modo, out_last=1):
    """ScaleAndRound of [B][Q+P][N] down to the P towers (out_last) or the Q towers."""
    x = np.ascontiguousarray(x, np.uint64)
    B = x.size // ((self.sq + self.sp) * self.n)
    o, sI = (self.P.q, self.sq) if out_last else (self.Q.q, self.sp)
    out = np.empty((B, len(o), self.n), np.uint64)
    frac, modo = np.ascontiguousarray(frac, np.float64), np.ascontiguousarray(modo, np.uint64)
    self.o.L.orc_scale_and_round(_p(out), _p(x), B, self.n, sI, _p(o), len(o), out_last, _p(frac), _p(modo))
    return out


OracleBasis.exact_tables = _exact_tables
OracleBasis.switch_exact = _switch_exact
OracleBasis.expand = _expand
OracleBasis.scale_and_round = _scale_and_round


def _expand_reverse(self, x, fmt, result_fmt):
    """ExpandCRTBasisReverseOrder: [B][Q][N] -> [B][P+Q][N], the P towers first."""
    x = np.ascontiguousarray(x, np.uint64)
    B = x.size // (self.sq * self.n)
    out = np.empty((B, self.sq + self.sp, self.n), np.uint64)
    self.o.L.orc_expand_crt_basis_reverse(self.Q.h, self.P.h, _p(out), _p(x), B, fmt, result_fmt)
    return out


OracleBasis.expand_reverse = _expand_reverse


def _pl_over_q_tables(self, q, pl):
    """m_negRlQHatInvModq [sq] and m_qInvModr [sq][spl] (bfvrns-cryptoparameters.cpp:501-523)."""
    q, pl = np.ascontiguousarray(q, np.uint64), np.ascontiguousarray(pl, np.uint64)
    neg, qi = np.empty(len(q), np.uint64), np.empty((len(q), len(pl)), np.uint64)
    self.L.orc_pl_over_q_tables(_p(q), len(q), _p(pl), len(pl), _p(neg), _p(qi))
    return neg, qi


def _fast_expand_pl_over_q(self, x, n, q, ql, pl):
    """FastExpandCRTBasisPloverQ: [B][len(q)][N] -> [B][len(ql)+len(pl)][N]."""
    q, ql, pl = (np.ascontiguousarray(v, np.uint64) for v in (q, ql, pl))
    neg, qi = self.pl_over_q_tables(q, pl)
    x = np.ascontiguousarray(x, np.uint64)
    B = x.size // (len(q) * n)
    out = np.empty((B, len(ql) + len(pl), n), np.uint64)
    self.L.orc_fast_expand_pl_over_q(_p(out), _p(x), B, n, _p(q), len(q), _p(ql), len(ql), _p(pl), len(pl), _p(neg), _p(qi))
    return out


def _expand_ql_hat(self, x, n, q, sl, c):
    q, c = np.ascontiguousarray(q, np.uint64), np.ascontiguousarray(c, np.uint64)
    x = np.ascontiguousarray(x, np.uint64)
    B = x.size // (sl * n)
    out = np.empty((B, len(q), n), np.uint64)
    self.L.orc_expand_ql_hat(_p(out), _p(x), B, n, _p(q), len(q), sl, _p(c))
    return out


def _scale_round_p_over_q(self, x, n, q, pinv_modq):
    q, c = np.ascontiguousarray(q, np.uint64), np.ascontiguousarray(pinv_modq, np.uint64)
    x = np.ascontiguousarray(x, np.uint64)
    B = x.size // ((len(q) + 1) * n)
    out = np.empty((B, len(q), n), np.uint64)
    self.L.orc_scale_round_p_over_q(_p(out), _p(x), B, n, _p(q), len(q), _p(c))
    return out


Oracle.pl_over_q_tables = _pl_over_q_tables
Oracle.fast_expand_pl_over_q = _fast_expand_pl_over_q
Oracle.expand_ql_hat = _expand_ql_hat
Oracle.scale_round_p_over_q = _scale_round_p_over_q


def _rescale(self, n, q, psi, x, fmt, t=0):
    """DropLastElementAndScale (t = 0) / ModReduce (t > 0) over all len(q) towers."""
    q, psi = np.ascontiguousarray(q, np.uint64), np.ascontiguousarray(psi, np.uint64)
    full, rest, last = self.ctx(n, q, psi), self.ctx(n, q[:-1].copy(), psi[:-1].copy()), self.ctx(n, q[-1:].copy(), psi[-1:].copy())
    x = np.ascontiguousarray(x, np.uint64)
    B = x.size // (len(q) * n)
    out = np.empty((B, len(q) - 1, n), np.uint64)
    self.L.orc_rescale(full.h, rest.h, last.h, _p(out), _p(x), B, fmt, t)
    return out


def _rescale_tables(self, q):
    q = np.ascontiguousarray(q, np.uint64)
    a, b = np.empty(len(q) - 1, np.uint64), np.empty(len(q) - 1, np.uint64)
    self.L.orc_rescale_tables(_p(q), len(q), _p(a), _p(b))
    return a, b


Oracle.rescale = _rescale
Oracle.rescale_tables = _rescale_tables
Oracle.basis = lambda self, n, q, psi, sq: OracleBasis(self, n, q, psi, sq)


def load():
    if not os.path.exists(ORACLE_SO):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle"), "oracle"])
    return Oracle(ctypes.CDLL(ORACLE_SO))
Oracle.behz = lambda self, n, q, psi, nq, t: OracleBehz(self, n, q, psi, nq, t)
