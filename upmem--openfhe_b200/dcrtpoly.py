"""Host-side handles over the C ABI, shaped like the reference's objects on this path.

`DCRTPolyBatch` is a device-resident batch of DCRTPoly (layout [batch][tower][N] uint64) with the
operator surface and error behaviour of the reference's DCRTPolyImpl for add / sub / mul /
SwitchFormat (lattice/hal/default/dcrtpoly.h:142-200, dcrtpoly-impl.h:398-416, 2518-2524):
`Times` throws unless both operands are in EVALUATION format, `Plus`/`Minus` throw on format or
tower-count mismatch.  Every operation is one CUDA launch over all towers and batch items.
numpy is used only to carry host data to and from the C ABI.
"""
import ctypes
import numpy as np

from . import _native as N

EVALUATION = 0      # lbcrypto::Format (utils/inttypes.h)
COEFFICIENT = 1


def _ptr(a):
    return a.ctypes.data_as(ctypes.c_void_p)


def gen_params(nbits, n, towers):
    """q_i, psi_i as benchmark/src/poly-benchmark-16k.cpp:82-101 generates them."""
    q = np.zeros(towers, np.uint64)
    psi = np.zeros(towers, np.uint64)
    N.check(N.lib.fhe_gen_params(nbits, n, towers, _ptr(q), _ptr(psi)))
    return q, psi


class Backend:
    """One B200: stream pool + device arena (what PimManager::getPim hands out in the reference)."""

    def __init__(self, device=0, opts=""):
        h = ctypes.c_void_p()
        N.check(N.lib.fhe_init(device, opts.encode() if opts else None, ctypes.byref(h)))
        self.h = h
        self.device = device

    def close(self):
        if self.h:
            N.lib.fhe_shutdown(self.h)
            self.h = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    @property
    def num_streams(self):
        return N.lib.fhe_num_streams(self.h)

    def alloc(self, nbytes):
        p = ctypes.c_uint64()
        N.check(N.lib.fhe_alloc(self.h, nbytes, ctypes.byref(p)))
        return p.value

    def free(self, p):
        N.check(N.lib.fhe_free(self.h, p))

    def host_alloc(self, nwords):
        """Pinned host buffer viewed as a numpy uint64 array."""
        p = ctypes.c_void_p()
        N.check(N.lib.fhe_host_alloc(self.h, nwords * 8, ctypes.byref(p)))
        buf = (ctypes.c_uint64 * nwords).from_address(p.value)
        arr = np.frombuffer(buf, dtype=np.uint64)
        arr.flags.writeable = True
        return arr, p.value

    def host_free(self, p):
        N.check(N.lib.fhe_host_free(self.h, ctypes.c_void_p(p)))

    def h2d(self, dptr, arr, stream=0):
        arr = np.ascontiguousarray(arr, dtype=np.uint64)
        N.check(N.lib.fhe_h2d(self.h, dptr, _ptr(arr), arr.nbytes, stream))
        self.sync(stream)               # the host array only has to outlive this call
        return arr

    def d2d(self, dst, src, nbytes, stream=0):
        N.check(N.lib.fhe_d2d(self.h, dst, src, nbytes, stream))

    def d2h(self, dptr, nwords, stream=0, out=None):
        if out is None:
            out = np.empty(nwords, np.uint64)
        N.check(N.lib.fhe_d2h(self.h, _ptr(out), dptr, nwords * 8, stream))
        self.sync(stream)
        return out

    def sync(self, stream=-2):
        N.check(N.lib.fhe_sync(self.h, stream))

    def timer_start(self, stream=0):
        N.check(N.lib.fhe_timer_start(self.h, stream))

    def timer_stop(self, stream=0):
        ms = ctypes.c_float()
        N.check(N.lib.fhe_timer_stop(self.h, stream, ctypes.byref(ms)))
        return ms.value

    def vec_op(self, op, dst, a, b, n, q, stream=0):
        N.check(N.lib.fhe_vec_op(self.h, op, dst, a, b, n, int(q), stream))

    def vec_op_scalar(self, op, dst, a, s, n, q, stream=0):
        N.check(N.lib.fhe_vec_op_scalar(self.h, op, dst, a, int(s), n, int(q), stream))

    def vec_modop(self, kind, dst, src, n, mod, p=0, q=0, stream=0):
        """kind: enum mod_kernel of pim/common.h (0 SWITCHMODULUS .. 12 DIVIDEANDROUNDEQ)."""
        N.check(N.lib.fhe_vec_modop(self.h, kind, dst, src, n, int(mod), int(p), int(q), stream))

    def checksum(self, dptr, nwords, stream=0):
        """(sum mod 2^64, xor) of a device buffer; equals sharding.checksum of the same words."""
        out = (ctypes.c_uint64 * 2)()
        N.check(N.lib.fhe_checksum(self.h, dptr, nwords, out, stream))
        return int(out[0]), int(out[1])

    def imad_probe(self, blocks, threads, iters):
        ms = ctypes.c_float()
        N.check(N.lib.fhe_imad_probe(self.h, blocks, threads, iters, ctypes.byref(ms)))
        return ms.value


class Ring:
    """ILDCRTParams on the device: N, moduli, roots, Barrett constants and twiddle tables in HBM."""

    def __init__(self, backend, n, q, psi=None):
        self.be = backend
        self.n = int(n)
        self.q = np.ascontiguousarray(q, dtype=np.uint64)
        self.psi = None if psi is None else np.ascontiguousarray(psi, dtype=np.uint64)
        self.towers = len(self.q)
        h = ctypes.c_void_p()
        N.check(N.lib.fhe_ring_create(backend.h, self.n, self.towers, _ptr(self.q),
                                      None if self.psi is None else _ptr(self.psi), ctypes.byref(h)))
        self.h = h

    def close(self):
        if self.h:
            N.lib.fhe_ring_destroy(self.h)
            self.h = None

    def words(self, batch):
        return batch * self.towers * self.n

    def table(self, tower, which):
        out = np.empty(self.n, np.uint64)
        N.check(N.lib.fhe_ring_table(self.h, tower, which, _ptr(out)))
        return out

    # raw device-pointer forms (what bench.py times)
    def op(self, op, dst, a, b, batch, stream=0):
        N.check(N.lib.fhe_dcrt_op(self.h, op, dst, a, b, batch, stream))

    def ntt(self, x, batch, inverse=False, stream=0, algo=0):
        N.check(N.lib.fhe_ntt_algo(self.h, x, batch, 1 if inverse else 0, algo, stream))

    def poly_mul(self, c, a, b, batch, stream=0):
        N.check(N.lib.fhe_poly_mul(self.h, c, a, b, batch, stream))

    def ct_multiply(self, c, a, b, batch, transform=0, stream=0):
        """c (3 device pointers) = tensor product of ciphertext batches a, b (2 pointers each);
        transform as fhe_ct_multiply: 0 none, 1 NTT in / INTT out, 2 INTT in / NTT out."""
        arr = lambda ps: (ctypes.c_uint64 * len(ps))(*ps)
        N.check(N.lib.fhe_ct_multiply(self.h, arr(c), arr(a), arr(b), batch, transform, stream))

    def fill_uniform(self, dst, batch, seed, stream=0):
        N.check(N.lib.fhe_fill_uniform(self.h, dst, batch, seed, stream))

    # host-operand forms (end-to-end path)
    def op_host(self, op, dst, a, b, batch):
        N.check(N.lib.fhe_dcrt_op_host(self.h, op, _ptr(dst), _ptr(a), _ptr(b), batch))

    def multi_op_host(self, ops, dsts, a, b, batch):
        """dsts[i] = a (ops[i]) b with ONE upload of a and b (PimData residency pattern)."""
        n = len(ops)
        c_ops = (ctypes.c_int * n)(*ops)
        c_dsts = (ctypes.c_void_p * n)(*[d.ctypes.data for d in dsts])
        N.check(N.lib.fhe_dcrt_multi_op_host(self.h, c_ops, n, c_dsts, _ptr(a), _ptr(b), batch))

    def chain_host(self, ops, a, b, batch):
        """a (ops[i])= b in sequence, in place on the host buffer `a` (the reference's += / -= / *=);
        one upload of a and b, one download of a."""
        n = len(ops)
        N.check(N.lib.fhe_dcrt_chain_host(self.h, (ctypes.c_int * n)(*ops), n, _ptr(a), _ptr(b), batch))

    def ntt_host(self, x, batch, inverse=False):
        N.check(N.lib.fhe_ntt_host(self.h, _ptr(x), batch, 1 if inverse else 0))

    # wire format (utils/serial.h:99-115, SerType::BINARY)
    def serial_bytes(self):
        """Bytes of one serialised DCRTPoly over this ring."""
        return int(N.lib.fhe_serial_bytes(self.h))

    def serialize(self, src, batch, fmt, stream=0):
        """`batch` device-resident polys -> `batch` consecutive streams, each what the reference's
        Serial::Serialize(poly, os, SerType::BINARY) writes (dcrtpoly.h:349-354)."""
        out = np.empty(self.serial_bytes() * batch, np.uint8)
        N.check(N.lib.fhe_serialize(self.h, _ptr(out), src, batch, fmt, stream))
        self.be.sync(stream)
        return out

    def deserialize(self, dst, blob, batch, stream=0):
        """Serial::Deserialize of `batch` consecutive streams into HBM; returns (format, bytes consumed).
        Raises if a stream does not belong to this ring (N, moduli, roots) or is malformed."""
        blob = np.frombuffer(blob, np.uint8) if not isinstance(blob, np.ndarray) else np.ascontiguousarray(blob, np.uint8)
        fmt, used = ctypes.c_int(), ctypes.c_size_t()
        N.check(N.lib.fhe_deserialize(self.h, dst, _ptr(blob), blob.size, batch, ctypes.byref(fmt), ctypes.byref(used), stream))
        self.be.sync(stream)
        return fmt.value, used.value


class DCRTPolyBatch:
    """`batch` DCRTPolys resident in HBM; operators mirror DCRTPolyImpl."""

    def __init__(self, ring, batch, fmt=EVALUATION, data=None, stream=0):
        self.ring, self.batch, self.format, self.stream = ring, int(batch), fmt, stream
        self.nwords = ring.words(self.batch)
        self.dptr = ring.be.alloc(max(self.nwords, 1) * 8)
        if data is not None:
            data = np.ascontiguousarray(data, dtype=np.uint64).reshape(-1)
            if data.size != self.nwords:
                raise ValueError("data must hold batch*towers*N words")
            ring.be.h2d(self.dptr, data, stream)

    def __del__(self):
        try:
            if self.dptr and self.ring.be.h:
                self.ring.be.free(self.dptr)
        except Exception:
            pass
        self.dptr = 0

    def to_host(self):
        out = self.ring.be.d2h(self.dptr, self.nwords, self.stream)
        return out.reshape(self.batch, self.ring.towers, self.ring.n)

    def Serialize(self):
        """The batch as consecutive binary streams of the reference's wire format (one per poly)."""
        return self.ring.serialize(self.dptr, self.batch, self.format, self.stream)

    @classmethod
    def Deserialize(cls, ring, blob, batch, stream=0):
        """A device-resident batch from `batch` consecutive streams (the format comes from the streams)."""
        p = cls(ring, batch, stream=stream)
        p.format, _ = ring.deserialize(p.dptr, blob, batch, stream)
        return p

    def clone(self):
        c = DCRTPolyBatch(self.ring, self.batch, self.format, stream=self.stream)
        N.check(N.lib.fhe_d2d(self.ring.be.h, c.dptr, self.dptr, self.nwords * 8, self.stream))
        return c

    # -- checks as in dcrtpoly.h:164-173,186-194 ------------------------------------------------
    def _check(self, rhs, what, need_eval=False):
        if need_eval and (self.format != EVALUATION or rhs.format != EVALUATION):
            raise RuntimeError(f"{what}: operator* for DCRTPolyImpl supported only in Format::EVALUATION")
        if self.ring.n != rhs.ring.n:
            raise RuntimeError(f"{what}: ring dimension mismatch")
        if self.format != rhs.format:
            raise RuntimeError(f"{what}: format mismatch")
        if self.ring.towers != rhs.ring.towers:
            raise RuntimeError(f"{what}: tower size mismatch")
        if self.ring.q[0] != rhs.ring.q[0]:
            raise RuntimeError(f"{what}: modulus mismatch")
        if self.batch != rhs.batch:
            raise RuntimeError(f"{what}: batch mismatch")

    def _binary(self, rhs, op, what, need_eval=False):
        self._check(rhs, what, need_eval)
        out = DCRTPolyBatch(self.ring, self.batch, self.format, stream=self.stream)
        self.ring.op(op, out.dptr, self.dptr, rhs.dptr, self.batch, self.stream)
        return out

    def _inplace(self, rhs, op):
        # operator+= / -= / *= do no checks in the reference (dcrtpoly-impl.h:410-416, dcrtpoly.h:142-148)
        self.ring.op(op, self.dptr, self.dptr, rhs.dptr, self.batch, self.stream)
        return self

    def Plus(self, rhs):
        return self._binary(rhs, N.OP_ADD, "DCRTPolyImpl::Plus")

    def Minus(self, rhs):
        return self._binary(rhs, N.OP_SUB, "DCRTPolyImpl::Minus")

    def Times(self, rhs):
        return self._binary(rhs, N.OP_MUL, "DCRTPolyImpl::Times", need_eval=True)

    __add__, __sub__, __mul__ = Plus, Minus, Times

    def __iadd__(self, rhs):
        return self._inplace(rhs, N.OP_ADD)

    def __isub__(self, rhs):
        return self._inplace(rhs, N.OP_SUB)

    def __imul__(self, rhs):
        return self._inplace(rhs, N.OP_MUL)

    def AutomorphismTransform(self, k, precomp=None):
        """dcrtpoly-impl.h:348-369.  precomp: device address of the PrecomputeAutoMap table (EVALUATION only,
        poly-impl.h:368-380); without it the index map is computed in the kernel."""
        if k % 2 == 0:
            raise RuntimeError("Automorphism index not odd")
        out = DCRTPolyBatch(self.ring, self.batch, self.format, stream=self.stream)
        if precomp is not None:
            if self.format != EVALUATION:
                raise RuntimeError("Automorphism Poly Format not EVALUATION or not power-of-two")
            N.check(N.lib.fhe_automorphism_map(self.ring.h, out.dptr, self.dptr, self.batch, precomp, self.stream))
        else:
            N.check(N.lib.fhe_automorphism(self.ring.h, out.dptr, self.dptr, self.batch, k, self.format, self.stream))
        return out

    def SwitchFormat(self):
        """dcrtpoly-impl.h:2518-2524: COEFFICIENT -> forward NTT, EVALUATION -> inverse."""
        if self.format == COEFFICIENT:
            self.ring.ntt(self.dptr, self.batch, inverse=False, stream=self.stream)
            self.format = EVALUATION
        else:
            self.ring.ntt(self.dptr, self.batch, inverse=True, stream=self.stream)
            self.format = COEFFICIENT

    def SetFormat(self, fmt):
        if self.format != fmt:
            self.SwitchFormat()


# ---- RNS basis conversion and scaling -----------------------------------------
class RingView(Ring):
    """Towers [first, first+count) of a parent Ring, sharing its tables (fhe_ring_view)."""

    def __init__(self, parent, first, count):
        self.be, self.n, self.parent = parent.be, parent.n, parent
        self.q = parent.q[first:first + count].copy()
        self.psi = None if parent.psi is None else parent.psi[first:first + count].copy()
        self.towers = count
        h = ctypes.c_void_p()
        N.check(N.lib.fhe_ring_view(parent.h, first, count, ctypes.byref(h)))
        self.h = h


class Basis:
    """Tables of DCRTPolyImpl::ApproxSwitchCRTBasis from ring `src` to ring `dst` (dcrtpoly-impl.h:1034-1082) and
    the two composites built on it.  Operands are raw device addresses, layouts [batch][towers][N]."""

    def __init__(self, src, dst, qhat_inv_modq=None, qhat_modp=None):
        self.src, self.dst = src, dst
        h = ctypes.c_void_p()
        a = None if qhat_inv_modq is None else _ptr(np.ascontiguousarray(qhat_inv_modq, np.uint64))
        b = None if qhat_modp is None else _ptr(np.ascontiguousarray(qhat_modp, np.uint64))
        N.check(N.lib.fhe_basis_create(src.h, dst.h, a, b, ctypes.byref(h)))
        self.h = h

    def close(self):
        if self.h:
            N.lib.fhe_basis_destroy(self.h)
            self.h = None

    def tables(self):
        hi = np.empty(self.src.towers, np.uint64)
        hm = np.empty((self.src.towers, self.dst.towers), np.uint64)
        N.check(N.lib.fhe_basis_tables(self.h, _ptr(hi), _ptr(hm)))
        return hi, hm

    def approx_switch(self, dst, src, batch, stream=0):
        N.check(N.lib.fhe_approx_switch_crt_basis(self.h, dst, src, batch, stream))

    def switch_crt_basis(self, dst, src, batch, stream=0):
        """DCRTPolyImpl::SwitchCRTBasis (dcrtpoly-impl.h:1178-1230), the exact form."""
        N.check(N.lib.fhe_switch_crt_basis(self.h, dst, src, batch, stream))

    def expand_crt_basis(self, dst_qp, src_q, batch, fmt, result_fmt, stream=0):
        """self = Q -> P.  DCRTPolyImpl::ExpandCRTBasis (dcrtpoly-impl.h:1311-1358)."""
        N.check(N.lib.fhe_expand_crt_basis(self.h, dst_qp, src_q, batch, fmt, result_fmt, stream))

    def expand_crt_basis_reverse_order(self, dst_pq, src_q, batch, fmt, result_fmt, stream=0):
        """self = Q -> P.  DCRTPolyImpl::ExpandCRTBasisReverseOrder (dcrtpoly-impl.h:1361-1410): P towers first."""
        N.check(N.lib.fhe_expand_crt_basis_reverse_order(self.h, dst_pq, src_q, batch, fmt, result_fmt, stream))

    def fast_expand_crt_basis_pl_over_q(self, pl_to_ql, dst_qlpl, src_q, batch, stream=0):
        """self = Q -> Pl built with the scheme's tables m_negRlQHatInvModq / m_qInvModr, pl_to_ql the exact switch back.
        DCRTPolyImpl::FastExpandCRTBasisPloverQ (dcrtpoly-impl.h:1413-1468)."""
        N.check(N.lib.fhe_fast_expand_crt_basis_pl_over_q(self.h, pl_to_ql.h, dst_qlpl, src_q, batch, stream))

    def approx_mod_up(self, dst_qp, src_q, batch, fmt, stream=0):
        """self = Q -> P.  DCRTPolyImpl::ApproxModUp (dcrtpoly-impl.h:1085-1133)."""
        N.check(N.lib.fhe_approx_mod_up(self.h, dst_qp, src_q, batch, fmt, stream))

    def approx_mod_down(self, dst_q, src_qp, batch, t=0, stream=0):
        """self = P -> Q.  DCRTPolyImpl::ApproxModDown (dcrtpoly-impl.h:1136-1175)."""
        N.check(N.lib.fhe_approx_mod_down(self.h, dst_q, src_qp, batch, t, stream))


class TowerScale:
    """One constant per tower of `ring`: DCRTPolyImpl::ExpandCRTBasisQlHat (dcrtpoly-impl.h:1514-1534, c = m_QlHatModq[l])
    and DCRTPolyImpl::ScaleAndRoundPOverQ (dcrtpoly-impl.h:2052-2066, c = pInvModq)."""

    def __init__(self, ring, c):
        self.ring = ring
        c = np.ascontiguousarray(c, np.uint64)
        if c.size != ring.towers:
            raise ValueError("one constant per tower")
        h = ctypes.c_void_p()
        N.check(N.lib.fhe_tower_scale_create(ring.h, c.ctypes.data, ctypes.byref(h)))
        self.h = h

    def close(self):
        if self.h:
            N.lib.fhe_tower_scale_destroy(self.h)
            self.h = None

    def expand_crt_basis_ql_hat(self, dst_q, src_ql, size_ql, batch, stream=0):
        N.check(N.lib.fhe_expand_crt_basis_ql_hat(self.h, size_ql, dst_q, src_ql, batch, stream))

    def scale_and_round_p_over_q(self, dst_q, src_qp, batch, stream=0):
        N.check(N.lib.fhe_scale_and_round_p_over_q(self.h, dst_q, src_qp, batch, stream))

    def times(self, dst, src, batch, stream=0):
        """DCRTPolyImpl::Times(const std::vector<NativeInteger>&) (dcrtpoly-impl.h:617-627); dst may be src."""
        N.check(N.lib.fhe_dcrt_times_towers(self.h, dst, src, batch, stream))


class ScaleRound:
    """DCRTPolyImpl::ScaleAndRound(paramsOutput, ...) (dcrtpoly-impl.h:1876-1960): `ring` is the input basis, the output
    moduli are its last (out_last) or first size_out towers; frac / modo are the scheme's tables."""

    def __init__(self, ring, size_out, frac, modo, out_last=True):
        self.ring, self.size_out = ring, size_out
        frac, modo = np.ascontiguousarray(frac, np.float64), np.ascontiguousarray(modo, np.uint64)
        if frac.size != ring.towers - size_out or modo.size != size_out * (ring.towers - size_out + 1):
            raise ValueError("frac must hold sI values and modo size_out x (sI + 1)")
        h = ctypes.c_void_p()
        N.check(N.lib.fhe_scale_round_create(ring.h, size_out, 1 if out_last else 0, frac.ctypes.data, modo.ctypes.data,
                                             ctypes.byref(h)))
        self.h = h

    def close(self):
        if self.h:
            N.lib.fhe_scale_round_destroy(self.h)
            self.h = None

    def scale_and_round(self, dst, src, batch, stream=0):
        N.check(N.lib.fhe_scale_and_round(self.h, dst, src, batch, stream))


class Rescale:
    """DCRTPolyImpl::DropLastElementAndScale / ModReduce (dcrtpoly-impl.h:746-812) for a ring's last tower."""

    def __init__(self, ring, qlql=None, qlinv=None):
        self.ring = ring
        h = ctypes.c_void_p()
        a = None if qlql is None else _ptr(np.ascontiguousarray(qlql, np.uint64))
        b = None if qlinv is None else _ptr(np.ascontiguousarray(qlinv, np.uint64))
        N.check(N.lib.fhe_rescale_create(ring.h, a, b, ctypes.byref(h)))
        self.h = h

    def close(self):
        if self.h:
            N.lib.fhe_rescale_destroy(self.h)
            self.h = None

    def tables(self):
        a, b = np.empty(self.ring.towers - 1, np.uint64), np.empty(self.ring.towers - 1, np.uint64)
        N.check(N.lib.fhe_rescale_tables(self.h, _ptr(a), _ptr(b)))
        return a, b

    def drop_last_element_and_scale(self, dst, src, batch, fmt, stream=0):
        N.check(N.lib.fhe_drop_last_element_and_scale(self.h, dst, src, batch, fmt, stream))

    def mod_reduce(self, dst, src, batch, t, fmt, stream=0):
        N.check(N.lib.fhe_mod_reduce(self.h, dst, src, batch, t, fmt, stream))


class BfvHpsMult:
    """The RNS core of BFVrns EvalMult with the HPS technique (src/pke/lib/scheme/bfvrns/bfvrns-leveledshe.cpp:193-207,
    322-356) composed from the device calls: ExpandCRTBasis of the four input polynomials Q -> QR, the fused 2 x 2 tensor
    product over QR, inverse NTT, ScaleAndRound QR -> R, SwitchCRTBasis R -> Q.  `ring_qr` holds the Q towers followed by
    the R towers; frac / modr are the scheme's t*R/(Q*R) tables (bfvrns-cryptoparameters.cpp:278-311).
    Inputs: a = (a0, a1), b = (b0, b1) device addresses over Q in EVALUATION; outputs c = (c0, c1, c2) over Q in COEFFICIENT."""

    def __init__(self, ring_qr, size_q, frac, modr):
        self.ring = ring_qr
        self.sq, self.sr = size_q, ring_qr.towers - size_q
        self.rq, self.rr = RingView(ring_qr, 0, size_q), RingView(ring_qr, size_q, self.sr)
        self.q2r, self.r2q = Basis(self.rq, self.rr), Basis(self.rr, self.rq)
        self.scale = ScaleRound(ring_qr, self.sr, frac, modr, out_last=True)
        self.batch, self.bufs = 0, []

    def _reserve(self, batch):
        if batch <= self.batch:
            return
        be = self.ring.be
        for p in self.bufs:
            be.free(p)
        wqr = self.ring.words(batch) * 8
        self.bufs = [be.alloc(wqr) for _ in range(7)] + [be.alloc(self.rr.words(batch) * 8)]
        self.batch = batch

    def __call__(self, c, a, b, batch, stream=0):
        self._reserve(batch)
        e, t, r = self.bufs[0:4], self.bufs[4:7], self.bufs[7]
        for dst, src in zip(e, (a[0], a[1], b[0], b[1])):
            self.q2r.expand_crt_basis(dst, src, batch, EVALUATION, EVALUATION, stream)
        self.ring.ct_multiply(t, e[0:2], e[2:4], batch, 0, stream)
        for k in range(3):
            self.ring.ntt(t[k], batch, inverse=True, stream=stream)
            self.scale.scale_and_round(r, t[k], batch, stream)
            self.r2q.switch_crt_basis(c[k], r, batch, stream)

    def close(self):
        for p in self.bufs:
            self.ring.be.free(p)
        self.bufs, self.batch = [], 0
        for o in (self.scale, self.q2r, self.r2q, self.rq, self.rr):
            o.close()


class BfvHpsPOverQMult:
    """The RNS core of BFVrns EvalMult with the HPSPOVERQ technique (bfvrns-leveledshe.cpp:208-235, 322-334, 357-366)
    composed from the device calls: a -> ExpandCRTBasis Q -> QR; b -> inverse NTT, FastExpandCRTBasisPloverQ (b*R/Q over
    [Q | R]), forward NTT; the fused 2 x 2 tensor product over QR; inverse NTT; ScaleAndRound QR -> Q.  `ring_qr` holds the
    Q towers followed by as many R towers; neg / qinv_modr are m_negRlQHatInvModq / m_qInvModr
    (bfvrns-cryptoparameters.cpp:501-523), frac / modq the t*Q/(QR) tables (:531-557) = hps_scale_tables(r, q, t).
    Inputs: a = (a0, a1), b = (b0, b1) device addresses over Q in EVALUATION; outputs c = (c0, c1, c2) over Q in COEFFICIENT."""

    def __init__(self, ring_qr, size_q, neg, qinv_modr, frac, modq):
        if ring_qr.towers != 2 * size_q:
            raise ValueError("HPSPOVERQ needs as many R towers as Q towers")
        self.ring, self.sq = ring_qr, size_q
        self.rq, self.rr = RingView(ring_qr, 0, size_q), RingView(ring_qr, size_q, size_q)
        self.q2r, self.r2q = Basis(self.rq, self.rr), Basis(self.rr, self.rq)
        self.q2r_over = Basis(self.rq, self.rr, neg, qinv_modr)
        self.scale = ScaleRound(ring_qr, size_q, frac, modq, out_last=False)
        self.batch, self.bufs = 0, []

    def _reserve(self, batch):
        if batch <= self.batch:
            return
        be = self.ring.be
        for p in self.bufs:
            be.free(p)
        wqr = self.ring.words(batch) * 8
        self.bufs = [be.alloc(wqr) for _ in range(7)] + [be.alloc(self.rq.words(batch) * 8)]
        self.batch = batch

    def __call__(self, c, a, b, batch, stream=0):
        self._reserve(batch)
        e, t, coef = self.bufs[0:4], self.bufs[4:7], self.bufs[7]
        be = self.ring.be
        for dst, src in zip(e[0:2], a):
            self.q2r.expand_crt_basis(dst, src, batch, EVALUATION, EVALUATION, stream)
        for dst, src in zip(e[2:4], b):
            be.d2d(coef, src, self.rq.words(batch) * 8, stream)
            self.rq.ntt(coef, batch, inverse=True, stream=stream)
            self.q2r_over.fast_expand_crt_basis_pl_over_q(self.r2q, dst, coef, batch, stream)
            self.ring.ntt(dst, batch, inverse=False, stream=stream)
        self.ring.ct_multiply(t, e[0:2], e[2:4], batch, 0, stream)
        for k in range(3):
            self.ring.ntt(t[k], batch, inverse=True, stream=stream)
            self.scale.scale_and_round(c[k], t[k], batch, stream)

    def close(self):
        for p in self.bufs:
            self.ring.be.free(p)
        self.bufs, self.batch = [], 0
        for o in (self.scale, self.q2r, self.r2q, self.q2r_over, self.rq, self.rr):
            o.close()


def hps_p_over_q_tables(q, r):
    """m_negRlQHatInvModq [sizeQ] = q_i - [R * (Q/q_i)^-1]_{q_i} and m_qInvModr [sizeQ][sizeR] = [q_i^-1]_{r_j}, as
    src/pke/lib/scheme/bfvrns/bfvrns-cryptoparameters.cpp:501-523 computes them (BigInteger there, Python integers here)."""
    q, r = [int(v) for v in q], [int(v) for v in r]
    Q, R = 1, 1
    for v in q:
        Q *= v
    for v in r:
        R *= v
    neg = np.array([qi - (R * pow(Q // qi, -1, qi)) % qi for qi in q], np.uint64)
    qinv = np.array([[pow(qi, -1, rj) for rj in r] for qi in q], np.uint64)
    return neg, qinv


def hps_scale_tables(q, r, t):
    """The ScaleAndRound tables of the HPS multiplication, t*R/(Q*R) from QR down to R, as
    src/pke/lib/scheme/bfvrns/bfvrns-cryptoparameters.cpp:278-311 computes them (BigInteger there, Python integers
    here): frac[i] = double([(QR/q_i)^-1 * R * t]_{q_i}) / double(q_i);
    modr[j][i] = floor(t * R * [(QR/q_i)^-1]_{q_i} / q_i) mod r_j, modr[j][sizeQ] the same with r_j in place of q_i."""
    q, r = [int(v) for v in q], [int(v) for v in r]
    R = 1
    for v in r:
        R *= v
    QR = R
    for v in q:
        QR *= v
    frac = np.array([float((pow(QR // qi, -1, qi) * R * t) % qi) / float(qi) for qi in q], np.float64)
    modr = np.empty((len(r), len(q) + 1), np.uint64)
    for j, rj in enumerate(r):
        for i, qi in enumerate(q):
            modr[j, i] = ((t * R * pow(QR // qi, -1, qi)) // qi) % rj
        modr[j, len(q)] = ((t * R * pow(QR // rj, -1, rj)) // rj) % rj
    return frac, modr


class Group:
    """Several B200s of one node driven by this process (fhe_group_*): what PimManager::getPim(nr_dpus) hands out when
    `nr_dpus` devices are asked for (src/core/include/pim/PimManager.h:23-29).  The batch is split in contiguous
    blocks, one per device; NCCL inside libfhe_b200.so carries only the optional gather."""

    def __init__(self, ndev=0, devices=None, opts=""):
        h = ctypes.c_void_p()
        devs = None if devices is None else (ctypes.c_int * len(devices))(*devices)
        N.check(N.lib.fhe_group_create(ndev, devs, opts.encode() if opts else None, ctypes.byref(h)))
        self.h = h
        self.size = N.lib.fhe_group_size(h)

    def close(self):
        if self.h:
            N.lib.fhe_group_destroy(self.h)
            self.h = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def backend(self, rank):
        """A Backend view of one device of the group (not owning: do not close it)."""
        be = Backend.__new__(Backend)
        be.h = ctypes.c_void_p(N.lib.fhe_group_ctx(self.h, rank))
        be.device = N.lib.fhe_device(be.h)
        return be

    def partition(self, batch, rank):
        first, count = ctypes.c_size_t(), ctypes.c_size_t()
        N.check(N.lib.fhe_group_partition(self.h, batch, rank, ctypes.byref(first), ctypes.byref(count)))
        return first.value, count.value

    def sync(self):
        N.check(N.lib.fhe_group_sync(self.h))


class GroupRing:
    """The same ring on every device of a Group, and the hot path over a batch split across them."""

    def __init__(self, group, n, q, psi=None):
        self.group = group
        self.n = int(n)
        self.q = np.ascontiguousarray(q, dtype=np.uint64)
        self.psi = None if psi is None else np.ascontiguousarray(psi, dtype=np.uint64)
        self.towers = len(self.q)
        h = ctypes.c_void_p()
        N.check(N.lib.fhe_group_ring_create(group.h, self.n, self.towers, _ptr(self.q),
                                            None if self.psi is None else _ptr(self.psi), ctypes.byref(h)))
        self.h = h

    def close(self):
        if self.h:
            N.lib.fhe_group_ring_destroy(self.h)
            self.h = None

    def _arr(self, ptrs):
        return (ctypes.c_uint64 * self.group.size)(*ptrs)

    def words(self, batch):
        return batch * self.towers * self.n

    def alloc(self, batch):
        out = (ctypes.c_uint64 * self.group.size)()
        N.check(N.lib.fhe_group_alloc(self.h, batch, out))
        return list(out)

    def free(self, bufs):
        N.check(N.lib.fhe_group_free(self.h, self._arr(bufs)))

    def fill_uniform(self, bufs, batch, seed, stream=0):
        """Seeded synthetic residues in every block (block r uses seed + 7919 * first_r)."""
        for r in range(self.group.size):
            first, count = self.group.partition(batch, r)
            if count:
                ring = ctypes.c_void_p(N.lib.fhe_group_ring_at(self.h, r))
                N.check(N.lib.fhe_fill_uniform(ring, bufs[r], count, seed + 7919 * first, stream))

    def scatter(self, bufs, host, batch, stream=0):
        host = np.ascontiguousarray(host, dtype=np.uint64)
        N.check(N.lib.fhe_group_scatter(self.h, self._arr(bufs), _ptr(host), batch, stream))
        self.group.sync()

    def gather(self, bufs, batch, stream=0):
        out = np.empty((batch, self.towers, self.n), np.uint64)
        N.check(N.lib.fhe_group_gather(self.h, _ptr(out), self._arr(bufs), batch, stream))
        self.group.sync()
        return out

    def op(self, op, dst, a, b, batch, stream=0):
        N.check(N.lib.fhe_group_dcrt_op(self.h, op, self._arr(dst), self._arr(a), self._arr(b), batch, stream))

    def ntt(self, x, batch, inverse=False, stream=0):
        N.check(N.lib.fhe_group_ntt(self.h, self._arr(x), batch, 1 if inverse else 0, stream))

    def all_gather(self, dst, src, batch, stream=0):
        N.check(N.lib.fhe_group_all_gather(self.h, self._arr(dst), self._arr(src), batch, stream))

    def chain_host(self, ops, a, b, batch):
        n = len(ops)
        N.check(N.lib.fhe_group_chain_host(self.h, (ctypes.c_int * n)(*ops), n, _ptr(a), None if b is None else _ptr(b), batch))


class Comm:
    """One rank of a NCCL communicator owned by libfhe_b200.so (one-process-per-GPU callers, e.g. under torchrun):
    rank 0 creates the id with Comm.unique_id() and the caller hands it to the other ranks."""

    @staticmethod
    def unique_id():
        buf = ctypes.create_string_buffer(128)
        N.check(N.lib.fhe_comm_unique_id(buf, 128))
        return buf.raw

    def __init__(self, backend, nranks, rank, uid):
        h = ctypes.c_void_p()
        N.check(N.lib.fhe_comm_create(backend.h, nranks, rank, ctypes.c_char_p(uid), ctypes.byref(h)))
        self.h, self.nranks, self.rank = h, nranks, rank

    def all_gather(self, dst, src, words, stream=0):
        """dst (nranks * words device words) receives every rank's `words` words, on this library's stream."""
        N.check(N.lib.fhe_comm_all_gather(self.h, dst, src, words, stream))

    def close(self):
        if self.h:
            N.lib.fhe_comm_destroy(self.h)
            self.h = None


class KeySwitch:
    """Hybrid key switching at one level (KeySwitchHYBRID, keyswitch-hybrid.cpp:324-480) as a device composition.
    `ring_qlp`: the size_ql current towers of Q followed by P; alpha: towers per digit."""

    def __init__(self, ring_qlp, size_ql, alpha):
        self.ring, self.size_ql, self.alpha = ring_qlp, int(size_ql), int(alpha)
        h = ctypes.c_void_p()
        N.check(N.lib.fhe_keyswitch_create(ring_qlp.h, self.size_ql, self.alpha, ctypes.byref(h)))
        self.h = h
        self.digits = N.lib.fhe_keyswitch_digits(h)

    def close(self):
        if self.h:
            N.lib.fhe_keyswitch_destroy(self.h)
            self.h = None

    def precompute(self, digits, c, batch, stream=0):
        """EvalKeySwitchPrecomputeCore: c over Ql (EVALUATION) -> digits [digits][batch][T][N]."""
        N.check(N.lib.fhe_keyswitch_precompute(self.h, digits, c, batch, stream))

    def inner(self, ctilda0, ctilda1, digits, evk_b, evk_a, batch, stream=0):
        """EvalFastKeySwitchCoreExt: the inner product with the key (given over the towers of ring_qlp)."""
        N.check(N.lib.fhe_keyswitch_inner(self.h, ctilda0, ctilda1, digits, evk_b, evk_a, batch, stream))

    def fast(self, ct0, ct1, digits, evk_b, evk_a, batch, t=0, stream=0):
        """EvalFastKeySwitchCore (keyswitch-hybrid.cpp:415-437): inner product with the key, then ApproxModDown of both halves."""
        N.check(N.lib.fhe_keyswitch_fast(self.h, ct0, ct1, digits, evk_b, evk_a, batch, int(t), stream))

    def add_to(self, acc0, acc1, c, evk_b, evk_a, batch, t=0, stream=0):
        """acc0 += ct0, acc1 += ct1 of the key switch of c (what the schemes do with KeySwitchCore's result), one composition."""
        N.check(N.lib.fhe_keyswitch_hybrid_add(self.h, acc0, acc1, c, evk_b, evk_a, batch, int(t), stream))

    def __call__(self, ct0, ct1, c, evk_b, evk_a, batch, t=0, stream=0):
        """KeySwitchCore: ct0, ct1 over Ql."""
        N.check(N.lib.fhe_keyswitch_hybrid(self.h, ct0, ct1, c, evk_b, evk_a, batch, int(t), stream))


class CrtDecompose:
    """DCRTPolyImpl::CRTDecompose(0) (dcrtpoly-impl.h:266-288): src [batch][T][N] -> digits [T][batch][T][N], EVALUATION."""

    def __init__(self, ring):
        self.ring = ring
        h = ctypes.c_void_p()
        N.check(N.lib.fhe_crt_decompose_create(ring.h, ctypes.byref(h)))
        self.h = h

    def close(self):
        if self.h:
            N.lib.fhe_crt_decompose_destroy(self.h)
            self.h = None

    def __call__(self, dst_digits, src, batch, fmt=EVALUATION, stream=0):
        N.check(N.lib.fhe_crt_decompose(self.h, dst_digits, src, batch, fmt, stream))


class Behz:
    """The BEHZ set of DCRTPolyImpl (dcrtpoly-impl.h:2005-2050, 2069-2420) over `ring_qbsk` = size_q towers of Q followed
    by Bsk = B u {msk}; every table of bfvrns-cryptoparameters.cpp:724-891 is derived on the host from the moduli and t."""
    NTABLES = 17

    def __init__(self, ring_qbsk, size_q, t):
        self.ring, self.nq, self.nb, self.t = ring_qbsk, size_q, ring_qbsk.towers - size_q, int(t)
        h = ctypes.c_void_p()
        N.check(N.lib.fhe_behz_create(ring_qbsk.h, size_q, self.t, ctypes.byref(h)))
        self.h = h

    def close(self):
        if self.h:
            N.lib.fhe_behz_destroy(self.h)
            self.h = None

    def table(self, which):
        n = ctypes.c_size_t()
        N.check(N.lib.fhe_behz_table(self.h, which, None, ctypes.byref(n)))
        out = np.zeros(n.value, np.uint64)
        N.check(N.lib.fhe_behz_table(self.h, which, _ptr(out), None))
        return out

    def q_to_bsk_montgomery(self, dst_qbsk, src_q, batch, fmt, stream=0):
        N.check(N.lib.fhe_fast_base_conv_q_to_bsk_montgomery(self.h, dst_qbsk, src_q, batch, fmt, stream))

    def fast_rns_floor_q(self, x_qbsk, batch, update_q=True, stream=0):
        N.check(N.lib.fhe_fast_rns_floor_q(self.h, x_qbsk, batch, 1 if update_q else 0, stream))

    def fast_rns_floor_q_noalloc(self, x_qbsk, batch, update_q=True, stream=0):
        """Same result with the intermediate in the library's static 8 MiB device array: no allocation (fhe_b200.h)."""
        N.check(N.lib.fhe_fast_rns_floor_q_noalloc(self.h, x_qbsk, batch, 1 if update_q else 0, stream))

    def fast_base_conv_sk(self, dst_q, src_qbsk, batch, stream=0):
        N.check(N.lib.fhe_fast_base_conv_sk(self.h, dst_q, src_qbsk, batch, stream))

    def scale_and_round(self, dst, src_q, batch, stream=0):
        N.check(N.lib.fhe_scale_and_round_behz(self.h, dst, src_q, batch, stream))


class BfvBehzMult:
    """The RNS core of BFVrns EvalMult with the BEHZ technique (src/pke/lib/scheme/bfvrns/bfvrns-leveledshe.cpp:275-297,
    322-334, 367-388) composed from the device calls: FastBaseConvqToBskMontgomery of the four input polynomials Q -> QBsk,
    the fused 2 x 2 tensor product over QBsk, inverse NTT, FastRNSFloorq, FastBaseConvSK back to Q.
    Inputs: a = (a0, a1), b = (b0, b1) device addresses over Q in EVALUATION; outputs c = (c0, c1, c2) over Q in COEFFICIENT."""

    def __init__(self, ring_qbsk, size_q, t):
        self.ring = ring_qbsk
        self.behz = Behz(ring_qbsk, size_q, t)
        self.batch, self.bufs = 0, []

    def _reserve(self, batch):
        if batch <= self.batch:
            return
        be = self.ring.be
        for p in self.bufs:
            be.free(p)
        self.bufs = [be.alloc(self.ring.words(batch) * 8) for _ in range(7)]
        self.batch = batch

    def __call__(self, c, a, b, batch, stream=0):
        self._reserve(batch)
        e, t = self.bufs[0:4], self.bufs[4:7]
        for dst, src in zip(e, (a[0], a[1], b[0], b[1])):
            self.behz.q_to_bsk_montgomery(dst, src, batch, EVALUATION, stream)
        self.ring.ct_multiply(t, e[0:2], e[2:4], batch, 0, stream)
        for k in range(3):
            self.ring.ntt(t[k], batch, inverse=True, stream=stream)
            self.behz.fast_rns_floor_q(t[k], batch, update_q=False, stream=stream)     # the towers of Q are dead from here on
            self.behz.fast_base_conv_sk(c[k], t[k], batch, stream)

    def close(self):
        for p in self.bufs:
            self.ring.be.free(p)
        self.bufs, self.batch = [], 0
        self.behz.close()


class DecryptScale:
    """BFV decryption's scaling step, DCRTPolyImpl::ScaleAndRound(t, ...) -> NativePoly (dcrtpoly-impl.h:1537-1816):
    src over `ring` (COEFFICIENT) -> [batch][N] values modulo t.  tables = (modt, b_modt, frac, b_frac) as the scheme
    holds them (bfvrns-cryptoparameters.cpp:447-494), or None to have them derived from the moduli."""

    def __init__(self, ring, t, tables=None):
        self.ring, self.t = ring, int(t)
        h = ctypes.c_void_p()
        if tables is None:
            args = (None, None, None, None)
        else:
            self._keep = [np.ascontiguousarray(tables[0], np.uint64), np.ascontiguousarray(tables[1], np.uint64),
                          np.ascontiguousarray(tables[2], np.float64), np.ascontiguousarray(tables[3], np.float64)]
            args = (_ptr(self._keep[0]), _ptr(self._keep[1]), _ptr(self._keep[2]), _ptr(self._keep[3]))
        N.check(N.lib.fhe_decrypt_scale_create(ring.h, self.t, args[0], args[1], args[2], args[3], ctypes.byref(h)))
        self.h = h

    def close(self):
        if self.h:
            N.lib.fhe_decrypt_scale_destroy(self.h)
            self.h = None

    def tables(self):
        T = self.ring.towers
        modt, bmodt = np.zeros(T, np.uint64), np.zeros(T, np.uint64)
        frac, bfrac = np.zeros(T, np.float64), np.zeros(T, np.float64)
        br = ctypes.c_int()
        N.check(N.lib.fhe_decrypt_scale_tables(self.h, _ptr(modt), _ptr(bmodt), _ptr(frac), _ptr(bfrac), ctypes.byref(br)))
        return modt, bmodt, frac, bfrac, br.value

    def __call__(self, dst, src, batch, stream=0):
        N.check(N.lib.fhe_scale_and_round_native(self.h, dst, src, batch, stream))
