"""B200 backend for the UPMEM--OpenFHE intercept surface: tower-wise modular add / sub / mul of RNS
polynomials in the evaluation domain and the forward / inverse negacyclic NTT, as hand-written
sm_100a kernels behind a C ABI (include/fhe_b200.h).

The directory name is not a Python identifier; import it with
    importlib.import_module("upmem--openfhe_b200")
Importing fails if lib/libfhe_b200.so has not been built: there is no CPU fallback.
"""
from . import _native
from ._native import FheError, LIB_PATH, OP_ADD, OP_INTT, OP_MUL, OP_NTT, OP_SUB, SIGNATURES
from .dcrtpoly import (COEFFICIENT, EVALUATION, Backend, Basis, Behz, CrtDecompose, BfvBehzMult, BfvHpsMult, BfvHpsPOverQMult, Comm, DCRTPolyBatch, DecryptScale, Group, GroupRing, KeySwitch, Rescale, Ring, RingView, ScaleRound, TowerScale,
                       gen_params, hps_p_over_q_tables, hps_scale_tables)

__all__ = ["Backend", "Group", "GroupRing", "Comm", "KeySwitch", "DecryptScale", "Ring", "RingView", "Basis", "Rescale", "ScaleRound", "TowerScale", "Behz", "CrtDecompose", "BfvBehzMult", "BfvHpsMult", "BfvHpsPOverQMult", "hps_scale_tables", "hps_p_over_q_tables", "DCRTPolyBatch", "gen_params", "FheError", "EVALUATION", "COEFFICIENT",
           "OP_ADD", "OP_SUB", "OP_MUL", "OP_NTT", "OP_INTT", "LIB_PATH", "SIGNATURES"]
