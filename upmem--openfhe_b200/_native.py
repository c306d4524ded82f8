"""ctypes binding of the C ABI declared in include/fhe_b200.h.

This is the reference-side binding a maintainer would add for a Python caller; the C++ binding is
include/pim/*.h.  There is no fallback: if the CUDA library is missing, importing fails loudly, and
if no B200 is visible every compute call fails with FHE_ENODEV.
"""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libfhe_b200.so")

u64 = ctypes.c_uint64
u32 = ctypes.c_uint32
vp = ctypes.c_void_p
sz = ctypes.c_size_t
ci = ctypes.c_int
pu64 = ctypes.POINTER(ctypes.c_uint64)

# name -> (restype, argtypes): exactly the symbols include/fhe_b200.h declares
SIGNATURES = {
    "fhe_last_error": (ctypes.c_char_p, []),
    "fhe_version": (ctypes.c_char_p, []),
    "fhe_init": (ci, [ci, ctypes.c_char_p, ctypes.POINTER(vp)]),
    "fhe_shutdown": (ci, [vp]),
    "fhe_device": (ci, [vp]),
    "fhe_num_streams": (ci, [vp]),
    "fhe_sync": (ci, [vp, ci]),
    "fhe_launch_count": (ctypes.c_long, []),
    "fhe_alloc": (ci, [vp, sz, ctypes.POINTER(u64)]),
    "fhe_free": (ci, [vp, u64]),
    "fhe_arena_stats": (ci, [vp, ctypes.POINTER(sz), ctypes.POINTER(sz)]),
    "fhe_host_alloc": (ci, [vp, sz, ctypes.POINTER(vp)]),
    "fhe_host_free": (ci, [vp, vp]),
    "fhe_h2d": (ci, [vp, u64, vp, sz, ci]),
    "fhe_d2h": (ci, [vp, vp, u64, sz, ci]),
    "fhe_d2d": (ci, [vp, u64, u64, sz, ci]),
    "fhe_timer_start": (ci, [vp, ci]),
    "fhe_timer_stop": (ci, [vp, ci, ctypes.POINTER(ctypes.c_float)]),
    "fhe_gen_params": (ci, [u32, u32, u32, vp, vp]),
    "fhe_root_of_unity": (u64, [u64, u64]),
    "fhe_ring_create": (ci, [vp, u32, u32, vp, vp, ctypes.POINTER(vp)]),
    "fhe_ring_destroy": (ci, [vp]),
    "fhe_ring_n": (u32, [vp]),
    "fhe_ring_towers": (u32, [vp]),
    "fhe_ring_table": (ci, [vp, u32, ci, vp]),
    "fhe_dcrt_op": (ci, [vp, ci, u64, u64, u64, sz, ci]),
    "fhe_dcrt_op_scalars": (ci, [vp, ci, u64, u64, vp, sz, ci]),
    "fhe_ntt_forward": (ci, [vp, u64, sz, ci]),
    "fhe_ntt_inverse": (ci, [vp, u64, sz, ci]),
    "fhe_ntt_algo": (ci, [vp, u64, sz, ci, ci, ci]),
    "fhe_poly_mul": (ci, [vp, u64, u64, u64, sz, ci]),
    "fhe_ct_tensor_product": (ci, [vp, pu64, pu64, pu64, sz, ci]),
    "fhe_ct_multiply": (ci, [vp, pu64, pu64, pu64, sz, ci, ci]),
    "fhe_automorphism": (ci, [vp, u64, u64, sz, u32, ci, ci]),
    "fhe_precompute_automap": (ci, [u32, u32, vp]),
    "fhe_automorphism_map": (ci, [vp, u64, u64, sz, u64, ci]),
    "fhe_ring_view": (ci, [vp, u32, u32, ctypes.POINTER(vp)]),
    "fhe_basis_create": (ci, [vp, vp, vp, vp, ctypes.POINTER(vp)]),
    "fhe_basis_destroy": (ci, [vp]),
    "fhe_basis_tables": (ci, [vp, vp, vp]),
    "fhe_approx_switch_crt_basis": (ci, [vp, u64, u64, sz, ci]),
    "fhe_approx_mod_up": (ci, [vp, u64, u64, sz, ci, ci]),
    "fhe_approx_mod_down": (ci, [vp, u64, u64, sz, u64, ci]),
    "fhe_switch_crt_basis": (ci, [vp, u64, u64, sz, ci]),
    "fhe_expand_crt_basis": (ci, [vp, u64, u64, sz, ci, ci, ci]),
    "fhe_expand_crt_basis_reverse_order": (ci, [vp, u64, u64, sz, ci, ci, ci]),
    "fhe_fast_expand_crt_basis_pl_over_q": (ci, [vp, vp, u64, u64, sz, ci]),
    "fhe_tower_scale_create": (ci, [vp, vp, ctypes.POINTER(vp)]),
    "fhe_tower_scale_destroy": (ci, [vp]),
    "fhe_expand_crt_basis_ql_hat": (ci, [vp, u32, u64, u64, sz, ci]),
    "fhe_scale_and_round_p_over_q": (ci, [vp, u64, u64, sz, ci]),
    "fhe_dcrt_times_towers": (ci, [vp, u64, u64, sz, ci]),
    "fhe_scale_round_create": (ci, [vp, u32, ci, vp, vp, ctypes.POINTER(vp)]),
    "fhe_scale_round_destroy": (ci, [vp]),
    "fhe_scale_and_round": (ci, [vp, u64, u64, sz, ci]),
    "fhe_rescale_create": (ci, [vp, vp, vp, ctypes.POINTER(vp)]),
    "fhe_rescale_destroy": (ci, [vp]),
    "fhe_rescale_tables": (ci, [vp, vp, vp]),
    "fhe_drop_last_element_and_scale": (ci, [vp, u64, u64, sz, ci, ci]),
    "fhe_mod_reduce": (ci, [vp, u64, u64, sz, u64, ci, ci]),
    "fhe_serial_bytes": (sz, [vp]),
    "fhe_serial_template": (sz, [u32, u32, vp, vp, ci, vp, sz, vp]),
    "fhe_serial_parse": (ci, [u32, u32, vp, vp, vp, sz, vp, ctypes.POINTER(ci), ctypes.POINTER(sz)]),
    "fhe_serialize": (ci, [vp, vp, u64, sz, ci, ci]),
    "fhe_deserialize": (ci, [vp, u64, vp, sz, sz, ctypes.POINTER(ci), ctypes.POINTER(sz), ci]),
    "fhe_fill_uniform": (ci, [vp, u64, sz, u64, ci]),
    "fhe_vec_op": (ci, [vp, ci, u64, u64, u64, sz, u64, ci]),
    "fhe_vec_op_scalar": (ci, [vp, ci, u64, u64, u64, sz, u64, ci]),
    "fhe_vec_modop": (ci, [vp, ci, u64, u64, sz, u64, u64, u64, ci]),
    "fhe_checksum": (ci, [vp, u64, sz, pu64, ci]),
    "fhe_launch": (ci, [vp, ci, vp, u64]),
    "fhe_dcrt_op_host": (ci, [vp, ci, vp, vp, vp, sz]),
    "fhe_dcrt_multi_op_host": (ci, [vp, ctypes.POINTER(ci), ci, ctypes.POINTER(vp), vp, vp, sz]),
    "fhe_dcrt_chain_host": (ci, [vp, ctypes.POINTER(ci), ci, vp, vp, sz]),
    "fhe_ntt_host": (ci, [vp, vp, sz, ci]),
    "fhe_imad_probe": (ci, [vp, ci, ci, ci, ctypes.POINTER(ctypes.c_float)]),
    "fhe_decrypt_scale_create": (ci, [vp, u64, vp, vp, vp, vp, ctypes.POINTER(vp)]),
    "fhe_decrypt_scale_destroy": (ci, [vp]),
    "fhe_decrypt_scale_tables": (ci, [vp, vp, vp, vp, vp, ctypes.POINTER(ci)]),
    "fhe_scale_and_round_native": (ci, [vp, u64, u64, sz, ci]),
    "fhe_crt_decompose_create": (ci, [vp, ctypes.POINTER(vp)]),
    "fhe_crt_decompose_destroy": (ci, [vp]),
    "fhe_crt_decompose": (ci, [vp, u64, u64, sz, ci, ci]),
    "fhe_behz_create": (ci, [vp, u32, u64, ctypes.POINTER(vp)]),
    "fhe_behz_destroy": (ci, [vp]),
    "fhe_behz_table": (ci, [vp, ci, vp, ctypes.POINTER(sz)]),
    "fhe_fast_base_conv_q_to_bsk_montgomery": (ci, [vp, u64, u64, sz, ci, ci]),
    "fhe_fast_rns_floor_q": (ci, [vp, u64, sz, ci, ci]),
    "fhe_fast_rns_floor_q_noalloc": (ci, [vp, u64, sz, ci, ci]),
    "fhe_fast_base_conv_sk": (ci, [vp, u64, u64, sz, ci]),
    "fhe_scale_and_round_behz": (ci, [vp, u64, u64, sz, ci]),
    "fhe_keyswitch_create": (ci, [vp, u32, u32, ctypes.POINTER(vp)]),
    "fhe_keyswitch_destroy": (ci, [vp]),
    "fhe_keyswitch_digits": (u32, [vp]),
    "fhe_keyswitch_precompute": (ci, [vp, u64, u64, sz, ci]),
    "fhe_keyswitch_inner": (ci, [vp, u64, u64, u64, u64, u64, sz, ci]),
    "fhe_keyswitch_hybrid": (ci, [vp, u64, u64, u64, u64, u64, sz, u64, ci]),
    "fhe_keyswitch_hybrid_add": (ci, [vp, u64, u64, u64, u64, u64, sz, u64, ci]),
    "fhe_keyswitch_fast": (ci, [vp, u64, u64, u64, u64, u64, sz, u64, ci]),
    "fhe_group_create": (ci, [ci, vp, ctypes.c_char_p, ctypes.POINTER(vp)]),
    "fhe_group_destroy": (ci, [vp]),
    "fhe_group_size": (ci, [vp]),
    "fhe_group_ctx": (vp, [vp, ci]),
    "fhe_group_sync": (ci, [vp]),
    "fhe_group_partition": (ci, [vp, sz, ci, ctypes.POINTER(sz), ctypes.POINTER(sz)]),
    "fhe_group_launch": (ci, [vp, ci, vp, pu64]),
    "fhe_group_ring_create": (ci, [vp, u32, u32, vp, vp, ctypes.POINTER(vp)]),
    "fhe_group_ring_destroy": (ci, [vp]),
    "fhe_group_ring_at": (vp, [vp, ci]),
    "fhe_group_alloc": (ci, [vp, sz, pu64]),
    "fhe_group_free": (ci, [vp, pu64]),
    "fhe_group_scatter": (ci, [vp, pu64, vp, sz, ci]),
    "fhe_group_gather": (ci, [vp, vp, pu64, sz, ci]),
    "fhe_group_dcrt_op": (ci, [vp, ci, pu64, pu64, pu64, sz, ci]),
    "fhe_group_ntt": (ci, [vp, pu64, sz, ci, ci]),
    "fhe_group_all_gather": (ci, [vp, pu64, pu64, sz, ci]),
    "fhe_group_chain_host": (ci, [vp, ctypes.POINTER(ci), ci, vp, vp, sz]),
    "fhe_comm_unique_id": (ci, [vp, sz]),
    "fhe_comm_create": (ci, [vp, ci, ci, vp, ctypes.POINTER(vp)]),
    "fhe_comm_destroy": (ci, [vp]),
    "fhe_comm_all_gather": (ci, [vp, u64, u64, sz, ci]),
}

FHE_OK, FHE_EINVAL, FHE_ECUDA, FHE_ENOMEM, FHE_ENODEV, FHE_EUNSUPPORTED = 0, -1, -2, -3, -4, -5
OP_ADD, OP_SUB, OP_MUL = 0, 1, 2
OP_NTT, OP_INTT = 3, 4          # fhe_dcrt_chain_host only: SwitchFormat forward / inverse


class FheError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"fhe_b200 error {code}: {msg}")
        self.code = code


def load():
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: build it with `make lib` (or __graft_entry__.build()). "
            "This backend has no CPU fallback.")
    lib = ctypes.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)          # AttributeError if the library does not export it
        fn.restype = res
        fn.argtypes = args
    return lib


lib = load()


def check(rc):
    if rc != 0:
        raise FheError(rc, lib.fhe_last_error().decode())
