"""ctypes binding of libluxb200.so (include/luxb200.h).  No fallback: if the CUDA library is missing or no GPU is
present, calls raise LuxB200Error — the product never routes through a CPU path."""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
SO_PATH = os.environ.get("LUXB200_SO") or os.path.join(HERE, "libluxb200.so")   # override: A/B builds when profiling

VP, U64, U32, INT = C.c_void_p, C.c_uint64, C.c_uint32, C.c_int

ERRORS = {-1: "INVALID_ARG", -2: "NO_DEVICE", -3: "CUDA", -4: "UNSUPPORTED", -5: "NO_KEYS", -6: "ALLOC"}

METHOD = {"AP": 1, "GINX": 2, "LMKCDEY": 3}
GATE = {"OR": 0, "AND": 1, "NOR": 2, "NAND": 3, "XOR": 4, "XNOR": 5, "MAJORITY": 6, "AND3": 7, "OR3": 8, "AND4": 9,
        "OR4": 10, "XOR_FAST": 11, "XNOR_FAST": 12, "CMUX": 13}


class LuxB200Error(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"luxb200 error {code} ({ERRORS.get(code, '?')}): {msg}")
        self.code = code


class Params(C.Structure):
    _fields_ = [("n", U32), ("N", U32), ("q", U64), ("Q", U64), ("qKS", U64), ("baseKS", U32), ("baseG", U32),
                ("baseR", U32), ("numAutoKeys", U32), ("method", U32), ("reserved", U32), ("psi", U64)]


# every symbol include/luxb200.h declares, with its signature
SIGNATURES = {
    "luxb200_last_error": (C.c_char_p, []),
    "luxb200_version": (INT, []),
    "luxb200_device_count": (INT, []),
    "luxb200_last_prime": (U64, [U32, U64]),
    "luxb200_previous_prime": (U64, [U64, U64]),
    "luxb200_root_of_unity": (U64, [U64, U64]),
    "luxb200_ctx_create": (INT, [C.POINTER(Params), INT, C.POINTER(VP)]),
    "luxb200_ctx_destroy": (None, [VP]),
    "luxb200_ctx_info": (INT, [VP, VP]),
    "luxb200_bsk_len": (U64, [VP]),
    "luxb200_ksk_len": (U64, [VP]),
    "luxb200_load_bsk": (INT, [VP, VP, INT]),
    "luxb200_load_ksk": (INT, [VP, VP, VP, INT]),
    "luxb200_btkeygen": (INT, [VP, VP, U64, VP]),
    "luxb200_export_keys": (INT, [VP, VP, VP, VP]),
    "luxb200_synchronize": (INT, [VP]),
    "luxb200_evalbingate_batch": (INT, [VP, INT, VP, VP, VP, VP, VP, VP, U64, INT]),
    "luxb200_evalbingate_batch_ex": (INT, [VP, INT, VP, VP, VP, VP, VP, VP, U64, INT, U64, INT]),
    "luxb200_switch_to_qn_batch": (INT, [VP, VP, VP, VP, VP, U64]),
    "luxb200_broadcast_keys": (INT, [VP, INT]),
    "luxb200_host_alloc": (INT, [U64, C.POINTER(VP)]),
    "luxb200_host_free": (INT, [VP]),
    "luxb200_evalfloor_batch": (INT, [VP, VP, VP, VP, VP, U64, U64, U64, U32]),
    "luxb200_evalsign_batch": (INT, [VP, INT, VP, VP, VP, VP, U64, U64, U64, INT]),
    "luxb200_evaldecomp_batch": (INT, [VP, INT, VP, VP, VP, VP, VP, U32, VP, U64, U64, U64]),
    "luxb200_evaldecomp_count": (U32, [U64, U64, U64]),
    "luxb200_share_ksk": (INT, [VP, VP]),
    "luxb200_evalbingate_multi_batch": (INT, [VP, INT, U32, VP, VP, VP, VP, U64, U64]),
    "luxb200_bootstrap_batch": (INT, [VP, VP, VP, VP, VP, U64, U64, INT]),
    "luxb200_evalfunc_batch": (INT, [VP, VP, VP, VP, U64, U64, VP, VP, U64, U64]),
    "luxb200_blind_rotate_batch": (INT, [VP, VP, U64, VP, U64]),
    "luxb200_external_product_batch": (INT, [VP, VP, VP, VP, U64]),
    "luxb200_dev_alloc": (INT, [INT, U64, C.POINTER(VP)]),
    "luxb200_dev_free": (INT, [INT, VP]),
    "luxb200_dev_copy": (INT, [INT, VP, VP, U64, INT]),
    "luxb200_dev_mem_info": (INT, [INT, C.POINTER(U64), C.POINTER(U64)]),
    "luxb200_keyswitch_batch": (INT, [VP, VP, VP, VP, VP, U64]),
    "luxb200_modswitch_batch": (INT, [VP, U64, U64, VP, VP, U64]),
    "luxb200_ntt_plan_create": (INT, [U32, U32, VP, VP, INT, C.POINTER(VP)]),
    "luxb200_ntt_plan_destroy": (None, [VP]),
    "luxb200_ntt_batch": (INT, [VP, VP, U64, INT]),
    "luxb200_ntt_batch_dev": (INT, [VP, VP, U64, INT, INT]),
    "luxb200_ntt_plan_sync": (INT, [VP]),
    "luxb200_ntt_last_ms": (C.c_float, [VP]),
    "luxb200_rns_plan_create": (INT, [U32, U32, VP, U32, VP, INT, C.POINTER(VP)]),
    "luxb200_rns_plan_destroy": (None, [VP]),
    "luxb200_approx_switch_crt_basis": (INT, [VP, VP, VP, U64, INT]),
    "luxb200_approx_switch_crt_basis_static": (INT, [VP, VP, VP, U64, INT]),
    "luxb200_approx_mod_switch": (INT, [VP, VP, VP, U64, INT]),
    "luxb200_approx_mod_up_dev": (INT, [VP, VP, VP, U64]),
    "luxb200_approx_mod_down_dev": (INT, [VP, VP, VP, U64]),
    "luxb200_approx_mod_switch_bgv": (INT, [VP, VP, VP, U64, INT, U64]),
    "luxb200_hks_plan_create": (INT, [U32, U32, VP, U32, VP, U32, U32, INT, C.POINTER(VP)]),
    "luxb200_hks_plan_destroy": (None, [VP]),
    "luxb200_hks_num_digits": (U32, [VP]),
    "luxb200_hks_keyswitch": (INT, [VP, VP, VP, VP, VP, U64, U64]),
    "luxb200_hks_keyswitch_dev": (INT, [VP, VP, VP, VP, VP, U64, U64]),
    "luxb200_rns_keyswitch_inner": (INT, [VP, VP, VP, VP, VP, U64, U32, U32]),
    "luxb200_rns_keyswitch_inner_dev": (INT, [VP, VP, VP, VP, VP, U64, U32, U32]),
    "luxb200_approx_mod_down_bgv_dev": (INT, [VP, VP, VP, U64, U64]),
    "luxb200_evalbingate_batch_dev": (INT, [VP, INT, VP, VP, VP, VP, VP, VP, U64]),
    "luxb200_circuit_batch": (INT, [VP, U32, VP, U32, VP, U32, VP, VP, VP, VP, U64]),
    "luxb200_circuit_batch_ex": (INT, [VP, U32, VP, U32, VP, U32, VP, U32, U64, VP, VP, VP, VP, U64]),
    "luxb200_circuit_last_stats": (INT, [VP, VP]),
    "luxb200_last_timing": (INT, [VP, VP]),
    "luxb200_launch_count": (U64, [VP]),
    "luxb200_intpipe_probe": (INT, [INT, C.POINTER(C.c_double), C.POINTER(C.c_double)]),
    "luxb200_intpipe_probe64": (INT, [INT, C.POINTER(C.c_double)]),
}

_lib = None


def load():
    """Load libluxb200.so and bind every declared symbol. Raises if the library was not built."""
    global _lib
    if _lib is None:
        if not os.path.exists(SO_PATH):
            raise LuxB200Error(-2, f"{SO_PATH} not built: run `python -m fhe_b200.build` (nvcc, sm_100a). "
                                   "There is no CPU fallback.")
        L = C.CDLL(SO_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def check(rc):
    if rc != 0:
        raise LuxB200Error(rc, load().luxb200_last_error().decode())


def ptr(x):
    """Host numpy array (C-contiguous) or raw integer device pointer -> void*."""
    if isinstance(x, np.ndarray):
        assert x.flags["C_CONTIGUOUS"]
        return x.ctypes.data_as(VP)
    return VP(int(x))
