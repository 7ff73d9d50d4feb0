"""ctypes binding of libscaldpc.so (the C ABI declared in include/scaldpc.h).

There is no CPU fallback: if the library is missing this module raises, and if no CUDA device is visible
every *_create call fails with ScaldpcError(SCALDPC_ERR_CUDA).
"""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libscaldpc.so")

# status codes (include/scaldpc.h)
OK = 0
ERR_ARG, ERR_PROB_SUM, ERR_NO_MAX, ERR_NO_CONFIG, ERR_DEGREE, ERR_STRUCTURE, ERR_RANGE, ERR_CUDA, ERR_UNSUPPORTED = (
    -1, -2, -3, -4, -5, -6, -7, -8, -9)
# decode flags
IN_LLR, DEVICE_PTRS, CN_EXACT, CN_GENERIC, NO_VERIFY, QARY_EARLY_STOP = 0x1, 0x2, 0x4, 0x8, 0x100, 0x200
BP_STATIC_BYTES = 8 << 20  # SCALDPC_BP_STATIC_BYTES: the static scratch of scaldpc_bp_decode_static
BP_PROBS_PER_TRIAL, BP_NO_EARLY_STOP, BP_NO_COMPACT = 0x10, 0x20, 0x40

# every exported symbol of include/scaldpc.h (tests/test_abi.py checks this list against the header)
SYMBOLS = [
    "scaldpc_last_error", "scaldpc_version", "scaldpc_device_count",
    "scaldpc_qary_create", "scaldpc_qary_create_csr", "scaldpc_qary_destroy", "scaldpc_qary_info", "scaldpc_qary_set_max_iter",
    "scaldpc_qary_decode", "scaldpc_qary_decode_observed", "scaldpc_qary_simulate", "scaldpc_qary_launch_count", "scaldpc_qary_time_kernel",
    "scaldpc_qary_last_run", "scaldpc_qary_set_risk_margin", "scaldpc_qary_last_iterations",
    "scaldpc_bp_create", "scaldpc_bp_destroy", "scaldpc_bp_info", "scaldpc_bp_decode", "scaldpc_bp_decode_scored", "scaldpc_bp_decode_nested", "scaldpc_bp_launch_count",
    "scaldpc_bp_create_multi", "scaldpc_bp_decode_multi", "scaldpc_bp_create_circulant", "scaldpc_bp_decode_static", "scaldpc_bp_last_early_copy", "scaldpc_qary_decode_one", "scaldpc_qary_decode_f64", "scaldpc_qary_one_stats",
    "scaldpc_bp_last_run", "scaldpc_bp_time_kernel",
]


class ScaldpcError(RuntimeError):
    """A failed libscaldpc call; `.code` is the SCALDPC_ERR_* value."""

    def __init__(self, code: int, message: str):
        super().__init__(f"libscaldpc error {code}: {message}")
        self.code = code


_lib = None


def lib() -> C.CDLL:
    """Load libscaldpc.so (built by sca_ldpc_b200._build / __graft_entry__.build()).  Fails loudly if absent."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: build it with `python -m sca_ldpc_b200._build` "
            "(nvcc, sm_100a).  sca_ldpc_b200 has no CPU or PyTorch fallback path.")
    L = C.CDLL(LIB_PATH)
    vp, i, u = C.c_void_p, C.c_int, C.c_uint
    ip = C.POINTER(C.c_int)
    L.scaldpc_last_error.restype = C.c_char_p
    L.scaldpc_last_error.argtypes = []
    L.scaldpc_version.restype = C.c_char_p
    L.scaldpc_version.argtypes = []
    L.scaldpc_device_count.restype = i
    L.scaldpc_device_count.argtypes = []

    L.scaldpc_qary_create.restype = i
    L.scaldpc_qary_create.argtypes = [vp, i, i, i, i, i, i, i, i, C.POINTER(vp)]
    L.scaldpc_qary_create_csr.restype = i
    L.scaldpc_qary_create_csr.argtypes = [i, i, vp, vp, vp, i, i, i, i, i, i, C.POINTER(vp)]
    L.scaldpc_qary_destroy.restype = None
    L.scaldpc_qary_destroy.argtypes = [vp]
    L.scaldpc_qary_info.restype = i
    L.scaldpc_qary_info.argtypes = [vp, ip, ip, ip, ip, ip, ip, ip]
    L.scaldpc_qary_set_max_iter.restype = i
    L.scaldpc_qary_set_max_iter.argtypes = [vp, i]
    L.scaldpc_qary_decode.restype = i
    L.scaldpc_qary_decode.argtypes = [vp, i, vp, vp, vp, vp, vp, vp, u, vp]
    L.scaldpc_qary_decode_observed.restype = i
    L.scaldpc_qary_decode_observed.argtypes = [vp, i, vp, i, vp, i, vp, vp, vp, vp, vp, vp, u, vp]
    L.scaldpc_qary_simulate.restype = i
    L.scaldpc_qary_simulate.argtypes = [vp, i, C.c_uint64, C.c_uint64, vp, i, vp, i, vp, i, vp, i, C.c_double, i,
                                        vp, vp, vp, vp, vp, vp, u, vp]
    L.scaldpc_qary_launch_count.restype = C.c_longlong
    L.scaldpc_qary_launch_count.argtypes = [vp]
    L.scaldpc_qary_last_run.restype = i
    L.scaldpc_qary_last_run.argtypes = [vp, C.POINTER(C.c_longlong), C.POINTER(C.c_longlong), C.POINTER(C.c_longlong)]
    L.scaldpc_qary_last_iterations.restype = i
    L.scaldpc_qary_last_iterations.argtypes = [vp, vp, i, u, vp]
    L.scaldpc_qary_set_risk_margin.restype = i
    L.scaldpc_qary_set_risk_margin.argtypes = [vp, C.c_float]
    L.scaldpc_qary_time_kernel.restype = i
    L.scaldpc_qary_time_kernel.argtypes = [vp, i, i, u, i, vp, C.POINTER(C.c_float)]

    L.scaldpc_bp_create.restype = i
    L.scaldpc_bp_create.argtypes = [i, i, vp, vp, i, i, C.POINTER(vp)]
    L.scaldpc_bp_destroy.restype = None
    L.scaldpc_bp_destroy.argtypes = [vp]
    L.scaldpc_bp_info.restype = i
    L.scaldpc_bp_info.argtypes = [vp, ip, ip, ip, ip]
    L.scaldpc_bp_decode.restype = i
    L.scaldpc_bp_decode.argtypes = [vp, i, vp, vp, vp, vp, vp, vp, u, vp]
    L.scaldpc_bp_decode_scored.restype = i
    L.scaldpc_bp_decode_scored.argtypes = [vp, i, i, vp, vp, vp, vp, i, vp, vp, vp, vp, u, vp]
    L.scaldpc_bp_decode_nested.restype = i
    L.scaldpc_bp_decode_nested.argtypes = [vp, i, vp, vp, vp, vp, vp, vp, vp, u, vp]
    L.scaldpc_bp_create_multi.restype = i
    L.scaldpc_bp_create_multi.argtypes = [i, i, i, vp, vp, i, i, C.POINTER(vp)]
    L.scaldpc_bp_create_circulant.restype = i
    L.scaldpc_bp_create_circulant.argtypes = [i, i, i, i, vp, vp, i, i, C.POINTER(vp)]
    L.scaldpc_qary_decode_f64.restype = i
    L.scaldpc_qary_decode_f64.argtypes = [vp, i, vp, vp, vp, vp, vp, vp, C.c_uint]
    L.scaldpc_qary_decode_one.restype = i
    L.scaldpc_qary_decode_one.argtypes = [vp, vp, vp, vp]
    L.scaldpc_qary_one_stats.restype = i
    L.scaldpc_qary_one_stats.argtypes = [vp, C.POINTER(C.c_longlong), C.POINTER(C.c_longlong)]
    L.scaldpc_bp_last_early_copy.restype = i
    L.scaldpc_bp_last_early_copy.argtypes = [vp, C.POINTER(i)]
    L.scaldpc_bp_decode_static.restype = i
    L.scaldpc_bp_decode_static.argtypes = [vp, i, vp, vp, vp, vp, vp, vp]
    L.scaldpc_bp_decode_multi.restype = i
    L.scaldpc_bp_decode_multi.argtypes = [vp, i, vp, vp, vp, vp, vp, i, vp, vp, vp, vp, vp, u, vp]
    L.scaldpc_bp_launch_count.restype = C.c_longlong
    L.scaldpc_bp_launch_count.argtypes = [vp]
    L.scaldpc_bp_last_run.restype = i
    L.scaldpc_bp_last_run.argtypes = [vp, ip, ip]
    L.scaldpc_bp_time_kernel.restype = i
    L.scaldpc_bp_time_kernel.argtypes = [vp, i, i, u, i, vp, C.POINTER(C.c_float)]
    _lib = L
    return L


def check(rc: int) -> int:
    if rc < 0:
        raise ScaldpcError(rc, lib().scaldpc_last_error().decode("utf-8", "replace"))
    return rc


def version() -> str:
    return lib().scaldpc_version().decode()


def device_count() -> int:
    n = lib().scaldpc_device_count()
    return max(n, 0)
