"""ctypes binding of libqnn_b200.so (include/qnn_b200.h).

There is no CPU implementation behind this module: if the shared library is missing or a call
fails, the error is raised -- nothing falls back to PyTorch ops or to ``oracle/``.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libqnn_b200.so")

QNN_OK = 0
SRC_CONST = -(2 ** 31)
GATE = {"RX": 0, "RY": 1, "RZ": 2, "U3": 3, "CNOT": 4, "CZ": 5, "CU3": 6}
GATE_ANGLES = {"RX": 1, "RY": 1, "RZ": 1, "U3": 3, "CNOT": 0, "CZ": 0, "CU3": 3}
ENCODER = {"none": 0, "amplitude": 1}
# qnn_circuit_create_ex flags (include/qnn_b200.h)
CREATE_GENERIC_ONLY, CREATE_WIDE_ONLY, CREATE_NARROW_ONLY, CREATE_16_PER_THREAD = 1, 2, 4, 8
VARIANT_FLAGS = {None: 0, "generic": CREATE_GENERIC_ONLY, "wide": CREATE_WIDE_ONLY, "narrow": CREATE_NARROW_ONLY,
                 "16-per-thread": CREATE_16_PER_THREAD}


class QnnOp(C.Structure):
    _fields_ = [("kind", C.c_int32), ("wire0", C.c_int32), ("wire1", C.c_int32),
                ("src", C.c_int32 * 3), ("value", C.c_float * 3)]


class QnatParams(C.Structure):
    _fields_ = [("sigma", C.c_float), ("lo", C.c_float), ("hi", C.c_float), ("eps", C.c_float),
                ("levels", C.c_int32), ("normalise", C.c_int32), ("quantise", C.c_int32),
                ("reserved", C.c_int32)]


class QnnError(RuntimeError):
    pass


_vp, _i32, _i64, _sz = C.c_void_p, C.c_int32, C.c_int64, C.c_size_t

# name -> (restype, argtypes); mirrors include/qnn_b200.h one to one (tests check the export list).
SIGNATURES = {
    "qnn_circuit_create": (C.c_int, [C.POINTER(QnnOp), _i32, _i32, _i32, _i32, _i32, C.POINTER(_vp)]),
    "qnn_circuit_create_ex": (C.c_int, [C.POINTER(QnnOp), _i32, _i32, _i32, _i32, _i32, C.c_uint32, C.POINTER(_vp)]),
    "qnn_circuit_destroy": (None, [_vp]),
    "qnn_circuit_n_qubits": (_i32, [_vp]),
    "qnn_circuit_n_inputs": (_i32, [_vp]),
    "qnn_circuit_n_weights": (_i32, [_vp]),
    "qnn_circuit_describe": (_sz, [_vp, C.c_char_p, _sz]),
    "qnn_circuit_plan": (C.c_char_p, [_vp, _i64]),
    "qnn_forward_launches": (_i32, [_vp, _i64]),
    "qnn_backward_launches": (_i32, [_vp, _i64]),
    "qnn_forward_workspace_bytes": (_sz, [_vp, _i64]),
    "qnn_backward_workspace_bytes": (_sz, [_vp, _i64]),
    "qnn_forward": (C.c_int, [_vp, _vp, _vp, _vp, _i64, _vp, _sz, _vp]),
    "qnn_backward": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp, _i64, _vp, _sz, _vp]),
    "qnn_state_bytes": (_sz, [_vp, _i64]),
    "qnn_forward_save": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _i64, _vp, _sz, _vp]),
    "qnn_backward_saved": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _i64, _vp, _sz, _vp]),
    "qnn_static_scratch_bytes": (_sz, []),
    "qnn_backward_static": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp, _i64, _vp]),
    "qnat_workspace_bytes": (_sz, [_i64, _i32]),
    "qnat_batch_sums": (C.c_int, [_vp, _vp, _i64, _i32, _vp, _sz, _vp]),
    "qnat_apply": (C.c_int, [_vp, _vp, _vp, _i64, _vp, _vp, _vp, _i64, _i32, C.POINTER(QnatParams), _vp]),
    "qnat_backward_sums": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _i64, _i32, C.POINTER(QnatParams), _vp, _sz, _vp]),
    "qnat_backward_apply": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _i64, _vp, _i64, _i32, C.POINTER(QnatParams), _vp]),
    "qnn_pruned_adamw_step": (C.c_int, [_vp, _vp, _vp, _vp, _i64, C.c_float, C.c_float, C.c_float, C.c_float,
                                        C.c_float, _i64, C.c_float, _vp, _vp]),
    "qnn_pruned_adamw_step_counted": (C.c_int, [_vp, _vp, _vp, _vp, _i64, C.c_float, C.c_float, C.c_float, C.c_float,
                                                C.c_float, _vp, C.c_float, _vp, _vp]),
    "qnn_head_forward": (C.c_int, [_vp, _vp, _vp, _vp, _i64, _i32, _i32, _vp]),
    "qnn_head_workspace_bytes": (_sz, [_i64, _i32, _i32]),
    "qnn_head_backward": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _i64, _i32, _i32, _vp, _sz, _vp]),
    "qnn_fp32_peak_probe": (C.c_int, [_i32, C.POINTER(C.c_double), C.POINTER(C.c_double)]),
    "qnn_last_error": (C.c_char_p, []),
    "qnn_abi_version": (_i32, []),
}

_lib = None


def lib():
    """Load the shared library once.  Raises if it has not been built (``__graft_entry__.build()``)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise QnnError(
                f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "(nvcc, sm_100a).  This package has no CPU or PyTorch fallback.")
        handle = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(handle, name)
            fn.restype, fn.argtypes = res, args
        _lib = handle
    return _lib


def check(rc: int, what: str) -> None:
    if rc != QNN_OK:
        msg = lib().qnn_last_error()
        raise QnnError(f"{what} failed (code {rc}): {msg.decode() if msg else ''}")
