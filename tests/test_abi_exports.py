"""CPU: the C-ABI library loads and exports every symbol include/*.h declares (no compute calls)."""
import ctypes
import glob
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    names = []
    for h in glob.glob(os.path.join(ROOT, "include", "*.h")):
        src = open(h).read()
        src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
        names += re.findall(r"QNN_API\s+[\w\s\*]+?\b(q\w+)\s*\(", src)
    return sorted(set(names))


def test_header_declares_entry_points():
    names = _declared()
    for must in ("qnn_circuit_create", "qnn_forward", "qnn_backward", "qnat_apply", "qnn_last_error"):
        assert must in names


def test_library_exports_every_declared_symbol():
    import qris_b200
    path = qris_b200._lib.LIB_PATH
    if not os.path.exists(path):
        import __graft_entry__
        __graft_entry__.build()
    lib = ctypes.CDLL(path)
    for name in _declared():
        assert hasattr(lib, name), f"{name} declared in include/ but not exported by {path}"
    assert set(qris_b200._lib.SIGNATURES) == set(_declared()), "ctypes signature table out of sync with the header"
    fn = lib.qnn_abi_version
    fn.restype = ctypes.c_int32
    assert fn() == 1


def test_invalid_arguments_are_reported_not_crashed():
    """Argument validation happens before any CUDA call, so it can run without a device."""
    import qris_b200
    L = qris_b200._lib
    lib = L.lib()
    out = ctypes.c_void_p()
    ops = (L.QnnOp * 1)()
    ops[0].kind, ops[0].wire0, ops[0].wire1 = L.GATE["RY"], 7, -1
    ops[0].src[0] = 0
    assert lib.qnn_circuit_create(ops, 1, 4, 4, 1, 0, ctypes.byref(out)) == -1
    assert b"wires" in lib.qnn_last_error()
    assert lib.qnn_circuit_create(ops, 1, 17, 4, 1, 0, ctypes.byref(out)) == -1
    assert b"n_qubits" in lib.qnn_last_error()
    ops[0].wire0 = 0
    ops[0].src[0] = 5
    assert lib.qnn_circuit_create(ops, 1, 4, 4, 1, 0, ctypes.byref(out)) == -1
    assert b"weight index" in lib.qnn_last_error()
    assert lib.qnn_forward(None, None, None, None, 4, None, 0, None) == -1
