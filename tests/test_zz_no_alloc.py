"""GPU: qnn_backward_static -- the backward pass for hosts that may not allocate device memory.

A real-time channel-estimation loop, a process whose HBM belongs entirely to someone else's pool, or a
bare C/FFI host cannot give ``qnn_backward`` a workspace.  ``qnn_backward_static`` keeps that scratch in
a fixed array inside the library instead.  This file checks, through the C ABI, that it returns what
``qnn_backward`` returns and what the CPU oracle computes (same tolerance as the other parity tests) on calls in a row that share the
array, and that a request larger than the array is refused with a readable message.  It is named to be
collected last, and depends on nothing the other tests leave behind.
"""
import numpy as np
import pytest
import torch

import qris_b200
from helpers import REL_TOL, assert_close, oracle_ops, seeded_inputs
from oracle import qnn_oracle as O

pytestmark = pytest.mark.gpu


def _backward_pair(circ, B, seed):
    L = qris_b200._lib
    lib = L.lib()
    dev = torch.device("cuda:0")
    x, w, gy = (t.to(dev) for t in seeded_inputs(circ, B, seed))
    h = circ.handle(0)
    st = torch.cuda.current_stream(dev).cuda_stream
    # existing entry point, caller-provided workspace
    gx_ref, gw_ref = torch.empty_like(x), torch.empty(circ.n_weights, device=dev)
    ws = torch.empty(lib.qnn_backward_workspace_bytes(h, B), dtype=torch.uint8, device=dev)
    L.check(lib.qnn_backward(h, x.data_ptr(), w.data_ptr(), gy.data_ptr(), gx_ref.data_ptr(), gw_ref.data_ptr(),
                             B, ws.data_ptr(), ws.numel(), st), "qnn_backward")
    # new entry point: inputs and outputs only
    gx, gw = torch.full_like(x, float("nan")), torch.full((circ.n_weights,), float("nan"), device=dev)
    L.check(lib.qnn_backward_static(h, x.data_ptr(), w.data_ptr(), gy.data_ptr(), gx.data_ptr(), gw.data_ptr(), B, st),
            "qnn_backward_static")
    torch.cuda.synchronize()
    # and the CPU oracle's gradients for the same inputs
    _, gxo, gwo = O.forward_and_grads(oracle_ops(circ), circ.n_qubits, x.cpu(), w.cpu(), gy.cpu())
    return (gx.cpu().numpy(), gw.cpu().numpy()), (gx_ref.cpu().numpy(), gw_ref.cpu().numpy()), (gxo.numpy(), gwo.numpy())


def test_static_scratch_matches_workspace_version_on_calls_in_a_row():
    lib = qris_b200._lib.lib()
    assert lib.qnn_static_scratch_bytes() == 4 << 20
    # different circuits and batches back to back: each later call finds the earlier call's leftovers
    for circ, B, seed in ((qris_b200.u3cu3_circuit(8, 2), 300, 1), (qris_b200.reference_circuit(6, 3), 4096, 2),
                          (qris_b200.u3cu3_circuit(8, 2), 37, 3)):
        (gx, gw), (gx_ref, gw_ref), (gx_or, gw_or) = _backward_pair(circ, B, seed)
        assert np.isfinite(gx).all() and np.isfinite(gw).all()
        assert_close(gx, gx_ref, rel=REL_TOL, scale=float(np.abs(gx_ref).max()), what="gx")
        assert_close(gw, gw_ref, rel=REL_TOL, scale=float(np.abs(gw_ref).max()), what="gw")
        assert_close(gx, gx_or, rel=REL_TOL, scale=float(np.abs(gx_or).max()), what="gx vs oracle")
        assert_close(gw, gw_or, rel=REL_TOL, scale=float(np.abs(gw_or).max()), what="gw vs oracle")


def test_request_beyond_the_array_is_refused():
    L = qris_b200._lib
    lib = L.lib()
    circ = qris_b200.u3cu3_circuit(15, 1)          # 64 samples in flight: 32 MiB of state vectors
    h = circ.handle(0)
    B = 64
    # the static entry point runs the generic kernel family: its scratch need is that family's workspace
    assert lib.qnn_backward_workspace_bytes(circ.generic().handle(0), B) > lib.qnn_static_scratch_bytes()
    x = torch.zeros(B, 15, device="cuda"); w = torch.zeros(circ.n_weights, device="cuda")
    gy = torch.zeros(B, 15, device="cuda"); gx = torch.empty_like(x); gw = torch.empty_like(w)
    rc = lib.qnn_backward_static(h, x.data_ptr(), w.data_ptr(), gy.data_ptr(), gx.data_ptr(), gw.data_ptr(), B, None)
    assert rc == -4
    msg = lib.qnn_last_error().decode()
    assert "QNN_STATIC_SCRATCH_BYTES" in msg and "qnn_backward" in msg
    assert lib.qnn_backward_static(None, None, None, None, None, None, B, None) == -1
