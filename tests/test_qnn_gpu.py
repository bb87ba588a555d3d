"""GPU parity tests: the CUDA path, called through the C ABI, against the CPU oracle on the same
seeded inputs; the committed golden vectors; size-independent properties at BASELINE sizes."""
import os

import numpy as np
import pytest
import torch

import qris_b200
from helpers import REL_TOL, assert_close, oracle_ops, seeded_inputs
from oracle import qnn_oracle as O

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(__file__), "golden", "qnn_golden.npz")


def _plan(circ, B):
    return qris_b200._lib.lib().qnn_circuit_plan(circ.handle(0), B).decode()


def run_gpu(circ, x, w, gy):
    dev = torch.device("cuda:0")
    xd, wd, gd = x.to(dev), w.to(dev), gy.to(dev)
    y = qris_b200.qnn_forward_raw(circ, xd, wd)
    gx, gw = qris_b200.qnn_backward_raw(circ, xd, wd, gd)
    torch.cuda.synchronize()
    return y.cpu().numpy(), gx.cpu().numpy(), gw.cpu().numpy()


def check_against_oracle(circ, B, seed=0):
    x, w, gy = seeded_inputs(circ, B, seed)
    y, gx, gw = run_gpu(circ, x, w, gy)
    yo, gxo, gwo = O.forward_and_grads(oracle_ops(circ), circ.n_qubits, x, w, gy, encoder=circ.encoder)
    assert_close(y, yo.numpy(), what="y")
    assert_close(gx, gxo.numpy(), scale=float(np.abs(gxo.numpy()).max()), what="gx")
    # gw sums B per-sample terms: fp32 error scales with the magnitude of what is summed
    assert_close(gw, gwo.numpy(), scale=float(np.abs(gwo.numpy()).max()), what="gw")


@pytest.mark.parametrize("n,L,B", [(1, 2, 5), (2, 1, 7), (3, 3, 33), (4, 3, 256), (6, 3, 256), (6, 3, 1),
                                    (8, 3, 64), (10, 3, 40), (10, 1, 3)])
def test_reference_circuit_parity(n, L, B):
    check_against_oracle(qris_b200.reference_circuit(n, L), B, seed=n * 100 + L)


@pytest.mark.parametrize("n,L,B", [(2, 2, 9), (5, 2, 64), (7, 1, 31), (10, 3, 24)])
def test_u3cu3_circuit_parity(n, L, B):
    check_against_oracle(qris_b200.u3cu3_circuit(n, L), B, seed=n * 100 + L)


@pytest.mark.parametrize("n,L,B", [(3, 2, 17), (8, 1, 12), (10, 1, 5)])
def test_amplitude_encoder_parity(n, L, B):
    check_against_oracle(qris_b200.u3cu3_circuit(n, L, encoder="amplitude"), B, seed=n)


@pytest.mark.parametrize("n,n_inputs,B", [(5, 32, 40), (6, 50, 33), (8, 256, 64), (8, 300, 7), (10, 1024, 20), (9, 1, 3)])
def test_amplitude_encoder_on_the_vm(n, n_inputs, B):
    """Rows shorter than 2^n are zero padded, longer ones truncated (their extra columns get zero
    gradient); 8 qubits x 256 inputs is the raw (2, 16, 8) pilot tensor of Run:345."""
    base = qris_b200.u3cu3_circuit(n, 2, encoder="amplitude")
    circ = qris_b200.Circuit(n, base.ops, n_inputs=n_inputs, n_weights=base.n_weights, encoder="amplitude")
    assert _plan(circ, B) == "vm-register"
    check_against_oracle(circ, B, seed=n + n_inputs)


@pytest.mark.parametrize("n,L,B", [(11, 2, 5), (12, 1, 3), (13, 1, 2)])
def test_cta_per_sample_sizes(n, L, B):
    check_against_oracle(qris_b200.u3cu3_circuit(n, L), B, seed=n)
    check_against_oracle(qris_b200.reference_circuit(n, L), B, seed=n + 1)


@pytest.mark.parametrize("n,B", [(14, 2), (14, 80), (15, 3), (16, 1), (16, 20)])
def test_largest_sizes(n, B):
    """n = 16 is QNN_MAX_QUBITS (BASELINE config 4's width): 512 KiB of state per sample, held in the
    registers of a thread-block cluster (2 / 4 / 8 CTAs for n = 14 / 15 / 16)."""
    circ = qris_b200.u3cu3_circuit(n, 1)
    assert _plan(circ, B) == "vm-register"
    check_against_oracle(circ, B, seed=n)


def test_config4_exact_circuit():
    """BASELINE configs[3] as bench.py runs it: 16 qubits, U3+CU3, EIGHT layers (768 weights), both the plain
    pair and the saved-state pair, against the oracle and the frozen golden vector of the same circuit."""
    circ = qris_b200.u3cu3_circuit(16, 8)
    assert _plan(circ, 2) == "vm-register" and circ.n_weights == 768
    check_against_oracle(circ, 2, seed=168)
    x, w, gy = (t.cuda() for t in seeded_inputs(circ, 3, seed=169))
    y1, state = qris_b200.qnn_forward_raw(circ, x, w, save_state=True)
    gx0, gw0 = qris_b200.qnn_backward_raw(circ, x, w, gy)
    gx1, gw1 = qris_b200.qnn_backward_raw(circ, x, w, gy, state=state)
    assert torch.equal(y1, qris_b200.qnn_forward_raw(circ, x, w))
    assert_close(gx1.cpu().numpy(), gx0.cpu().numpy(), scale=float(gx0.abs().max()), what="gx saved vs recompute")
    assert_close(gw1.cpu().numpy(), gw0.cpu().numpy(), scale=float(gw0.abs().max()), what="gw saved vs recompute")


@pytest.mark.parametrize("n,n_inputs,B", [(6, 50, 33), (5, 32, 40), (8, 300, 7), (9, 1, 3), (7, 128, 1), (10, 1024, 5)])
def test_amplitude_encoder_ragged_batches_are_finite(n, n_inputs, B):
    """A warp carries 2^(10-n) samples; the padding samples of a partly filled last warp must not leak NaNs
    into the warp-wide weight-gradient sums (plain pair and saved-state pair), and an all-zero row is the
    zero state: y = 0 and zero gradients, not NaN."""
    base = qris_b200.u3cu3_circuit(n, 2, encoder="amplitude")
    circ = qris_b200.Circuit(n, base.ops, n_inputs=n_inputs, n_weights=base.n_weights, encoder="amplitude")
    x, w, gy = seeded_inputs(circ, B, seed=n + B)
    xd, wd, gd = x.cuda(), w.cuda(), gy.cuda()
    yo, gxo, gwo = O.forward_and_grads(oracle_ops(circ), n, x, w, gy, encoder="amplitude")
    y, state = qris_b200.qnn_forward_raw(circ, xd, wd, save_state=True)
    for st in (None, state):
        gx, gw = qris_b200.qnn_backward_raw(circ, xd, wd, gd, state=st)
        assert torch.isfinite(gx).all() and torch.isfinite(gw).all() and torch.isfinite(y).all()
        assert_close(gw.cpu().numpy(), gwo.numpy(), scale=float(gwo.abs().max()), what="gw")
        assert_close(gx.cpu().numpy(), gxo.numpy(), scale=float(gxo.abs().max()), what="gx")
    xz = xd.clone(); xz[B // 2] = 0
    for c in (circ, circ.generic()):
        yz = qris_b200.qnn_forward_raw(c, xz, wd)
        gxz, gwz = qris_b200.qnn_backward_raw(c, xz, wd, gd)
        assert torch.isfinite(yz).all() and torch.isfinite(gxz).all() and torch.isfinite(gwz).all()
        assert yz[B // 2].abs().max() == 0 and gxz[B // 2].abs().max() == 0


@pytest.mark.parametrize("n,L,B", [(14, 6, 300), (15, 4, 150), (16, 4, 90)])
def test_cluster_exchanges_are_deterministic_under_load(n, L, B):
    """Many samples per cluster and two forward clusters per SM: the alternating shared-memory buffers, the second
    cluster barrier of the forward kernel and the L2 slices of the backward pass are reused hundreds of times.  A race
    in that protocol would show as run-to-run differences: three runs must agree bit for bit (plain pair and
    saved-state pair), sampled rows must match the oracle, and the result must not depend on which rows share a call."""
    circ = qris_b200.u3cu3_circuit(n, L)
    x, w, gy = (t.cuda() for t in seeded_inputs(circ, B, seed=n * L))
    runs = []
    for _ in range(3):
        y, state = qris_b200.qnn_forward_raw(circ, x, w, save_state=True)
        gx, gw = qris_b200.qnn_backward_raw(circ, x, w, gy, state=state)
        gx2, gw2 = qris_b200.qnn_backward_raw(circ, x, w, gy)
        runs.append((y.clone(), gx.clone(), gw.clone(), gx2.clone(), gw2.clone()))
    for r in runs[1:]:
        assert all(torch.equal(a, b) for a, b in zip(r, runs[0]))
    idx = torch.tensor([0, B // 2, B - 1])
    yo, gxo, _ = O.forward_and_grads(oracle_ops(circ), n, x[idx].cpu(), w.cpu(), gy[idx].cpu())
    assert_close(runs[0][0][idx].cpu().numpy(), yo.numpy(), what="rows y")
    assert_close(runs[0][1][idx].cpu().numpy(), gxo.numpy(), scale=float(gxo.abs().max()), what="rows gx")
    ys = qris_b200.qnn_forward_raw(circ, x[idx.cuda()].contiguous(), w)
    assert torch.equal(ys, runs[0][0][idx.cuda()])
    assert_close(runs[0][4].cpu().numpy(), runs[0][2].cpu().numpy(), rel=2e-5, scale=float(runs[0][2].abs().max()), what="gw recompute vs saved")


def test_deep_cluster_circuit_falls_back_to_l2_exchanges():
    """Programs whose tables leave no room for the 64 KB local-exchange buffer are compiled with L2
    exchanges only and still run on the VM."""
    c = qris_b200.u3cu3_circuit(14, 20)
    d = c.describe()
    assert d["ok"] and sum(x["local"] for x in d["exchanges"]) == 0
    assert _plan(c, 2) == "vm-register"
    check_against_oracle(c, 2, seed=7)
    # just below the switch: local exchanges are used and the program still fits
    c16 = qris_b200.u3cu3_circuit(14, 12)
    assert sum(x["local"] for x in c16.describe()["exchanges"]) > 0 and _plan(c16, 2) == "vm-register"


def test_largest_sizes_generic_family():
    for n, B in ((14, 2), (16, 1)):
        circ = qris_b200.u3cu3_circuit(n, 1).generic()
        assert _plan(circ, B) == "generic-cta-global"
        check_against_oracle(circ, B, seed=n)


def test_mixed_sources_reupload_and_constants():
    """Input columns used twice (data re-uploading), constants, shared weights, RX and CZ."""
    Op = qris_b200.Op
    ops = [Op("RX", (0,), (("x", 0),)), Op("RY", (1,), (("x", 1),)), Op("RZ", (2,), (("x", 0),)),
           Op("U3", (2,), (("x", 2), ("w", 0), ("c", 0.3))), Op("CZ", (0, 2)), Op("CNOT", (2, 1)),
           Op("RY", (0,), (("w", 1),)), Op("RY", (1,), (("w", 1),)), Op("CU3", (1, 0), (("w", 2), ("x", 1), ("w", 0))),
           Op("RX", (2,), (("c", -1.1),)), Op("RZ", (1,), (("w", 3),)), Op("RX", (1,), (("w", 4),))]
    circ = qris_b200.Circuit(3, ops, n_inputs=3, n_weights=5)
    check_against_oracle(circ, 50, seed=9)


def test_golden_vectors():
    g = np.load(GOLDEN)
    for name in sorted({k.split("/")[0] for k in g.files}):
        n, L, B, enc = (int(v) for v in g[f"{name}/meta"])
        circ = (qris_b200.reference_circuit(n, L) if name.startswith("ref")
                else qris_b200.u3cu3_circuit(n, L, encoder="amplitude" if enc else "angle"))
        y, gx, gw = run_gpu(circ, torch.tensor(g[f"{name}/x"]), torch.tensor(g[f"{name}/w"]), torch.tensor(g[f"{name}/gy"]))
        assert_close(y, g[f"{name}/y"], what=f"{name} y")
        assert_close(gx, g[f"{name}/gx"], scale=float(np.abs(g[f"{name}/gx"]).max()), what=f"{name} gx")
        assert_close(gw, g[f"{name}/gw"], scale=float(np.abs(g[f"{name}/gw"]).max()), what=f"{name} gw")


def test_survey_fixed_vector_and_closed_form():
    x, W = O.survey_fixed_inputs()
    fv = O.SURVEY_FIXED_VECTOR
    circ = qris_b200.reference_circuit(6, 3)
    y, gx, gw = run_gpu(circ, torch.tensor(x, dtype=torch.float32), torch.tensor(W.reshape(-1), dtype=torch.float32),
                        torch.tensor([fv["g"]], dtype=torch.float32))
    assert_close(y[0], np.array(fv["y"]), what="fixed y")
    gW = gw.reshape(3, 6, 2)
    assert abs(gx[0, 0] - fv["dx0"]) < 1e-5 and abs(gW[1, 2, 0] - fv["dW_1_2_0"]) < 2e-5
    assert abs(gW[2, 3, 0] - fv["dW_2_3_0"]) < 1e-5 and abs(gW[2, 3, 1]) < 1e-6
    # analytic L = 1 closed form, at a BASELINE-sized batch
    rng = np.random.default_rng(0)
    xx = rng.uniform(-1, 1, (4096, 6)).astype(np.float32); ww = rng.uniform(0, 2 * np.pi, (1, 6, 2)).astype(np.float32)
    c1 = qris_b200.reference_circuit(6, 1)
    y1 = qris_b200.qnn_forward_raw(c1, torch.tensor(xx).cuda(), torch.tensor(ww.reshape(-1)).cuda()).cpu().numpy()
    assert_close(y1, O.ref_circuit_L1_closed_form(xx, ww), what="closed form")


@pytest.mark.parametrize("kind,n,L,B", [("ref", 6, 3, 4096), ("u3", 10, 3, 65536), ("u3", 16, 8, 576)])
def test_full_size_properties(kind, n, L, B):
    """BASELINE configs 2, 3 and 4 at full batch: properties that need no oracle run of that size."""
    circ = qris_b200.reference_circuit(n, L) if kind == "ref" else qris_b200.u3cu3_circuit(n, L)
    x, w, gy = seeded_inputs(circ, B, seed=42)
    dev = torch.device("cuda:0")
    xd, wd, gd = x.to(dev), w.to(dev), gy.to(dev)
    y = qris_b200.qnn_forward_raw(circ, xd, wd)
    assert torch.isfinite(y).all() and y.abs().max() <= 1 + 1e-5
    gx, gw = qris_b200.qnn_backward_raw(circ, xd, wd, gd)
    # (1) per-sample independence: any row subset reproduces the same rows
    idx = torch.tensor([0, 1, B // 3, B // 2 + 1, B - 1], device=dev)
    ys = qris_b200.qnn_forward_raw(circ, xd[idx].contiguous(), wd)
    gxs, _ = qris_b200.qnn_backward_raw(circ, xd[idx].contiguous(), wd, gd[idx].contiguous())
    assert torch.equal(ys, y[idx]) and torch.allclose(gxs, gx[idx], atol=1e-6 if n <= 10 else 2e-5)
    # (2) those rows agree with the oracle
    yo, gxo, _ = O.forward_and_grads(oracle_ops(circ), n, x[idx.cpu()], w, gy[idx.cpu()])
    assert_close(ys.cpu().numpy(), yo.numpy(), what="rows y")
    assert_close(gxs.cpu().numpy(), gxo.numpy(), scale=float(gxo.abs().max()), what="rows gx")
    # (3) the weight gradient is additive over a split of the batch and linear in gy
    h = B // 2
    _, gwa = qris_b200.qnn_backward_raw(circ, xd[:h].contiguous(), wd, gd[:h].contiguous(), need_gx=False)
    _, gwb = qris_b200.qnn_backward_raw(circ, xd[h:].contiguous(), wd, gd[h:].contiguous(), need_gx=False)
    scale = float(gw.abs().max())
    assert_close((gwa + gwb).cpu().numpy(), gw.cpu().numpy(), rel=2e-5, scale=scale, what="gw additivity")
    _, gw2 = qris_b200.qnn_backward_raw(circ, xd, wd, (2.5 * gd).contiguous(), need_gx=False)
    assert_close(gw2.cpu().numpy(), 2.5 * gw.cpu().numpy(), rel=2e-5, scale=2.5 * scale, what="gw linearity")
    # (4) deterministic: the same call twice gives the same bits
    _, gw3 = qris_b200.qnn_backward_raw(circ, xd, wd, gd, need_gx=False)
    assert torch.equal(gw3, gw)
    if kind == "ref":   # (5) last layer's RZ angles have identically zero gradient
        assert gw.reshape(L, n, 2)[L - 1, :, 1].abs().max().item() <= 1e-3 * scale
    # (6) the batch reduction itself: gw of a contiguous 2,048-sample slice, taken from inside the full batch so
    #     that several CTAs and the finalize kernel take part, against the oracle's gw of the same rows
    lo, cnt = B // 4 + 3, (min(2048, B // 2) if n <= 10 else 3)      # (a 16-qubit, 8-layer oracle row takes seconds)
    _, gws = qris_b200.qnn_backward_raw(circ, xd[lo:lo + cnt].contiguous(), wd, gd[lo:lo + cnt].contiguous(), need_gx=False)
    _, _, gwo = O.forward_and_grads(oracle_ops(circ), n, x[lo:lo + cnt], w, gy[lo:lo + cnt])
    assert_close(gws.cpu().numpy(), gwo.numpy(), scale=float(gwo.abs().max()), what="gw of a 2048-row slice")
    # ... and the full-batch gw is that slice's gw plus the gw of the rest (additivity against an oracle-pinned part)
    rest = torch.cat([torch.arange(0, lo), torch.arange(lo + cnt, B)]).to(dev)
    _, gwr = qris_b200.qnn_backward_raw(circ, xd[rest].contiguous(), wd, gd[rest].contiguous(), need_gx=False)
    assert_close((gws + gwr).cpu().numpy(), gw.cpu().numpy(), rel=2e-5, scale=scale, what="gw = slice + rest")


def test_error_paths():
    circ = qris_b200.reference_circuit(4, 1)
    L = qris_b200._lib
    lib = L.lib()
    h = circ.handle(0)
    x = torch.zeros(8, 4, device="cuda"); w = torch.zeros(8, device="cuda"); y = torch.empty(8, 4, device="cuda")
    need = lib.qnn_backward_workspace_bytes(h, 8)
    assert need > 0
    rc = lib.qnn_backward(h, x.data_ptr(), w.data_ptr(), y.data_ptr(), None, w.data_ptr(), 8, None, 0, None)
    assert rc == -4 and b"workspace" in lib.qnn_last_error()
    rc = lib.qnn_forward(h, x.data_ptr(), w.data_ptr(), y.data_ptr(), 0, None, 0, None)
    assert rc == -1
    with pytest.raises(ValueError):
        qris_b200.qnn_apply(circ, torch.zeros(8, 5, device="cuda"), w)
    assert qris_b200.qnn_forward_raw(circ, torch.zeros(0, 4, device="cuda"), w).shape == (0, 4)


def test_kernel_family_selection_and_both_families_agree():
    """n in [5, 16] runs on the register-resident VM; a circuit created with QNN_CREATE_GENERIC_ONLY
    (``Circuit.generic()``) stays on the generic interpreter.  Both must match the oracle and each other."""
    for make in (lambda: qris_b200.reference_circuit(6, 3), lambda: qris_b200.u3cu3_circuit(10, 2),
                 lambda: qris_b200.u3cu3_circuit(7, 2), lambda: qris_b200.reference_circuit(9, 2)):
        fast = make()
        assert _plan(fast, 64).startswith("vm-register")
        slow = make().generic()
        assert _plan(slow, 64).startswith("generic")
        x, w, gy = seeded_inputs(fast, 77, seed=5)
        a, b = run_gpu(fast, x, w, gy), run_gpu(slow, x, w, gy)
        for u, v, name in zip(a, b, ("y", "gx", "gw")):
            assert_close(u, v, rel=2e-5, scale=float(np.abs(v).max()), what=name)
        check_against_oracle(slow, 33, seed=1)
    assert _plan(qris_b200.reference_circuit(2, 1), 8).startswith("generic")
    assert _plan(qris_b200.reference_circuit(4, 1), 8) == "vm-register-narrow"      # 3 and 4 qubits: narrow variant only
    assert _plan(qris_b200.reference_circuit(12, 1), 8) == "vm-register"
    assert _plan(qris_b200.reference_circuit(14, 1), 8) == "vm-register"


@pytest.mark.parametrize("n", [5, 6, 7, 8, 9, 10])
@pytest.mark.parametrize("B", [1, 3, 32, 1000])
def test_vm_all_widths_and_ragged_batches(n, B):
    """Every lane/sample geometry of the VM (2^(n-5) lanes per sample), including partially filled warps."""
    check_against_oracle(qris_b200.reference_circuit(n, 2).pinned("wide"), B, seed=n * 7 + B)
    check_against_oracle(qris_b200.u3cu3_circuit(n, 1).pinned("wide"), B, seed=n * 11 + B)


@pytest.mark.parametrize("n", [3, 4, 5, 6, 7, 8])
@pytest.mark.parametrize("B", [1, 3, 256, 4096])
def test_narrow_variant_parity(n, B):
    """8 amplitudes per thread (a warp carries 256): the variant small batches run on, at the reference's own
    sizes (n = 6, B = 256 per call, Run:341-352; B = 4096 is BASELINE configs[1]) and for 3 and 4 qubits, which
    only this variant takes.  Plain pair and saved-state pair against the oracle."""
    for circ in (qris_b200.reference_circuit(n, 3).pinned("narrow"), qris_b200.u3cu3_circuit(n, 2).pinned("narrow")):
        assert _plan(circ, B) == "vm-register-narrow"
        check_against_oracle(circ, B, seed=n * 13 + B)
        x, w, gy = (t.cuda() for t in seeded_inputs(circ, B, seed=n + B))
        y1, state = qris_b200.qnn_forward_raw(circ, x, w, save_state=True)
        gx0, gw0 = qris_b200.qnn_backward_raw(circ, x, w, gy)
        gx1, gw1 = qris_b200.qnn_backward_raw(circ, x, w, gy, state=state)
        assert state is not None and torch.equal(y1, qris_b200.qnn_forward_raw(circ, x, w))
        assert_close(gx1.cpu().numpy(), gx0.cpu().numpy(), scale=float(gx0.abs().max()), what="gx saved vs recompute")
        assert_close(gw1.cpu().numpy(), gw0.cpu().numpy(), scale=float(gw0.abs().max()), what="gw saved vs recompute")


def test_variant_is_chosen_by_batch_size_and_variants_agree():
    """Unpinned circuits of up to 8 qubits pick the narrow variant for batches that would leave warp slots of the
    wide one empty, the wide one for large batches; both give the same numbers."""
    circ = qris_b200.reference_circuit(6, 3)
    assert _plan(circ, 256) == "vm-register-narrow" and _plan(circ, 4096) == "vm-register-narrow"
    assert _plan(circ, 65536) == "vm-register"
    assert _plan(qris_b200.reference_circuit(10, 3), 256) == "vm-register"          # no narrow variant above 8 qubits
    x, w, gy = seeded_inputs(circ, 1000, seed=3)
    a, b = run_gpu(circ.pinned("wide"), x, w, gy), run_gpu(circ.pinned("narrow"), x, w, gy)
    for u, v, name in zip(a, b, ("y", "gx", "gw")):
        assert_close(u, v, rel=2e-5, scale=float(np.abs(v).max()), what=name)


@pytest.mark.parametrize("n,B", [(5, 70), (8, 33), (9, 5), (10, 1), (10, 37), (11, 6), (12, 3)])
def test_sixteen_amplitudes_per_thread_variant(n, B):
    """The occupancy experiment (QNN_CREATE_16_PER_THREAD): half the state per thread, twice the warps; n = 10 then
    spans two warps per sample.  Same results as the oracle."""
    for circ in (qris_b200.u3cu3_circuit(n, 2).pinned("16-per-thread"), qris_b200.reference_circuit(n, 2).pinned("16-per-thread")):
        assert _plan(circ, B) == "vm-register-16"
        check_against_oracle(circ, B, seed=n * 5 + B)
        x, w, gy = (t.cuda() for t in seeded_inputs(circ, B, seed=n))
        _, state = qris_b200.qnn_forward_raw(circ, x, w, save_state=True)
        gx0, gw0 = qris_b200.qnn_backward_raw(circ, x, w, gy)
        gx1, gw1 = qris_b200.qnn_backward_raw(circ, x, w, gy, state=state)
        assert_close(gw1.cpu().numpy(), gw0.cpu().numpy(), scale=float(gw0.abs().max()), what="gw saved vs recompute")


@pytest.mark.parametrize("n", [11, 12, 13])
@pytest.mark.parametrize("B", [1, 5, 40])
def test_vm_multi_warp_samples(n, B):
    """n = 11..13 on the VM: one sample per 2^(n-10) warps, group-wide exchanges and cross-warp sums."""
    assert _plan(qris_b200.u3cu3_circuit(n, 1), B) == "vm-register"
    check_against_oracle(qris_b200.u3cu3_circuit(n, 1), B, seed=n + B)
    check_against_oracle(qris_b200.reference_circuit(n, 2), B, seed=n * 3 + B)


def test_vm_irregular_circuit():
    Op = qris_b200.Op
    n = 7
    ops = [Op("RY", (i,), (("x", i),)) for i in range(n)]
    ops += [Op("CNOT", (0, 6)), Op("RZ", (0,), (("w", 0),)), Op("RZ", (6,), (("w", 1),)), Op("RX", (6,), (("w", 2),)),
            Op("CZ", (1, 5)), Op("CNOT", (5, 1)), Op("CNOT", (1, 5)), Op("U3", (5,), (("w", 3), ("c", 0.4), ("w", 4))),
            Op("CU3", (6, 2), (("w", 5), ("w", 6), ("w", 7))), Op("CNOT", (2, 3)), Op("RY", (3,), (("x", 0),)),
            Op("CZ", (3, 4)), Op("RZ", (4,), (("c", 1.3),)), Op("CNOT", (4, 0)), Op("CNOT", (0, 4)), Op("RY", (0,), (("w", 8),)),
            Op("RX", (2,), (("w", 2),)), Op("CU3", (0, 1), (("c", 0.2), ("w", 6), ("c", -0.7)))]
    circ = qris_b200.Circuit(n, ops, n_inputs=n, n_weights=9)
    assert _plan(circ, 10).startswith("vm-register")
    for variant in ("wide", "narrow", "16-per-thread"):
        check_against_oracle(circ.pinned(variant), 45, seed=2)


@pytest.mark.parametrize("make,B", [(lambda: qris_b200.u3cu3_circuit(10, 2), 100), (lambda: qris_b200.reference_circuit(6, 3), 300),
                                     (lambda: qris_b200.reference_circuit(7, 2), 33), (lambda: qris_b200.reference_circuit(4, 2), 20),
                                     (lambda: qris_b200.reference_circuit(2, 2), 20), (lambda: qris_b200.reference_circuit(6, 3).pinned("wide"), 300)])
def test_backward_from_saved_state_matches_recompute(make, B):
    """qnn_forward_save / qnn_backward_saved (the autograd pair) against qnn_forward / qnn_backward."""
    circ = make()
    x, w, gy = (t.cuda() for t in seeded_inputs(circ, B, seed=11))
    y0 = qris_b200.qnn_forward_raw(circ, x, w)
    y1, state = qris_b200.qnn_forward_raw(circ, x, w, save_state=True)
    assert torch.equal(y0, y1)
    lib = qris_b200._lib.lib()
    nbytes = lib.qnn_state_bytes(circ.handle(0), B)
    if circ.n_qubits < 3:
        assert state is None and nbytes == 0          # generic kernels recompute
    else:
        assert state is not None and state.numel() * 4 == nbytes and nbytes >= B * (8 << circ.n_qubits)
    gx0, gw0 = qris_b200.qnn_backward_raw(circ, x, w, gy)
    gx1, gw1 = qris_b200.qnn_backward_raw(circ, x, w, gy, state=state)
    assert_close(gx1.cpu().numpy(), gx0.cpu().numpy(), scale=float(gx0.abs().max()), what="gx")
    assert_close(gw1.cpu().numpy(), gw0.cpu().numpy(), scale=float(gw0.abs().max()), what="gw")
