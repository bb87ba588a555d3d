#!/usr/bin/env python
"""Benchmark of the QNN hot path: circuit-samples/sec, forward + backward.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--workload NAME]

Workload (default ``cfg3``, BASELINE.json configs[2], the configuration the metric "samples/sec at
1/2/4/8 B200, % FP32 roofline" is quoted on): 10-qubit U3+CU3 circuit, 3 layers, RY angle encoder,
QuantumNAT weight noise (sigma 0.01) + post-measurement normalise / noise / 5-level quantise, batch
65,536 PER GPU (weak scaling; at N=1 this is exactly BASELINE's batch).  ``cfg2`` is configs[1] (the
reference's n=6, L=3 circuit at batch 4096) -- launch-bound by construction.

One step = weight noise -> qnn_forward -> QuantumNAT forward -> QuantumNAT backward -> qnn_backward
(adjoint VJP: gx and batch-reduced gw) [-> one NCCL all-reduce of gw when N > 1], on inputs resident in
HBM (`value`); `e2e` is the same work driven through the nn.Module surface with pinned-host inputs
copied in and the loss read back every step.  Prints ONE JSON line (rank 0).
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

FP32_PEAK_NOMINAL_TFLOPS = 148 * 128 * 2 * 1.965e9 / 1e12   # 74.45; SURVEY.md 8(d), BASELINE.md 4

WORKLOADS = {
    # name: (circuit kind, n_qubits, n_layers, per-GPU batch, QuantumNAT pipeline, description)
    "cfg3": ("u3cu3", 10, 3, 65536, True,
             "BASELINE configs[2]: 10-qubit U3+CU3 L=3, QuantumNAT weight noise + normalise/noise/5-level quantise, batch 65536 per GPU"),
    "cfg2": ("ref", 6, 3, 4096, False,
             "BASELINE configs[1]: reference RY/RZ/CNOT circuit n=6 L=3 (P=128 pipeline's QNN), batch 4096 per GPU"),
    "ref10": ("ref", 10, 3, 65536, False, "reference RY/RZ/CNOT circuit at n=10 L=3, batch 65536 per GPU"),
    "cfg4": ("u3cu3", 16, 8, 576, False,
             "BASELINE configs[3]: 16-qubit U3+CU3 L=8 (a sample spans a cluster of 8 CTAs), adjoint backward, batch 576 per GPU"),
}


# ---- algorithmic work (flops of the circuit as the reference executes it) -----------
def algorithmic_flops(ops, n):
    """(forward, backward) real flops per sample; FMA = 2.  Per amplitude: RX/RY/RZ 6, U3 14, CU3 7,
    CNOT/CZ 0; measurement 3+n; cotangent seeding n+2; one complex inner product per angle: 8 (4 for a
    controlled gate, which touches half the amplitudes)."""
    per_amp = {"RX": 6, "RY": 6, "RZ": 6, "U3": 14, "CU3": 7, "CNOT": 0, "CZ": 0}
    N = 1 << n
    gates = sum(per_amp[k] for k, _, _ in ops)
    inner = sum((4 if k == "CU3" else 8) * len(p) for k, _, p in ops)
    fwd = gates * N + (3 + n) * N
    bwd = (n + 2) * N + 2 * gates * N + inner * N
    return fwd, bwd


def hbm_bytes_per_sample(n, saved_state=True):
    """fwd: read x, write y; bwd: read x, gy, write gx (4 bytes each, n columns); plus, when the adjoint
    pass starts from the saved final state, one write and one read of 2^n complex64 amplitudes."""
    return 20 * n + (2 * 8 * (1 << n) if saved_state else 0)


# ---- clocks ---------------------------------------------------------------------------------------
class ClockSampler:
    FIELDS = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc, self.marks = index, [], None, []

    def __enter__(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits", "-lms", "100",
                 "-i", str(self.index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            parts = [p.strip() for p in line.split(",")]
            if len(parts) >= 7:
                self.rows.append((time.time(), parts))

    def mark(self):
        """Call at the start and at the end of the timed region."""
        self.marks.append(time.time())

    def __exit__(self, *exc):
        if self.proc is not None:
            time.sleep(0.15)
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:
                self.proc.kill()

    def summary(self):
        sm, mx, reasons = [], [], set()
        rows = self.rows
        window = "timed region"
        if len(self.marks) >= 2:
            inside = [r for r in rows if self.marks[0] <= r[0] <= self.marks[-1] + 0.1]
            if inside:
                rows = inside
            else:   # region shorter than the sampling period: use the samples taken under the same load (warm-up)
                window = "warm-up + timed region (timed region shorter than the sampling period)"
        for _, r in rows:
            try:
                sm.append(float(r[0])); mx.append(float(r[1]))
            except ValueError:
                continue
            for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[3:7]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"], "samples": 0}
        return {"sm_mhz": statistics.median(sm), "sm_max_mhz": max(mx), "reasons": sorted(reasons), "samples": len(sm),
                "window": window}


# ---- CPU arm: the reference's execution model restated (oracle/, complex64, all host cores) ----------
def cpu_step_fn(kind, n, L, qnat, B, seed=0):
    """One fwd+bwd of the workload on the CPU the way default.qubit + torch backprop runs it
    (Estimators_QuantumNAT_onchipQNN.py:148: one batched tensor op per gate, autograd backward)."""
    import torch
    from oracle import qnat_oracle as QO
    from oracle import qnn_oracle as O
    ops = O.ref_circuit_ops(n, L) if kind == "ref" else O.u3cu3_circuit_ops(n, L)
    g = torch.Generator().manual_seed(seed)
    x = (2 * torch.rand(B, n, generator=g) - 1).float()
    w = (6.2831853 * torch.rand(O.n_weights_of(ops), generator=g)).float()
    gout = torch.randn(B, n, generator=g).float()

    def step():
        xr = x.clone().requires_grad_(True)
        wr = (QO.weight_noise(w, 0.01) if qnat else w.clone()).requires_grad_(True)
        y = O.simulate(ops, n, xr, wr, cdtype=torch.complex64)
        if qnat:
            noise = torch.randn(B, n)
            out, yn, mean, rstd = QO.qnat_forward(y.detach(), noise, sigma=0.01, dtype=torch.float32)
            gy = QO.qnat_backward(gout, y.detach(), yn, mean, rstd, dtype=torch.float32)
        else:
            gy = gout
        y.backward(gy)
        return xr.grad, wr.grad
    return step


def pick_threads(kind, n, L, qnat):
    """All host threads, unless oversubscription makes that slower (containers with a CPU quota below
    the visible core count): probe cpu_count and cpu_count/2 and keep the faster.  Reported as `cores`."""
    import torch
    ncpu = os.cpu_count() or 1
    best, best_t = ncpu, None
    for nt in sorted({ncpu, max(1, ncpu // 2)}, reverse=True):
        torch.set_num_threads(nt)
        step = cpu_step_fn(kind, n, L, qnat, 64)
        step()
        ts = []
        for _ in range(2):
            t0 = time.perf_counter(); step(); ts.append(time.perf_counter() - t0)
        if best_t is None or min(ts) < best_t:
            best, best_t = nt, min(ts)
    torch.set_num_threads(best)
    return best, best_t / 64


def time_cpu(kind, n, L, qnat, target_s=12.0, max_B=4096, steps=3):
    """Bounded sample: size B so that one step takes about target_s/4, then time `steps` steps."""
    import torch
    threads, per_sample = pick_threads(kind, n, L, qnat)
    B = int(min(max_B, max(32, (target_s / 4) / max(per_sample, 1e-6))))
    B = 1 << (B.bit_length() - 1)
    step = cpu_step_fn(kind, n, L, qnat, B)
    step()
    ts = []
    for _ in range(steps):
        t0 = time.perf_counter(); step(); ts.append(time.perf_counter() - t0)
    med = statistics.median(ts)
    return {"value": B / med, "unit": "samples/s", "cores": threads, "kind": "port", "batch": B,
            "sample": f"B={B} samples of the same workload per step (complex64 torch, one tensor op per gate, "
                      f"autograd backward), median of {steps} steps of {med:.2f} s"}


def run_reference(args, wl):
    kind, n, L, B_gpu, qnat, desc = WORKLOADS[wl]
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    import torch
    # bounded sample per step so that warmup + steps end within minutes
    threads, per_sample = pick_threads(kind, n, L, qnat)
    budget_s = 150.0 / max(1, args.steps + args.warmup)
    B = int(min(4096, max(32, min(budget_s, 4.0) / max(per_sample, 1e-6))))
    B = 1 << (B.bit_length() - 1)
    step = cpu_step_fn(kind, n, L, qnat, B)
    for _ in range(args.warmup):
        step()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step()
    dt = (time.perf_counter() - t0) / args.steps
    value = B / dt
    line = {
        "impl": "reference", "metric": "qnn_samples_per_sec_fwd_bwd", "value": value, "unit": "samples/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32 (complex64 state)",
        "data": "synthetic", "config": {"workload": desc, "cpu_sample_batch": B},
        "cpu_baseline": {"value": value, "unit": "samples/s", "cores": torch.get_num_threads(), "kind": "port",
                         "sample": f"B={B} per step; PennyLane (the reference's simulator) is not installed, so the "
                                   "oracle's restatement of its execution model is timed"},
        "e2e": {"value": value, "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)
    return 0


# ---- GPU arm ----------------------------------------------------------------------------------------
def run_b200(args, wl):
    import torch
    import torch.distributed as dist
    import torch.nn.functional as F

    import qris_b200
    from qris_b200.parallel import FlatGradBucket

    kind, n, L, B, qnat, desc = WORKLOADS[wl]
    if args.batch:
        B = args.batch
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the product has no CPU path; use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    lib = qris_b200._lib.lib()

    circ = qris_b200.reference_circuit(n, L) if kind == "ref" else qris_b200.u3cu3_circuit(n, L)
    ops = [(o.kind, o.wires, o.params) for o in circ.ops]
    f_fwd, f_bwd = algorithmic_flops(ops, n)
    torch.manual_seed(1234 + rank)
    x = (2 * torch.rand(B, n, device=dev) - 1)
    gout = torch.randn(B, n, device=dev)
    torch.manual_seed(99)
    w = 2 * torch.pi * torch.rand(circ.n_weights, device=dev)
    gw_bucket = torch.zeros(circ.n_weights, device=dev)
    head = qris_b200.QuantumNATHead(sigma=0.01, lo=-2.0, hi=2.0, levels=5).to(dev).train() if qnat else None
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)     # > 126 MB L2
    h = circ.handle(local_rank)
    launches_per_step = lib.qnn_forward_launches(h, B) + lib.qnn_backward_launches(h, B) + (6 if qnat else 0)

    def step():
        wn = w + 0.01 * torch.randn_like(w) if qnat else w           # weight-space noise (Est:186-188)
        y, state = qris_b200.qnn_forward_raw(circ, x, wn, save_state=True)   # keeps psi_final for the adjoint pass
        if qnat:
            yq = y.detach().requires_grad_(True)
            out = head(yq)
            out.backward(gout)
            gy = yq.grad
        else:
            gy = gout
        gx, _ = qris_b200.qnn_backward_raw(circ, x, wn, gy, need_gx=True, gw_out=gw_bucket, state=state)
        if world > 1:
            dist.all_reduce(gw_bucket)
        return gx

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    with ClockSampler(local_rank) as clocks:                          # started early: nvidia-smi needs a moment
        for _ in range(max(args.warmup, 3)):
            step()
        barrier()
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
        barrier()
        clocks.mark()
        for s, e in ev:
            flush.zero_()                                             # L2 flush, outside the timed events
            s.record(); step(); e.record()
        barrier()
        clocks.mark()
        step_ms = [s.elapsed_time(e) for s, e in ev]
        total_ms = sum(step_ms)
        # dominant kernel (adjoint backward) and the forward kernel, timed alone on the same stream
        gy_fixed = torch.randn(B, n, device=dev)
        k_bwd, k_fwd = [], []
        for _ in range(max(5, min(args.steps, 20))):
            flush.zero_()
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s2, e2 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s2.record(); _, st_k = qris_b200.qnn_forward_raw(circ, x, w, save_state=True); e2.record()
            s.record(); qris_b200.qnn_backward_raw(circ, x, w, gy_fixed, need_gx=True, gw_out=gw_bucket, state=st_k); e.record()
            torch.cuda.synchronize()
            k_bwd.append(s.elapsed_time(e)); k_fwd.append(s2.elapsed_time(e2))
    t = torch.tensor([total_ms], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms = t.item()
    ms_per_step = total_ms / args.steps
    value = B * world / (ms_per_step * 1e-3)

    # ---- end to end through the module surface: pinned host -> device, loss read back ----
    torch.manual_seed(5)
    layer = qris_b200.QNNLayer(circ).to(dev)
    clf = torch.nn.Linear(n, 3).to(dev)
    params = list(layer.parameters()) + list(clf.parameters())
    bucket = FlatGradBucket(params)
    xh = (2 * torch.rand(B, n) - 1).pin_memory()
    lh = torch.randint(0, 3, (B,)).pin_memory()
    e2e_launches = launches_per_step

    def e2e_step():
        bucket.zero()
        xd = xh.to(dev, non_blocking=True).requires_grad_(True)
        ld = lh.to(dev, non_blocking=True)
        if qnat:                                                      # Est:176-196 on the layer's own weights
            with torch.no_grad():
                saved = layer.weights.data.clone()
                layer.weights.data.add_(0.01 * torch.randn_like(layer.weights))
        y = layer(xd)
        if qnat:
            with torch.no_grad():
                layer.weights.data.copy_(saved)
            y = head(y)
        loss = F.nll_loss(F.log_softmax(clf(y), dim=1), ld)
        loss.backward()
        bucket.all_reduce()
        return loss.item()                                            # device -> host read every step

    for _ in range(max(args.warmup, 3)):
        e2e_step()
    barrier()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    for s, e in ev:
        flush.zero_()
        s.record(); e2e_step(); e.record()
    barrier()
    t = torch.tensor([sum(s.elapsed_time(e) for s, e in ev)], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_value = B * world / (t.item() / args.steps * 1e-3)

    if rank == 0:
        tf = 0.0
        try:
            import ctypes
            val, unused = ctypes.c_double(), ctypes.c_double()
            if lib.qnn_fp32_peak_probe(4096, ctypes.byref(val), ctypes.byref(unused)) == 0:
                tf = val.value
        except Exception:
            pass
        bwd_ms, fwd_ms = statistics.mean(k_bwd), statistics.mean(k_fwd)
        achieved = f_bwd * B / (bwd_ms * 1e-3) / 1e12
        traffic = None
        tpath = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(tpath):
            try:
                traffic = json.load(open(tpath)).get(wl, {}).get("bwd_dram_bytes_per_launch")
            except Exception:
                traffic = None
        cpu = time_cpu(kind, n, L, qnat) if (world == 1 and not args.no_cpu) else None
        line = {
            "metric": "qnn_samples_per_sec_fwd_bwd", "value": value, "unit": "samples/s", "n_gpus": world,
            "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms_per_step,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32 (complex64 state)",
            "data": "synthetic",
            "config": {"workload": desc, "name": wl, "n_qubits": n, "n_layers": L, "batch_per_gpu": B,
                       "global_batch": B * world, "parallelism": f"dp{world}",
                       "l2": "flushed between timed steps (256 MiB memset, outside the per-step CUDA events)",
                       "plan": lib.qnn_circuit_plan(h, B).decode()},
            "clocks": clocks.summary(),
            "e2e": {"value": e2e_value, "unit": "samples/s", "h2d_bytes_per_step": xh.numel() * 4 + lh.numel() * 8,
                    "d2h_bytes_per_step": 4},
            "gpu_launches": launches_per_step * args.steps + e2e_launches * args.steps,
            "roofline": {
                "bound": "fp32", "kernel": "qnn_backward_saved (adjoint VJP from the saved final state: uncompute + gradients)",
                "achieved": achieved, "peak": FP32_PEAK_NOMINAL_TFLOPS, "unit": "TFLOP/s",
                "frac": achieved / FP32_PEAK_NOMINAL_TFLOPS,
                "peak_source": "nominal 148 SM x 128 lanes x 2 x 1.965 GHz (MEASURED_PEAKS.json has no FP32 entry)",
                "peak_measured": tf, "frac_of_measured": (achieved / tf) if tf else None,
                "traffic": traffic, "algorithmic_flops_per_sample": {"fwd": f_fwd, "bwd": f_bwd},
                "kernel_ms": {"bwd": bwd_ms, "fwd": fwd_ms},
                "fwd_achieved": f_fwd * B / (fwd_ms * 1e-3) / 1e12,
                "step_achieved": (f_fwd + f_bwd) * B / (ms_per_step * 1e-3) / 1e12,
                "step_frac": (f_fwd + f_bwd) * B / (ms_per_step * 1e-3) / 1e12 / FP32_PEAK_NOMINAL_TFLOPS,
                "hbm_gbs": hbm_bytes_per_sample(n) * B / (ms_per_step * 1e-3) / 1e9,
            },
        }
        if cpu is not None:
            line["cpu_baseline"] = cpu
        emit(line)
    if world > 1:
        dist.destroy_process_group()
    return 0


_RESULT_OUT = None


def emit(line):
    out = _RESULT_OUT if _RESULT_OUT is not None else sys.stdout
    out.write(json.dumps(line) + "\n")
    out.flush()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="cfg3", choices=sorted(WORKLOADS))
    ap.add_argument("--batch", type=int, default=0, help="override the per-GPU batch")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    args = ap.parse_args()
    # The contract is ONE JSON line on stdout.  Libraries write there too (NCCL prints its version line at the
    # first communicator), so everything but the result line is sent to stderr: fd 1 is pointed at fd 2 for
    # the run and the JSON line goes to a saved copy of the original stdout.
    global _RESULT_OUT
    sys.stdout.flush()
    _RESULT_OUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    if args.impl == "reference":
        return run_reference(args, args.workload)
    return run_b200(args, args.workload)


if __name__ == "__main__":
    sys.exit(main())
