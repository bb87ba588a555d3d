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
copied in and the loss read back every step.  Prints ONE JSON line (rank 0).  The device-timed step (and the
forward / backward calls timed alone for the roofline block) is recorded once as a CUDA graph and replayed, so that
small workloads measure the kernels and not the host's per-call cost (`--eager` launches every step from Python
instead; the all-reduce of a multi-GPU step always stays outside the recording).  The e2e step is the user's call:
`qris_b200.GraphedCall` around the module's forward + backward, with the pinned-host copies, the all-reduce and the
loss read outside it.

Besides the headline workload the line carries a ``secondary`` list of short device-timed records, each with
its own roofline block, so that every number quoted in README/DESIGN comes out of a driver run: ``cfg4``
(BASELINE configs[3]: 16 qubits, 8 layers), ``cfg2`` (configs[1]: n=6, batch 4096), ``cfg1`` (configs[0], the
as-shipped training step: nine forwards of 256 + one backward + optimiser through ``runner.qsc_step``, eager and
as a replayed CUDA graph), and for N > 1 ``cfg3_strong`` (BASELINE's batch of 65,536 split over the N GPUs).

CPU legs (``cpu_baseline`` and ``--impl reference``) time ``oracle/``'s restatement of the reference's execution
model (PennyLane's default.qubit + torch backprop: one tensor op per gate, autograd backward) with one fixed
sample batch per workload (``CPU_BATCH``).  ``import pennylane`` is probed and the outcome reported in
``cpu_baseline.pennylane``; it is not installed in this image, so the reference's own simulator is never timed.
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


# One CPU sample batch per workload, used by BOTH CPU legs (cpu_baseline inside the default run and the
# --impl reference arm), so the two figures of one box are the same measurement.
CPU_BATCH = {"cfg3": 2048, "cfg2": 4096, "ref10": 2048, "cfg4": 8}


def pennylane_probe():
    try:
        import pennylane  # noqa: F401
        return f"installed ({getattr(pennylane, '__version__', '?')}) but not used: this arm times the oracle's restatement"
    except Exception as e:
        return f"not installed ({type(e).__name__}); the oracle's restatement of its execution model is timed"


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
        if int(os.environ.get("RANK", "0")) != 0:      # one sampler per job: rank 0's GPU stands for the box
            return self
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


def time_cpu(wl, steps=3, warmup=1):
    """Bounded sample: CPU_BATCH[wl] samples per step, `steps` timed steps after `warmup`."""
    kind, n, L, _, qnat, _ = WORKLOADS[wl]
    threads, _ = pick_threads(kind, n, L, qnat)
    B = CPU_BATCH[wl]
    step = cpu_step_fn(kind, n, L, qnat, B)
    for _ in range(warmup):
        step()
    ts = []
    for _ in range(steps):
        t0 = time.perf_counter(); step(); ts.append(time.perf_counter() - t0)
    med = statistics.median(ts)
    return {"value": B / med, "unit": "samples/s", "cores": threads, "kind": "port", "batch": B,
            "pennylane": pennylane_probe(),
            "sample": f"B={B} samples of the same workload per step (complex64 torch, one tensor op per gate, "
                      f"autograd backward), median of {steps} steps of {med:.2f} s"}, med


def run_reference(args, wl):
    kind, n, L, B_gpu, qnat, desc = WORKLOADS[wl]
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    # the same bounded sample per step as cpu_baseline; exactly the requested warm-up and step counts
    steps = max(1, args.steps)
    cpu, med = time_cpu(wl, steps=steps, warmup=max(0, args.warmup))
    value = cpu["value"]
    line = {
        "impl": "reference", "metric": "qnn_samples_per_sec_fwd_bwd", "value": value, "unit": "samples/s",
        "n_gpus": args.gpus, "steps": steps, "warmup": max(0, args.warmup), "ms_per_step": med * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32 (complex64 state)",
        "data": "synthetic", "config": {"workload": desc, "name": wl, "cpu_sample_batch": cpu["batch"]},
        "cpu_baseline": cpu,
        "e2e": {"value": value, "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)
    return 0


# ---- GPU arm ----------------------------------------------------------------------------------------
def traffic_record(wl):
    """DRAM bytes per launch of the dominant kernel as recorded from an `ncu --set full` capture
    (profiles/traffic.json): NOT measured in this run, and labelled so."""
    tpath = os.path.join(ROOT, "profiles", "traffic.json")
    try:
        rec = json.load(open(tpath)).get(wl)
    except Exception:
        rec = None
    if not rec:
        return None, "no ncu capture recorded for this workload"
    return rec.get("bwd_dram_bytes_per_launch"), "ncu --set full, " + rec.get("source", "profiles/traffic.json") + " (not measured in this run)"


class Ctx:
    """Per-process state shared by the workload measurements."""

    def __init__(self):
        import torch
        import torch.distributed as dist
        self.torch, self.dist = torch, dist
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.rank = int(os.environ.get("RANK", "0"))
        self.local_rank = int(os.environ.get("LOCAL_RANK", "0"))
        if not torch.cuda.is_available():
            raise SystemExit("bench.py: no CUDA device (the product has no CPU path; use --impl reference for the CPU arm)")
        torch.cuda.set_device(self.local_rank)
        self.dev = torch.device("cuda", self.local_rank)
        if self.world > 1:
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            dist.init_process_group("nccl", device_id=self.dev)
        self.flush = torch.empty(256 << 20, dtype=torch.uint8, device=self.dev)     # > 126 MB L2
        self.eager = False

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def max_over_ranks(self, ms):
        t = self.torch.tensor([ms], device=self.dev, dtype=self.torch.float64)
        if self.world > 1:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return t.item()


def record_graph(torch, fn):
    """Record `fn` (kernel launches on static tensors) once into a CUDA graph and return its replay; the eager
    function itself if recording is not possible.  Launch-bound steps (small batches: a few kernels of tens of
    microseconds) are then bound by the device, not by the host's per-call cost."""
    try:
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(2):
                fn()
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=side):
            fn()
        g.replay()
        torch.cuda.synchronize()
        return g.replay, "recorded once as a CUDA graph and replayed"
    except Exception as e:                                     # keep measuring, and say so in the record
        torch.cuda.synchronize()
        return fn, f"eager (graph recording failed: {type(e).__name__})"


def measure_workload(ctx, wl, B, steps, warmup, clocks=None):
    """Device-timed fwd+bwd step of one workload on inputs resident in HBM; returns the record pieces."""
    import qris_b200
    torch, dist, dev = ctx.torch, ctx.dist, ctx.dev
    kind, n, L, _, qnat, desc = WORKLOADS[wl]
    lib = qris_b200._lib.lib()
    circ = qris_b200.reference_circuit(n, L) if kind == "ref" else qris_b200.u3cu3_circuit(n, L)
    ops = [(o.kind, o.wires, o.params) for o in circ.ops]
    f_fwd, f_bwd = algorithmic_flops(ops, n)
    torch.manual_seed(1234 + ctx.rank)
    x = (2 * torch.rand(B, n, device=dev) - 1)
    gout = torch.randn(B, n, device=dev)
    torch.manual_seed(99)
    w = 2 * torch.pi * torch.rand(circ.n_weights, device=dev)
    gw_bucket = torch.zeros(circ.n_weights, device=dev)
    head = qris_b200.QuantumNATHead(sigma=0.01, lo=-2.0, hi=2.0, levels=5).to(dev).train() if qnat else None
    h = circ.handle(ctx.local_rank)
    launches_per_step = lib.qnn_forward_launches(h, B) + lib.qnn_backward_launches(h, B) + (6 if qnat else 0)

    def compute():
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
        return gx

    replay, launch_mode = record_graph(torch, compute) if not ctx.eager else (compute, "eager (--eager)")

    def step():
        with torch.cuda.nvtx.range("bench/step"):
            replay()
            if ctx.world > 1:                                    # the collective stays outside the recording
                with torch.cuda.nvtx.range("allreduce/gw"):
                    dist.all_reduce(gw_bucket)

    for _ in range(max(warmup, 3)):
        step()
    ctx.barrier()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    ctx.barrier()
    if clocks is not None:
        clocks.mark()
    for s_, e_ in ev:
        ctx.flush.zero_()                                             # L2 flush, outside the timed events
        s_.record(); step(); e_.record()
    ctx.barrier()
    if clocks is not None:
        clocks.mark()
    total_ms = ctx.max_over_ranks(sum(s_.elapsed_time(e_) for s_, e_ in ev))
    ms_per_step = total_ms / steps
    # dominant kernel (adjoint backward) and the forward kernel, timed alone on the same stream
    gy_fixed = torch.randn(B, n, device=dev)
    keep = {}

    def fwd_only():
        keep["state"] = qris_b200.qnn_forward_raw(circ, x, w, save_state=True)[1]

    def bwd_only():
        qris_b200.qnn_backward_raw(circ, x, w, gy_fixed, need_gx=True, gw_out=gw_bucket, state=keep["state"])
    fwd_replay = record_graph(torch, fwd_only)[0] if not ctx.eager else fwd_only
    bwd_replay = record_graph(torch, bwd_only)[0] if not ctx.eager else bwd_only
    k_bwd, k_fwd = [], []
    for _ in range(max(5, min(steps, 20))):
        ctx.flush.zero_()
        s1, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s2, e2 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s2.record(); fwd_replay(); e2.record()
        s1.record(); bwd_replay(); e1.record()
        torch.cuda.synchronize()
        k_bwd.append(s1.elapsed_time(e1)); k_fwd.append(s2.elapsed_time(e2))
    bwd_ms, fwd_ms = statistics.mean(k_bwd), statistics.mean(k_fwd)
    achieved = f_bwd * B / (bwd_ms * 1e-3) / 1e12
    traffic, traffic_source = traffic_record(wl)
    step_tf = (f_fwd + f_bwd) * B / (ms_per_step * 1e-3) / 1e12
    roofline = {
        "bound": "fp32", "kernel": "qnn_backward_saved (adjoint VJP from the saved final state: uncompute + gradients)",
        "achieved": achieved, "peak": FP32_PEAK_NOMINAL_TFLOPS, "unit": "TFLOP/s",
        "frac": achieved / FP32_PEAK_NOMINAL_TFLOPS,
        "peak_source": "nominal 148 SM x 128 lanes x 2 x 1.965 GHz (MEASURED_PEAKS.json has no FP32 entry)",
        "traffic": traffic, "traffic_source": traffic_source,
        "algorithmic_flops_per_sample": {"fwd": f_fwd, "bwd": f_bwd},
        "kernel_ms": {"bwd": bwd_ms, "fwd": fwd_ms},
        "fwd_achieved": f_fwd * B / (fwd_ms * 1e-3) / 1e12,
        "step_achieved": step_tf, "step_frac": step_tf / FP32_PEAK_NOMINAL_TFLOPS,
        "hbm_gbs": hbm_bytes_per_sample(n) * B / (ms_per_step * 1e-3) / 1e9,
        "saved_state_bytes_per_outstanding_forward": int(lib.qnn_state_bytes(h, B)),
    }
    return {"circ": circ, "value": B * ctx.world / (ms_per_step * 1e-3), "ms_per_step": ms_per_step, "roofline": roofline,
            "launches_per_step": launches_per_step, "head": head, "desc": desc, "n": n, "L": L, "launch": launch_mode,
            "plan": lib.qnn_circuit_plan(h, B).decode()}


def measure_e2e(ctx, m, B, qnat, steps, warmup):
    """The same metric end to end through the module surface (the call a user makes): every step copies its inputs
    from pinned host memory, runs QNNLayer (+ QuantumNAT head) + a linear classifier + loss forward and backward
    into the flat gradient bucket, all-reduces it, and reads the loss back."""
    import torch.nn.functional as F

    import qris_b200
    from qris_b200.parallel import FlatGradBucket
    torch, dev, n = ctx.torch, ctx.dev, m["n"]
    circ, head = m["circ"], m["head"]
    torch.manual_seed(5)
    layer = qris_b200.QNNLayer(circ).to(dev)
    clf = torch.nn.Linear(n, 3).to(dev)
    params = list(layer.parameters()) + list(clf.parameters())
    bucket = FlatGradBucket(params)
    xh = (2 * torch.rand(B, n) - 1).pin_memory()
    lh = torch.randint(0, 3, (B,)).pin_memory()

    def core(xd, ld):
        bucket.zero()
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
        return loss.detach()

    # the user's call: qris_b200.GraphedCall records forward + backward once and replays it (the host -> device
    # copies, the all-reduce and the loss read stay outside the recording); --eager runs the plain module calls
    x0 = torch.zeros(B, n, device=dev, requires_grad=True)
    l0 = torch.zeros(B, dtype=torch.long, device=dev)
    graphed = None if ctx.eager else qris_b200.GraphedCall(core, (x0, l0))

    def e2e_step():
        if graphed is not None:
            loss = graphed(xh, lh)                                    # pinned host -> the recording's input buffers
        else:
            loss = core(xh.to(dev, non_blocking=True).requires_grad_(True), lh.to(dev, non_blocking=True))
        bucket.all_reduce()
        return loss.item()                                            # device -> host read every step

    for _ in range(max(warmup, 3) + 2):                               # (three eager calls, the recording, a replay)
        e2e_step()
    ctx.barrier()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    for s_, e_ in ev:
        ctx.flush.zero_()
        s_.record(); e2e_step(); e_.record()
    ctx.barrier()
    e2e_ms = ctx.max_over_ranks(sum(s_.elapsed_time(e_) for s_, e_ in ev))
    return {"value": B * ctx.world / (e2e_ms / steps * 1e-3), "unit": "samples/s",
            "h2d_bytes_per_step": xh.numel() * 4 + lh.numel() * 8, "d2h_bytes_per_step": 4,
            "launch": "plain module calls (--eager)" if graphed is None else
                      "qris_b200.GraphedCall: forward + backward recorded once and replayed; copies, all-reduce and loss read outside"}


def secondary_record(ctx, wl, B, steps, scaling, name=None, with_e2e=False):
    m = measure_workload(ctx, wl, B, steps, 3)
    r = m["roofline"]
    extra = {"e2e": measure_e2e(ctx, m, B, WORKLOADS[wl][4], steps, 3)} if with_e2e else {}
    return {**extra, "name": name or wl, "workload": m["desc"], "metric": "qnn_samples_per_sec_fwd_bwd", "value": m["value"],
            "unit": "samples/s", "n_gpus": ctx.world, "steps": steps, "ms_per_step": m["ms_per_step"], "scaling": scaling,
            "batch_per_gpu": B, "global_batch": B * ctx.world, "plan": m["plan"], "launch": m["launch"],
            "roofline": {k: r[k] for k in ("bound", "achieved", "peak", "unit", "frac", "step_frac", "step_achieved", "kernel_ms",
                                           "algorithmic_flops_per_sample", "traffic", "traffic_source")}}


def cfg1_record(ctx, steps=30):
    """BASELINE configs[0], the step the reference ships (Run:341-369): nine forwards of 256 samples (n=6, L=3),
    one backward, pruning off, AdamW -- through runner.qsc_step on device-resident synthetic data; eager, with
    the nine users concatenated, and each of those replayed as a CUDA graph.  Host-timed around a device
    synchronise (the step is host-bound when eager), single process."""
    import qris_b200
    from qris_b200 import estimators as E
    from qris_b200 import optim, runner
    torch = ctx.torch
    r = runner.Y2HRunner(data_len=2560, batch_size_DML=256, device=ctx.dev, verbose=False)
    loader, _ = r.get_dataloader_DML()
    batches = [b for b in loader if b[0][0][0].shape[0] == 256]
    circ = qris_b200.reference_circuit(6, 3)
    f_fwd, f_bwd = algorithmic_flops([(o.kind, o.wires, o.params) for o in circ.ops], 6)
    modes = {}
    for concat, graph in ((False, False), (True, False), (False, True), (True, True)):
        torch.manual_seed(0)
        model = E.QSC_P128(use_quantumnat=False, use_gradient_pruning=False).to(ctx.dev).train()
        opt = optim.FlatPrunedAdamW(model.parameters(), lr=1e-3, weight_decay=0.01, device_step_count=graph)
        step = runner.GraphedQSCStep(r, model, opt, concat) if graph else (lambda b: r.qsc_step(model, opt, b, concat))
        for b in batches[:6]:
            step(b)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for i in range(steps):
            step(batches[i % len(batches)])
        torch.cuda.synchronize()
        dt = (time.perf_counter() - t0) / steps
        key = ("graph" if graph else "eager") + ("_concat_users" if concat else "")
        sps = 9 * 256 / dt
        modes[key] = {"steps_per_s": 1.0 / dt, "samples_per_s": sps, "ms_per_step": dt * 1e3,
                      "qnn_frac_of_fp32_roofline": (f_fwd + f_bwd) * sps / 1e12 / FP32_PEAK_NOMINAL_TFLOPS}
    return {"name": "cfg1", "workload": "BASELINE configs[0]: the as-shipped training step (Run:341-369): QSC_P128 n=6 L=3, nine "
                                        "user batches of 256, one backward, AdamW; whole model (CNN front end + QNN + classifier + optimiser)",
            "metric": "qsc_optimiser_steps_per_sec", "unit": "steps/s", "n_gpus": 1, "steps": steps, "scaling": "none",
            # runner.train_QSC_P128's default for the as-shipped flags (no weight noise, no batch statistics): the nine
            # user batches as one model call, launched eagerly; "eager" is the nine separate calls of Run:341-352
            "value": modes["eager_concat_users"]["steps_per_s"], "default_mode": "eager_concat_users", "modes": modes,
            "roofline": {"bound": "launch/host (71.0 kflop per sample: SURVEY.md H5)", "unit": "TFLOP/s", "peak": FP32_PEAK_NOMINAL_TFLOPS,
                         "frac": modes["graph_concat_users"]["qnn_frac_of_fp32_roofline"],
                         "algorithmic_flops_per_sample": {"fwd": f_fwd, "bwd": f_bwd}}}


def run_b200(args, wl):
    import torch
    import torch.nn.functional as F

    import qris_b200
    from qris_b200.parallel import FlatGradBucket

    ctx = Ctx()
    ctx.eager = bool(args.eager)
    dist, dev, world, rank = ctx.dist, ctx.dev, ctx.world, ctx.rank
    kind, n, L, B, qnat, desc = WORKLOADS[wl]
    if args.batch:
        B = args.batch
    lib = qris_b200._lib.lib()
    with ClockSampler(ctx.local_rank) as clocks:                      # started early: nvidia-smi needs a moment
        m = measure_workload(ctx, wl, B, args.steps, args.warmup, clocks)
    circ, head, ms_per_step, value = m["circ"], m["head"], m["ms_per_step"], m["value"]
    launches_per_step = m["launches_per_step"]
    flush, barrier = ctx.flush, ctx.barrier

    e2e = measure_e2e(ctx, m, B, qnat, args.steps, args.warmup)
    e2e_value = e2e["value"]

    # ---- secondary records: every other number the documentation quotes (all ranks take part) ----
    secondary = []
    if not args.no_secondary and wl == "cfg3" and not args.batch:
        if world > 1:     # BASELINE configs[2] as stated: batch 65,536 GLOBAL, split over the N GPUs
            secondary.append(secondary_record(ctx, "cfg3", WORKLOADS["cfg3"][3] // world, max(5, min(args.steps, 20)), "strong",
                                              name="cfg3_strong"))
        secondary.append(secondary_record(ctx, "cfg4", WORKLOADS["cfg4"][3], 5, "weak"))
        secondary.append(secondary_record(ctx, "cfg2", WORKLOADS["cfg2"][3], max(5, min(args.steps, 20)), "weak", with_e2e=True))
        if world == 1:
            secondary.append(cfg1_record(ctx))

    if rank == 0:
        tf = 0.0
        try:
            import ctypes
            val, unused = ctypes.c_double(), ctypes.c_double()
            if lib.qnn_fp32_peak_probe(4096, ctypes.byref(val), ctypes.byref(unused)) == 0:
                tf = val.value
        except Exception:
            pass
        roofline = m["roofline"]
        roofline["peak_measured"] = tf
        roofline["frac_of_measured"] = (roofline["achieved"] / tf) if tf else None
        cpu = time_cpu(wl)[0] if (world == 1 and not args.no_cpu) else None
        line = {
            "metric": "qnn_samples_per_sec_fwd_bwd", "value": value, "unit": "samples/s", "n_gpus": world,
            "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms_per_step,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32 (complex64 state)",
            "data": "synthetic",
            "config": {"workload": desc, "name": wl, "n_qubits": n, "n_layers": L, "batch_per_gpu": B,
                       "global_batch": B * world, "parallelism": f"dp{world}",
                       "l2": "flushed between timed steps (256 MiB memset, outside the per-step CUDA events)",
                       "plan": m["plan"], "launch": "device-timed step " + m["launch"]},
            "clocks": clocks.summary(),
            "e2e": e2e,
            "gpu_launches": 2 * launches_per_step * args.steps,
            "roofline": roofline,
            "secondary": secondary,
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
    ap.add_argument("--eager", action="store_true", help="launch every device-timed step eagerly instead of replaying a CUDA graph")
    ap.add_argument("--no-secondary", action="store_true", help="skip the secondary records (cfg4, cfg2, cfg1, strong scaling)")
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
