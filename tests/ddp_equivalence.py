"""Launched by tests/test_ddp_gpu.py as ``python -m torch.distributed.run --nproc-per-node N``: the data-parallel
step on N GPUs against the single-GPU step on the concatenated batch (the reference's
only multi-GPU mechanism is nn.DataParallel around the classical nets, Runner_P128_...py:144-148).

Every rank computes the single-GPU result for the WHOLE batch on its own device, then its shard of the same
batch through the same C-ABI calls, all-reduces, and compares.  Prints one ``DDP-OK`` line per check on rank 0
and exits non-zero on any mismatch.
"""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import qris_b200  # noqa: E402
from helpers import assert_close  # noqa: E402
from qris_b200 import estimators as E  # noqa: E402
from qris_b200 import runner  # noqa: E402
from qris_b200.parallel import FlatGradBucket, shard_range  # noqa: E402


def say(rank, msg):
    if rank == 0:
        print("DDP-OK " + msg, flush=True)


def kernel_level(rank, world, dev, circ, B, tag):
    """qnn_forward_save / qnn_backward_saved on this rank's rows + one all-reduce of gw == the whole batch on one GPU."""
    g = torch.Generator().manual_seed(11)
    x = (2 * torch.rand(B, circ.n_inputs, generator=g) - 1).to(dev)
    w = (2 * np.pi * torch.rand(circ.n_weights, generator=g)).to(dev)
    gy = torch.randn(B, circ.n_qubits, generator=g).to(dev)
    y1, st1 = qris_b200.qnn_forward_raw(circ, x, w, save_state=True)
    gx1, gw1 = qris_b200.qnn_backward_raw(circ, x, w, gy, state=st1)
    lo, hi = shard_range(B, rank, world)
    xs, gys = x[lo:hi].contiguous(), gy[lo:hi].contiguous()
    ys, sts = qris_b200.qnn_forward_raw(circ, xs, w, save_state=True)
    bucket = torch.zeros(circ.n_weights, device=dev)
    gxs, _ = qris_b200.qnn_backward_raw(circ, xs, w, gys, state=sts, gw_out=bucket)
    dist.all_reduce(bucket)
    assert torch.equal(ys, y1[lo:hi]), "y rows of the shard differ from the single-GPU rows"
    assert_close(gxs.cpu().numpy(), gx1[lo:hi].cpu().numpy(), rel=1e-6, scale=float(gx1.abs().max()), what="gx rows")
    assert_close(bucket.cpu().numpy(), gw1.cpu().numpy(), rel=1e-6, scale=float(gw1.abs().max()), what="all-reduced gw")
    say(rank, f"kernel-level {tag}: B={B} over {world} ranks (rows {lo}:{hi} here), max|dgw|/max|gw| = "
              f"{float((bucket - gw1).abs().max() / gw1.abs().max()):.2e}")


def qnat_global_statistics(rank, world, dev, B):
    """QuantumNATHead(process_group=WORLD) on shards == the head on the whole batch (ragged shards included)."""
    n = 10
    g = torch.Generator().manual_seed(5)
    y = (2 * torch.rand(B, n, generator=g) - 1).to(dev)
    noise = torch.randn(B, n, generator=g).to(dev)
    gout = torch.randn(B, n, generator=g).to(dev)
    single = qris_b200.QuantumNATHead(sigma=0.01, lo=-2.0, hi=2.0, levels=5).to(dev).train()
    shared = qris_b200.QuantumNATHead(sigma=0.01, lo=-2.0, hi=2.0, levels=5, process_group=dist.group.WORLD).to(dev).train()
    y1 = y.clone().requires_grad_(True)
    o1 = single(y1, noise); o1.backward(gout)
    lo, hi = shard_range(B, rank, world)
    y2 = y[lo:hi].clone().requires_grad_(True)
    o2 = shared(y2, noise[lo:hi].contiguous()); o2.backward(gout[lo:hi].contiguous())
    # quantised outputs identical, except where the pre-quantisation value sits on a rounding edge
    differ = (o2 != o1[lo:hi])
    assert int(differ.sum()) <= max(2, (hi - lo) * n // 20000), f"{int(differ.sum())} quantised outputs differ"
    assert_close(y2.grad.cpu().numpy(), y1.grad[lo:hi].cpu().numpy(), rel=1e-5, scale=float(y1.grad.abs().max()), what="qnat gy")
    say(rank, f"QuantumNAT global statistics: B={B} (shard {hi - lo} rows here), {int(differ.sum())} edge differences")


def module_level(rank, world, dev, B):
    """QSC_QuantumNAT (config 3's model, batch statistics shared across ranks) for one step == the single-GPU
    gradient of the global-batch mean loss.  The head's backward couples samples of different ranks, so every
    rank scales its loss by the GLOBAL row count (sum / B) and the bucket is summed, not averaged; this holds
    for ragged shards too."""
    torch.manual_seed(3)
    single = E.QSC_QuantumNAT(n_qubits=10, n_layers=2, noise_level=0.0, act_noise=0.0).to(dev).train()
    torch.manual_seed(3)
    shared = E.QSC_QuantumNAT(n_qubits=10, n_layers=2, noise_level=0.0, act_noise=0.0, process_group=dist.group.WORLD).to(dev).train()
    g = torch.Generator().manual_seed(8)
    x = torch.randn(B, 2, 16, 8, generator=g).to(dev)
    lab = torch.randint(0, 3, (B,), generator=g).to(dev)
    b1 = FlatGradBucket(single.parameters()); b1.zero()
    F.nll_loss(single(x), lab).backward()
    lo, hi = shard_range(B, rank, world)
    b2 = FlatGradBucket(shared.parameters()); b2.zero()
    (F.nll_loss(shared(x[lo:hi]), lab[lo:hi], reduction="sum") / B).backward()
    b2.all_reduce(average=False)
    # away from quantiser edges the two graphs are the same function; allow the few samples whose bin flipped
    err = (b2.flat - b1.flat).abs().max() / b1.flat.abs().max()
    assert float(err) < 5e-4, f"module-level gradient differs: {float(err):.3e}"
    say(rank, f"module-level QSC_QuantumNAT step: B={B} ({hi - lo} rows here), max|dg|/max|g| = {float(err):.2e}")


def module_level_row_weighted(rank, world, dev, B):
    """QSC_P128 (the reference's model: no batch statistics) with per-shard MEAN losses on ragged shards: the
    bucket weights each rank by its row count (``local_rows``) and reproduces the global-batch mean gradient."""
    torch.manual_seed(4)
    single = E.QSC_P128(use_quantumnat=False, use_gradient_pruning=False).to(dev).train()
    shared = E.QSC_P128(use_quantumnat=False, use_gradient_pruning=False).to(dev).train()
    shared.load_state_dict(single.state_dict())
    g = torch.Generator().manual_seed(9)
    x = torch.randn(B, 2, 16, 8, generator=g).to(dev)
    lab = torch.randint(0, 3, (B,), generator=g).to(dev)
    b1 = FlatGradBucket(single.parameters()); b1.zero()
    F.nll_loss(single(x), lab).backward()
    lo, hi = shard_range(B, rank, world)
    b2 = FlatGradBucket(shared.parameters()); b2.zero()
    F.nll_loss(shared(x[lo:hi]), lab[lo:hi]).backward()
    b2.all_reduce(local_rows=hi - lo)
    err = (b2.flat - b1.flat).abs().max() / b1.flat.abs().max()
    assert float(err) < 2e-5, f"row-weighted gradient differs: {float(err):.3e}"
    say(rank, f"module-level QSC_P128 step, row-weighted mean of ragged shards: B={B} ({hi - lo} rows here), "
              f"max|dg|/max|g| = {float(err):.2e}")


def runner_under_torchrun(rank, world, dev):
    """runner.Y2HRunner.train_QSC_P128 under torchrun: the rank-slicing loader and the bucket all-reduce execute;
    all ranks hold identical parameters afterwards and the loss is finite."""
    r = runner.Y2HRunner(data_len=400, batch_size_DML=64, n_epochs=1, device=dev, verbose=False,
                         workdir=os.path.join(ROOT, "gpurun_out", "ddp_workspace"))
    hist = r.train_QSC_P128(n_epochs=1, max_iters_per_epoch=4, save=False)
    flat = torch.cat([p.detach().reshape(-1) for p in r.model_QSC.parameters()])
    ref = flat.clone()
    dist.broadcast(ref, src=0)
    assert torch.equal(flat, ref), "ranks diverged"
    assert np.isfinite(hist["train_loss"][0]) and 0.0 <= hist["val_acc"][0] <= 1.0
    say(rank, f"runner.train_QSC_P128 under torchrun: loss {hist['train_loss'][0]:.4f}, val acc {hist['val_acc'][0]:.3f}, "
              f"parameters identical on {world} ranks")


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    # cuDNN's TF32 convolutions (torch's default) round differently for different batch shapes; the comparison
    # of a shard with the whole batch is about the data-parallel arithmetic, so keep the classical front end in fp32
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    dist.init_process_group("nccl", device_id=dev)
    try:
        kernel_level(rank, world, dev, qris_b200.u3cu3_circuit(10, 3), 8192, "cfg3 circuit")
        kernel_level(rank, world, dev, qris_b200.reference_circuit(6, 3), 2305, "reference circuit, ragged")
        kernel_level(rank, world, dev, qris_b200.u3cu3_circuit(14, 1), 6, "cluster geometry")
        qnat_global_statistics(rank, world, dev, 8192)
        qnat_global_statistics(rank, world, dev, 4097)
        module_level(rank, world, dev, 510)
        module_level(rank, world, dev, 257)
        module_level_row_weighted(rank, world, dev, 771)
        runner_under_torchrun(rank, world, dev)
        say(rank, "all checks passed")
    finally:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
