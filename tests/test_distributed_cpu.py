"""CPU, world_size 2 over gloo: the data-parallel host logic (shard ranges, flat gradient bucket,
one all-reduce per step).  Kernels are not involved; a small torch model stands in for the classifier."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from qris_b200.parallel import FlatGradBucket, shard_range


def test_shard_ranges_cover_the_batch():
    for n, world in ((10, 2), (7, 3), (65536, 8), (5, 8), (1, 2)):
        spans = [shard_range(n, r, world) for r in range(world)]
        assert spans[0][0] == 0 and spans[-1][1] == n
        assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
        assert all(hi >= lo for lo, hi in spans)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, out, n_rows=32, weighted=False):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        torch.manual_seed(0)                                   # same initial model on every rank
        model = torch.nn.Sequential(torch.nn.Linear(6, 5), torch.nn.Tanh(), torch.nn.Linear(5, 3))
        g = torch.Generator().manual_seed(1)
        x = torch.randn(n_rows, 6, generator=g); y = torch.randint(0, 3, (n_rows,), generator=g)
        lo, hi = shard_range(n_rows, rank, world)
        bucket = FlatGradBucket(model.parameters())
        for _ in range(2):                                     # zero() must re-arm the bucket
            bucket.zero()
            loss = torch.nn.functional.cross_entropy(model(x[lo:hi]), y[lo:hi])
            loss.backward()
            bucket.all_reduce(local_rows=(hi - lo) if weighted else None)
        # every parameter's .grad is still a view of the flat buffer
        assert all(p.grad.data_ptr() == bucket.flat[a:b].data_ptr() for p, (a, b) in zip(bucket.params, bucket.slices))
        out[rank] = bucket.flat.clone()
    finally:
        dist.destroy_process_group()


def test_two_rank_step_equals_single_process_step():
    world, port = 2, _free_port()
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(world, port, out), nprocs=world, join=True)
    torch.manual_seed(0)
    model = torch.nn.Sequential(torch.nn.Linear(6, 5), torch.nn.Tanh(), torch.nn.Linear(5, 3))
    g = torch.Generator().manual_seed(1)
    x = torch.randn(32, 6, generator=g); y = torch.randint(0, 3, (32,), generator=g)
    torch.nn.functional.cross_entropy(model(x), y).backward()
    ref = torch.cat([p.grad.reshape(-1) for p in model.parameters()])
    assert torch.allclose(out[0], out[1])                      # ranks agree after the all-reduce
    assert torch.allclose(out[0], ref, atol=1e-6)              # mean of equal-shard means == global mean


def test_ragged_shards_weighted_by_row_count_equal_single_process_step():
    """33 rows over 2 ranks = 17 + 16: the mean of shard means is NOT the global mean; with ``local_rows`` the
    bucket weights each rank by its row count and divides by the all-reduced total."""
    world, port = 2, _free_port()
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(world, port, out, 33, True), nprocs=world, join=True)
    torch.manual_seed(0)
    model = torch.nn.Sequential(torch.nn.Linear(6, 5), torch.nn.Tanh(), torch.nn.Linear(5, 3))
    g = torch.Generator().manual_seed(1)
    x = torch.randn(33, 6, generator=g); y = torch.randint(0, 3, (33,), generator=g)
    torch.nn.functional.cross_entropy(model(x), y).backward()
    ref = torch.cat([p.grad.reshape(-1) for p in model.parameters()])
    assert torch.allclose(out[0], out[1]) and torch.allclose(out[0], ref, atol=1e-6)


def test_bucket_without_process_group_is_a_noop():
    model = torch.nn.Linear(3, 2)
    bucket = FlatGradBucket(model.parameters())
    model(torch.ones(4, 3)).sum().backward()
    before = bucket.flat.clone()
    bucket.all_reduce()
    assert torch.equal(before, bucket.flat) and before.abs().sum() > 0
