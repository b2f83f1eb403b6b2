"""world_size-2 gloo tests (CPU) of the sweep-sharding host logic in tdssf/parallel.py: the shards partition a batch,
the confusion-matrix SUM equals the single-process matrix of the union (oracle), the gradient bucket averages exactly,
and a rank without gradients contributes zeros."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from tdssf import parallel
from oracle import np_oracle


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _run(fn, world_size, *args):
    port = _free_port()
    mp.spawn(_entry, args=(world_size, port, fn, args), nprocs=world_size, join=True)


def _entry(rank, world_size, port, fn, args):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world_size)
    try:
        fn(rank, world_size, *args)
    finally:
        dist.destroy_process_group()


# ---------------------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("n,w", [(0, 1), (1, 2), (8, 2), (7, 3), (5, 8), (64, 8)])
def test_shard_bounds_partition(n, w):
    spans = [parallel.shard_bounds(n, r, w) for r in range(w)]
    assert spans[0][0] == 0 and spans[-1][1] == n
    for (a0, a1), (b0, b1) in zip(spans, spans[1:]):
        assert a1 == b0 and a0 <= a1
    sizes = [hi - lo for lo, hi in spans]
    assert max(sizes) - min(sizes) <= 1 and sizes == sorted(sizes, reverse=True)


def test_shard_bounds_errors():
    with pytest.raises(ValueError):
        parallel.shard_bounds(4, 2, 2)
    with pytest.raises(ValueError):
        parallel.shard_bounds(-1, 0, 1)


def test_single_process_is_identity():
    assert parallel.world() == (0, 1)
    cm = torch.arange(16, dtype=torch.int64).view(4, 4)
    assert parallel.allreduce_confusion(cm.clone()).equal(cm)
    assert parallel.max_over_ranks(3.5) == 3.5
    assert parallel.shard_sweeps([1, 2, 3]) == [1, 2, 3]
    with pytest.raises(TypeError):
        parallel.allreduce_confusion(cm.float())


# ---------------------------------------------------------------------------------------------------------------
def _confusion_worker(rank, world_size, n_sweeps, pts, C):
    rng = np.random.default_rng(7)
    preds = rng.integers(1, C + 1, size=(n_sweeps, pts))
    labels = rng.integers(1, C + 1, size=(n_sweeps, pts))
    lo, hi = parallel.shard_bounds(n_sweeps, rank, world_size)
    assert parallel.shard_sweeps(torch.from_numpy(preds)).shape[0] == hi - lo
    mine = np_oracle.confusion_matrix(preds[lo:hi].ravel(), labels[lo:hi].ravel(), C)
    cm = torch.from_numpy(np.asarray(mine, dtype=np.int64))
    parallel.allreduce_confusion(cm)
    whole = np_oracle.confusion_matrix(preds.ravel(), labels.ravel(), C)
    assert np.array_equal(cm.numpy(), whole)       # bit-exact, every rank
    assert int(cm.sum()) == n_sweeps * pts
    n = parallel.sum_over_ranks(float((hi - lo) * pts))
    assert n == n_sweeps * pts
    assert parallel.max_over_ranks(float(rank)) == world_size - 1


def test_confusion_allreduce_gloo_ws2():
    _run(_confusion_worker, 2, 5, 1000, 16)


def _grad_worker(rank, world_size):
    torch.manual_seed(0)
    mlp = torch.nn.Sequential(torch.nn.Linear(448, 256), torch.nn.ReLU(), torch.nn.Linear(256, 64), torch.nn.ReLU(),
                              torch.nn.Linear(64, 16))
    # the whole batch, and this rank's sweeps of it
    x = torch.randn(4, 50, 448)
    y = torch.randn(4, 50, 16)
    lo, hi = parallel.shard_bounds(4, rank, world_size)
    loss = ((mlp(x[lo:hi]) - y[lo:hi]) ** 2).mean()
    loss.backward()
    n = parallel.allreduce_mean_grads(mlp.parameters())
    assert n == 448 * 256 + 256 + 256 * 64 + 64 + 64 * 16 + 16
    got = [p.grad.clone() for p in mlp.parameters()]
    mlp.zero_grad()
    # equal-size shards: the mean of per-shard mean losses is the whole-batch mean loss
    ((mlp(x) - y) ** 2).mean().backward()
    for g, p in zip(got, mlp.parameters()):
        torch.testing.assert_close(g, p.grad, rtol=1e-5, atol=1e-7)


def test_grad_bucket_allreduce_gloo_ws2():
    _run(_grad_worker, 2)


def _missing_grad_worker(rank, world_size):
    lin = torch.nn.Linear(3, 2)
    if rank == 0:
        lin.weight.grad = torch.full((2, 3), 4.0)
        lin.bias.grad = torch.full((2,), 2.0)
    parallel.allreduce_mean_grads(lin.parameters())
    assert torch.equal(lin.weight.grad, torch.full((2, 3), 2.0))
    assert torch.equal(lin.bias.grad, torch.full((2,), 1.0))


def test_rank_without_grads_contributes_zeros():
    _run(_missing_grad_worker, 2)


def _shared_flat_worker(rank, world_size):
    """Gradients that are back-to-back views of one buffer (what the bf16 backward hands to autograd) are reduced in place."""
    mlp = torch.nn.Sequential(torch.nn.Linear(8, 4), torch.nn.ReLU(), torch.nn.Linear(4, 4))
    params = list(mlp.parameters())
    flat = torch.randn(sum(p.numel() for p in params), generator=torch.Generator().manual_seed(100 + rank))
    mine = flat.clone()
    off = 0
    for p in params:
        p.grad = flat[off:off + p.numel()].view_as(p)
        off += p.numel()
    assert parallel._shared_flat_grad(params) is not None
    n = parallel.allreduce_mean_grads(params)
    assert n == flat.numel()
    other = torch.randn(flat.numel(), generator=torch.Generator().manual_seed(100 + 1 - rank))
    torch.testing.assert_close(flat, (mine + other) / 2)          # reduced in place: the views see the mean
    assert params[0].grad.untyped_storage().data_ptr() == flat.untyped_storage().data_ptr()
    # a gap between the views (or a foreign storage) falls back to the gather / scatter path
    params[1].grad = params[1].grad.clone()
    assert parallel._shared_flat_grad(params) is None


def test_shared_flat_grad_bucket_gloo_ws2():
    _run(_shared_flat_worker, 2)
