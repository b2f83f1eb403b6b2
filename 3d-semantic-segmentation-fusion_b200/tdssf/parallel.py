"""Sweep-sharded data parallelism for the fusion path: one process per GPU, no collective on the data path.

The reference is single-process (`main_train.py`, `utils/train.py:30-75`, `utils/test.py:60-95`); every row of the hot
path is independent per sweep, so N GPUs each take their own sweeps of the DataLoader batch.  Only two quantities ever
cross ranks, both outside the kernels and both tiny:

    * the int64 confusion matrix of the evaluation loop (utils/evaluation.py:35-42, summed over batches at
      utils/test.py:82-87) — SUM, exact in integers;
    * the fusion-MLP gradients of one training step (132 k fp32 parameters, mlp.{0,2,4}) — averaged, in ONE flat
      bucket (0.5 MB: launch-latency bound on NVSwitch, so one collective, not six).

`torch.distributed` is the plumbing (NCCL on the GPUs, gloo in the CPU tests).  The gradient bucket of CUDA ranks that
share a node does not go through NCCL: csrc/p2p.cu averages it with ONE kernel over NVLink peer memory (every rank's
exchange buffer mapped into every process through CUDA IPC, set up collectively at the first call; any failure anywhere,
several nodes, another backend or TDSSF_P2P=0 -> everybody uses ncclAllReduce).  At 0.5 MB the collective is launch-
latency bound: the kernel costs one launch from a ctypes call where the ProcessGroup route costs ~70 us of host time.
"""
from __future__ import annotations

from typing import Iterable, Sequence

import torch
import torch.distributed as dist


def world():
    """(rank, world_size) of the default group; (0, 1) when torch.distributed is not initialised."""
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    return 0, 1


def shard_bounds(n_items: int, rank: int, world_size: int) -> tuple[int, int]:
    """Contiguous [lo, hi) slice of `n_items` sweeps owned by `rank`: sizes differ by at most one, earlier ranks take
    the larger shares, every sweep is owned exactly once."""
    if world_size < 1 or not 0 <= rank < world_size:
        raise ValueError(f"bad rank/world_size {rank}/{world_size}")
    if n_items < 0:
        raise ValueError("n_items must be >= 0")
    base, extra = divmod(n_items, world_size)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def shard_sweeps(batch: Sequence, rank: int | None = None, world_size: int | None = None):
    """This rank's sweeps of a batch.  Tensors are sliced along dim 0 (views, no copy); lists/tuples (the reference's
    per-sweep calibration lists, utils/dataset.py collate) are sliced as lists."""
    r, w = world()
    rank = r if rank is None else rank
    world_size = w if world_size is None else world_size
    lo, hi = shard_bounds(len(batch), rank, world_size)
    return batch[lo:hi]


def allreduce_confusion(cm: torch.Tensor) -> torch.Tensor:
    """In-place SUM of the int64 (C,C) confusion matrix over ranks (exact); returns `cm`."""
    if cm.dtype != torch.int64:
        raise TypeError("confusion matrix must be int64")
    if world()[1] > 1:
        dist.all_reduce(cm, op=dist.ReduceOp.SUM)
    return cm


class _PeerExchange:
    """NVLink peer-memory exchange of one process group (tdssf_p2p_*): every rank's buffer mapped into every process.
    Set up once, collectively; `ok` is the group's common verdict (any failure anywhere -> everybody uses NCCL)."""

    def __init__(self, max_floats: int, device: torch.device):
        import ctypes as C
        from ._lib import lib
        self.lib, self.C = lib, C
        self.rank, self.world = world()
        self.max_floats = int(max_floats)
        self.device = device
        self.epoch = 0
        self.own = C.c_void_p()
        handle = (C.c_ubyte * 64)()
        good = 1
        with torch.cuda.device(device):
            if self.world > 16 or lib.tdssf_p2p_create(self.max_floats, C.byref(self.own), handle) != 0:
                good = 0
        mine = torch.tensor(list(bytes(handle)) + [good], dtype=torch.uint8, device=device)
        everyone = [torch.empty_like(mine) for _ in range(self.world)]
        dist.all_gather(everyone, mine)
        rows = [bytes(t.cpu().tolist()) for t in everyone]
        good = int(all(r[64] == 1 for r in rows))
        self.bufs = (C.c_void_p * self.world)()
        if good:
            with torch.cuda.device(device):
                for r, row in enumerate(rows):
                    if r == self.rank:
                        self.bufs[r] = self.own.value
                        continue
                    peer = C.c_void_p()
                    h = (C.c_ubyte * 64).from_buffer_copy(row[:64])
                    if lib.tdssf_p2p_open(h, C.byref(peer)) != 0:
                        good = 0
                        break
                    self.bufs[r] = peer.value
        verdict = torch.tensor([good], dtype=torch.int32, device=device)
        dist.all_reduce(verdict, op=dist.ReduceOp.MIN)
        self.ok = bool(int(verdict.item()))

    def allreduce_mean(self, flat: torch.Tensor) -> None:
        from ._lib import check
        self.epoch += 1
        with torch.cuda.device(self.device):
            stream = self.C.c_void_p(torch._C._cuda_getCurrentRawStream(self.device.index))
            check(self.lib.tdssf_p2p_allreduce_mean(self.bufs, self.world, self.rank, self.C.c_void_p(flat.data_ptr()), flat.numel(),
                                                    self.max_floats, self.epoch & 0xffffffff or 1, stream), "tdssf_p2p_allreduce_mean")


_peer_exchange: dict = {}   # (device index, bucket size) -> _PeerExchange of the default group


def _peers_for(n_floats: int, device: torch.device):
    """The default group's peer exchange for a bucket of n_floats, or None (NCCL) - decided collectively, once per size:
    NCCL backend, one node, TDSSF_P2P != 0, every rank's set-up succeeded."""
    import os
    if os.environ.get("TDSSF_P2P", "1") == "0" or dist.get_backend() != "nccl" or n_floats % 4 or n_floats == 0:
        return None
    key = (device.index, n_floats)
    if key not in _peer_exchange:
        _peer_exchange[key] = _PeerExchange(n_floats, device)
    ex = _peer_exchange[key]
    return ex if ex.ok else None


def allreduce_mean_grads(params: Iterable[torch.nn.Parameter]) -> int:
    """Average the gradients of `params` over ranks through one flat fp32 bucket; parameters without a gradient on
    this rank contribute zeros (a rank whose sweeps had no valid labels).  Returns the bucket's element count."""
    params = [p for p in params if p.requires_grad]
    if not params:
        return 0
    _, w = world()
    if w == 1:
        return sum(p.numel() for p in params)   # nothing to average
    dev = params[0].device
    total = sum(p.numel() for p in params)
    # CUDA ranks of one node: ONE kernel over NVLink peer memory (csrc/p2p.cu) instead of a library collective; the choice
    # depends only on the group and the bucket size, so every rank takes the same branch whatever its gradients look like
    peers = _peers_for(total, dev) if dev.type == "cuda" else None
    shared = _shared_flat_grad(params)
    if shared is not None and shared.numel() == total:
        # the bf16 backward hands out the six gradients as views of one buffer (tdssf.ops._flat_zero_grads) and autograd
        # keeps them: the bucket already exists, reduce it in place (one collective, no gather / scatter copies)
        if peers is not None and shared.data_ptr() % 16 == 0:
            peers.allreduce_mean(shared)
        elif peers is not None:
            tmp = shared.clone()
            peers.allreduce_mean(tmp)
            shared.copy_(tmp)
        elif dist.get_backend() == "nccl":
            dist.all_reduce(shared, op=dist.ReduceOp.AVG)
        else:
            dist.all_reduce(shared, op=dist.ReduceOp.SUM)
            shared.div_(w)
        return shared.numel()
    flat = torch.zeros(total, dtype=torch.float32, device=dev)
    off = 0
    for p in params:
        if p.grad is not None:
            flat[off:off + p.numel()].copy_(p.grad.reshape(-1))
        off += p.numel()
    if peers is not None:
        peers.allreduce_mean(flat)
    else:
        dist.all_reduce(flat, op=dist.ReduceOp.SUM)
        flat.div_(w)
    off = 0
    for p in params:
        g = flat[off:off + p.numel()].view_as(p)
        if p.grad is None:
            p.grad = g.clone()
        else:
            p.grad.copy_(g)
        off += p.numel()
    return flat.numel()


def _shared_flat_grad(params):
    """The gradients of `params` as ONE 1-D fp32 tensor if they lie back to back in memory (tdssf.ops._flat_zero_grads hands
    them out as views of one buffer, every size a multiple of four), else None."""
    grads = [p.grad for p in params]
    if any(g is None or g.dtype != torch.float32 or not g.is_contiguous() for g in grads):
        return None
    ptr, total = grads[0].data_ptr(), 0
    for g in grads:
        if g.data_ptr() != ptr + 4 * total:
            return None
        total += g.numel()
    g0 = grads[0]
    if g0.untyped_storage().nbytes() < 4 * (g0.storage_offset() + total):
        return None
    return g0.as_strided((total,), (1,))


def max_over_ranks(value: float, device=None) -> float:
    """MAX of a per-rank scalar (device times in bench.py)."""
    if world()[1] == 1:
        return float(value)
    t = torch.tensor([value], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def sum_over_ranks(value: float, device=None) -> float:
    """SUM of a per-rank scalar (loss / correct / point counters of the evaluation loop, utils/test.py:75-81)."""
    if world()[1] == 1:
        return float(value)
    t = torch.tensor([value], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return float(t.item())
