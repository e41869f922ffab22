"""Data-parallel plumbing for the ICP path: one process per GPU, `torch.distributed`.

The path shards by independent units -- every (scan, map) pair is its own ICP problem and dICP's
batch dimension never mixes samples (`external/dICP/dICP/ICP.py:130-250`) --
so ranks exchange nothing on the data path.  What is needed around it:

  shard_range      contiguous split of a batch of pairs over the ranks
  max_over_ranks   device-side timing / statistic reduced with MAX (bench.py)
  allreduce_mean_  one flat all-reduce of a list of gradient tensors (the UNet gradients of the
                   training step, ~7 MB: latency-bound over NVLink, so a single bucket)

Works with NCCL on GPUs and with gloo on CPU (the tests run world_size 2 over gloo).
"""
from __future__ import annotations

import torch
import torch.distributed as dist


def world() -> tuple[int, int]:
    """(rank, world_size); (0, 1) when no process group is initialised."""
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    return 0, 1


def shard_range(total: int, world_size: int, rank: int) -> tuple[int, int]:
    """[start, end) of the pairs rank `rank` owns: contiguous, sizes differ by at most one, the
    first `total % world_size` ranks take the extra pair."""
    if world_size <= 0 or not (0 <= rank < world_size):
        raise ValueError("bad rank / world_size")
    base, extra = divmod(total, world_size)
    start = rank * base + min(rank, extra)
    return start, start + base + (1 if rank < extra else 0)


def shard_batch(batch: dict, world_size: int, rank: int) -> dict:
    """Slice every array/tensor of a batch dict along dim 0 to this rank's pairs."""
    n = len(next(iter(batch.values())))
    s, e = shard_range(n, world_size, rank)
    return {k: v[s:e] for k, v in batch.items()}


def max_over_ranks(value: float, device=None) -> float:
    rank, ws = world()
    if ws == 1:
        return float(value)
    t = torch.tensor([float(value)], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def allreduce_mean_(tensors) -> None:
    """In-place mean over ranks of a list of same-dtype tensors through ONE flat buffer."""
    rank, ws = world()
    tensors = [t for t in tensors if t is not None]
    if ws == 1 or not tensors:
        return
    flat = torch.cat([t.reshape(-1) for t in tensors])
    dist.all_reduce(flat, op=dist.ReduceOp.SUM)
    flat /= ws
    off = 0
    for t in tensors:
        n = t.numel()
        t.copy_(flat[off:off + n].view_as(t))
        off += n
