"""Data-parallel plumbing for the ICP path: one process per GPU, `torch.distributed`.

The path shards by independent units -- every (scan, map) pair is its own ICP problem and dICP's
batch dimension never mixes samples (`external/dICP/dICP/ICP.py:130-250`) --
so ranks exchange nothing on the data path.  What is needed around it:

  shard_range      contiguous split of a batch of pairs over the ranks
  max_over_ranks   device-side timing / statistic reduced with MAX (bench.py)
  allreduce_mean_  one flat all-reduce of a list of gradient tensors (the UNet gradients of the
                   training step, ~7 MB: latency-bound over NVLink, so a single bucket)
  BucketedGradReducer  the same reduction started from gradient-ready hooks, bucket by bucket, while
                   backward is still running, in a launch order that is the same
                   on every rank whatever the timing
  shard_loss_scale / global_mean   weights that make unequal shards (batch % world != 0) reproduce
                   the single-process batch-mean loss and gradient

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


def shard_loss_scale(n_local: int, device=None) -> torch.Tensor:
    """Factor for this rank's batch-MEAN loss so that the plain mean over ranks of the gradients equals the
    gradient of the mean over the WHOLE batch: n_local * world / n_global (1 when the shards are equal).
    A 0-dim tensor; no host synchronisation."""
    rank, ws = world()
    if ws == 1:
        return torch.ones((), device=device)
    cnt = torch.tensor(float(n_local), device=device)
    dist.all_reduce(cnt, op=dist.ReduceOp.SUM)
    return float(n_local) * ws / cnt


def global_mean(value: torch.Tensor, n_local: int) -> torch.Tensor:
    """mean over the whole batch of a per-rank batch mean (weights n_local / n_global)"""
    rank, ws = world()
    if ws == 1:
        return value
    t = torch.stack((value.detach().reshape(()) * n_local, torch.tensor(float(n_local), device=value.device,
                                                                         dtype=value.dtype)))
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return t[0] / t[1]


class BucketedGradReducer:
    """Mean over ranks of the gradients of `params`, overlapped with backward.

    The parameters are grouped, in reverse registration order (the order gradients become ready in), into
    buckets of about `bucket_bytes`; a post-accumulate hook on every parameter counts its bucket down and a
    bucket is all-reduced asynchronously as soon as it AND every bucket before it are ready -- buckets are
    always launched in index order, so the sequence of collectives is identical on all ranks even when a
    parameter receives no gradient on some rank (its slot is zero-filled in `finish()`).  `finish()` launches
    what is left, waits, divides by the world size and copies the means back into `.grad`.
    One process per GPU; with a single process it does nothing."""

    def __init__(self, params, bucket_bytes: int = 2 << 20):
        self.params = [p for p in params if p.requires_grad]
        self.buckets, cur, size = [], [], 0
        for p in reversed(self.params):
            cur.append(p)
            size += p.numel() * p.element_size()
            if size >= bucket_bytes:
                self.buckets.append(cur); cur, size = [], 0
        if cur:
            self.buckets.append(cur)
        self.bucket_of = {id(p): i for i, b in enumerate(self.buckets) for p in b}
        self._reset()
        self.handles = [p.register_post_accumulate_grad_hook(self._ready) for p in self.params]

    def _reset(self):
        self.pending = [len(b) for b in self.buckets]
        self.launched = 0
        self.work = [None] * len(self.buckets)

    def _ready(self, p):
        if world()[1] == 1:
            return
        self.pending[self.bucket_of[id(p)]] -= 1
        self._launch_ready()

    def _launch_ready(self, force: bool = False):
        while self.launched < len(self.buckets) and (force or self.pending[self.launched] <= 0):
            b = self.buckets[self.launched]
            for p in b:
                if p.grad is None:
                    p.grad = torch.zeros_like(p)
            flat = torch.cat([p.grad.reshape(-1) for p in b])
            self.work[self.launched] = (flat, dist.all_reduce(flat, op=dist.ReduceOp.SUM, async_op=True))
            self.launched += 1

    def finish(self):
        ws = world()[1]
        if ws == 1:
            return
        self._launch_ready(force=True)
        for b, (flat, handle) in zip(self.buckets, self.work):
            handle.wait()
            flat /= ws
            off = 0
            for p in b:
                n = p.numel()
                p.grad.copy_(flat[off:off + n].view_as(p.grad))
                off += n
        self._reset()

    def remove(self):
        for h in self.handles:
            h.remove()
