"""Host -> device staging for the training / localisation loop.

The reference moves every batch to the GPU inside the loop body (`train_icp_weights.py:34-41`:
`batch_T['T_ml_init'].to(device)`, `icp_weight_policy.py:122-135`: `.to(self.device)` on the scan and
map tensors), i.e. the upload of batch k sits between the kernels of batch k-1 and batch k.  At c3
the map points alone are 38 MB per batch - about as long on PCIe as the whole ICP solve on the GPU.

`DevicePrefetcher` wraps the same iterator and hands out device copies: the upload of batch k+1 is
enqueued on a dedicated copy stream before batch k is returned, so it runs under the kernels of batch
k.  Every batch is still copied exactly once per step; nothing is cached across steps.

Large pinned tensors are uploaded in two halves on two copy streams: on the pool's B200 boxes (PCIe 5 x16) one
`cudaMemcpyAsync` stream sustains 31 GB/s, two sustain 47 GB/s (`profiles/h2d_probe.py`), which is the difference
between an upload that is longer than the c3 step (39.4 MB: 1.0 ms against 0.86 ms of kernels) and one that hides
under it (0.83 ms).
"""
from __future__ import annotations

import torch


def _map_tensors(obj, fn):
    if torch.is_tensor(obj):
        return fn(obj)
    if isinstance(obj, dict):
        return {k: _map_tensors(v, fn) for k, v in obj.items()}
    if isinstance(obj, (list, tuple)):
        return type(obj)(_map_tensors(v, fn) for v in obj)
    return obj


def _tensors(obj, out):
    if torch.is_tensor(obj):
        out.append(obj)
    elif isinstance(obj, dict):
        for v in obj.values():
            _tensors(v, out)
    elif isinstance(obj, (list, tuple)):
        for v in obj:
            _tensors(v, out)
    return out


def pin_batch(batch):
    """Page-lock every tensor of a (nested) host batch so that its upload is asynchronous."""
    return _map_tensors(batch, lambda t: t if t.is_cuda or t.is_pinned() else t.pin_memory())


class DevicePrefetcher:
    """Iterate over host batches (nested dicts / lists of tensors), yielding copies on `device`.

    With a CUDA device the copy of the NEXT batch is issued on a side stream when the current one is
    handed out; the consumer's stream waits on the copy's event, and the tensors are registered with
    the consumer's stream so the caching allocator does not recycle them early.  Host tensors should
    be pinned (`pin_batch`), otherwise the copy blocks the host.  With a CPU device it is a plain
    pass-through (host-side tests)."""

    SPLIT_BYTES = 4 << 20          # pinned tensors at least this large go up in two halves on two copy streams

    def __init__(self, iterable, device):
        self.device = torch.device(device)
        self._it = iter(iterable)
        self._cuda = self.device.type == "cuda"
        self._stream = torch.cuda.Stream(self.device) if self._cuda else None
        self._stream2 = torch.cuda.Stream(self.device) if self._cuda else None
        self._next = None
        self.bytes_uploaded = 0
        self._preload()

    def _upload(self, host):
        """device copy of a nested host batch; returns (device batch, events the consumer waits on)"""
        second = []                                      # (dst half, src half) for the second copy stream

        def up(t):
            if t.is_cuda:
                return t.to(self.device, non_blocking=True)
            nbytes = t.numel() * t.element_size()
            if nbytes >= self.SPLIT_BYTES and t.is_pinned() and t.is_contiguous():
                dst = torch.empty(t.shape, dtype=t.dtype, device=self.device)
                half = t.numel() // 2
                second.append((dst, dst.view(-1)[half:], t.view(-1)[half:], t.view(-1)[:half]))
                return dst
            return t.to(self.device, non_blocking=True)

        with torch.cuda.stream(self._stream):
            dev = _map_tensors(host, up)
            ready = torch.cuda.Event()
            ready.record(self._stream)                   # allocations (and the small copies) are ordered before stream 2
            for dst, _, _, src_lo in second:
                dst.view(-1)[:src_lo.numel()].copy_(src_lo, non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(self._stream)
        events = [ev]
        if second:
            self._stream2.wait_event(ready)
            with torch.cuda.stream(self._stream2):
                for dst, dst_hi, src_hi, _ in second:
                    dst_hi.copy_(src_hi, non_blocking=True)
                    dst.record_stream(self._stream2)
                ev2 = torch.cuda.Event()
                ev2.record(self._stream2)
            events.append(ev2)
        return dev, events

    def _preload(self):
        try:
            host = next(self._it)
        except StopIteration:
            self._next = None
            return
        if not self._cuda:
            self._next = (_map_tensors(host, lambda t: t.to(self.device)), None)
            return
        dev, events = self._upload(host)
        self.bytes_uploaded += sum(t.numel() * t.element_size() for t in _tensors(host, []) if not t.is_cuda)
        self._next = (dev, events)

    def __iter__(self):
        return self

    def __next__(self):
        if self._next is None:
            raise StopIteration
        dev, ev = self._next
        if self._cuda:
            cur = torch.cuda.current_stream(self.device)
            for e in ev:
                cur.wait_event(e)
            for t in _tensors(dev, []):
                t.record_stream(cur)
        self._preload()
        return dev
