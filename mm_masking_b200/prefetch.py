"""Host -> device staging for the training / localisation loop.

The reference moves every batch to the GPU inside the loop body (`train_icp_weights.py:34-41`:
`batch_T['T_ml_init'].to(device)`, `icp_weight_policy.py:122-135`: `.to(self.device)` on the scan and
map tensors), i.e. the upload of batch k sits between the kernels of batch k-1 and batch k.  At c3
the map points alone are 38 MB per batch - about as long on PCIe as the whole ICP solve on the GPU.

`DevicePrefetcher` wraps the same iterator and hands out device copies: the upload of batch k+1 is
enqueued on a dedicated copy stream before batch k is returned, so it runs under the kernels of batch
k.  Every batch is still copied exactly once per step; nothing is cached across steps.
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

    def __init__(self, iterable, device):
        self.device = torch.device(device)
        self._it = iter(iterable)
        self._cuda = self.device.type == "cuda"
        self._stream = torch.cuda.Stream(self.device) if self._cuda else None
        self._next = None
        self.bytes_uploaded = 0
        self._preload()

    def _preload(self):
        try:
            host = next(self._it)
        except StopIteration:
            self._next = None
            return
        if not self._cuda:
            self._next = (_map_tensors(host, lambda t: t.to(self.device)), None)
            return
        with torch.cuda.stream(self._stream):
            dev = _map_tensors(host, lambda t: t.to(self.device, non_blocking=True))
            ev = torch.cuda.Event()
            ev.record(self._stream)
        self.bytes_uploaded += sum(t.numel() * t.element_size() for t in _tensors(host, []) if not t.is_cuda)
        self._next = (dev, ev)

    def __iter__(self):
        return self

    def __next__(self):
        if self._next is None:
            raise StopIteration
        dev, ev = self._next
        if self._cuda:
            cur = torch.cuda.current_stream(self.device)
            cur.wait_event(ev)
            for t in _tensors(dev, []):
                t.record_stream(cur)
        self._preload()
        return dev
