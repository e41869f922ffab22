"""`mm_masking_b200.prefetch`: host batches come out unchanged, in order, one upload per batch."""
import pytest
import torch

from mm_masking_b200.prefetch import DevicePrefetcher, pin_batch


def _batches(count):
    for k in range(count):
        g = torch.Generator().manual_seed(k)
        yield {"loc_data": {"raw_pc": torch.rand((2, 50, 3), generator=g), "id": [k, "scan"]},
               "transforms": {"T_ml_init": torch.eye(4).repeat(2, 1, 1) * (k + 1)},
               "list": [torch.arange(5) + k, (torch.ones(3) * k,)]}


def _check(dev):
    got = list(DevicePrefetcher(_batches(4), dev))
    assert len(got) == 4
    for k, (b, ref) in enumerate(zip(got, _batches(4))):
        assert b["loc_data"]["id"] == [k, "scan"]
        assert b["loc_data"]["raw_pc"].device.type == torch.device(dev).type
        assert torch.equal(b["loc_data"]["raw_pc"].cpu(), ref["loc_data"]["raw_pc"])
        assert torch.equal(b["transforms"]["T_ml_init"].cpu(), ref["transforms"]["T_ml_init"])
        assert torch.equal(b["list"][0].cpu(), ref["list"][0]) and torch.equal(b["list"][1][0].cpu(), ref["list"][1][0])
        assert isinstance(b["list"][1], tuple)


def test_prefetcher_passes_host_batches_through_on_cpu():
    _check("cpu")
    assert list(DevicePrefetcher(iter(()), "cpu")) == []


@pytest.mark.gpu
def test_prefetcher_uploads_each_batch_once_and_in_order():
    _check("cuda:0")
    host = [pin_batch(b) for b in _batches(3)]
    assert host[0]["loc_data"]["raw_pc"].is_pinned()
    pre = DevicePrefetcher(iter(host), "cuda:0")
    per = sum(t.numel() * t.element_size() for t in (host[0]["loc_data"]["raw_pc"], host[0]["transforms"]["T_ml_init"],
                                                      host[0]["list"][0], host[0]["list"][1][0]))
    assert pre.bytes_uploaded == per                      # first batch in flight, nothing else yet
    # consume under load on the compute stream: results must still be the host values
    junk = torch.empty(64 << 20, dtype=torch.uint8, device="cuda:0")
    for k, b in enumerate(pre):
        junk.fill_(k)
        assert torch.equal((b["list"][0] * 2).cpu(), host[k]["list"][0] * 2)
    assert pre.bytes_uploaded == 3 * per
