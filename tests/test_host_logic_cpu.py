"""Host-side helpers of the Python mirror (no GPU)."""


def test_zero_outputs_are_aligned_views_of_one_block():
    """`ICP.py` carves its zero-initialised outputs from one buffer (one fill launch)"""
    import torch
    from mm_masking_b200.ops import _zeros_block
    specs = (((3, 10, 6), torch.float64), ((3, 10, 7), torch.float32), ((3,), torch.uint8), ((0, 4), torch.float32),
             ((1,), torch.int32))
    outs = _zeros_block("cpu", specs)
    base = outs[0].data_ptr()
    for o, (shape, dt) in zip(outs, specs):
        assert tuple(o.shape) == shape and o.dtype == dt and o.is_contiguous()
        assert o.numel() == 0 or ((o.data_ptr() - base) % 256 == 0 and float(o.double().abs().sum()) == 0.0)
    outs[0].fill_(1)                                   # no overlap
    assert all(float(o.double().abs().sum()) == 0.0 for o in outs[1:] if o.numel())
    assert [tuple(o.shape) for o in _zeros_block("cpu", (((0, 5, 6), torch.float32), ((0,), torch.uint8)))] == [(0, 5, 6), (0,)]
