"""world_size-2 gloo tests (CPU) of the data-parallel plumbing around the ICP path."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from mm_masking_b200 import dist as mdist


def test_shard_range_covers_the_batch_once():
    for total in (0, 1, 7, 32, 64, 257):
        for ws in (1, 2, 3, 4, 8):
            spans = [mdist.shard_range(total, ws, r) for r in range(ws)]
            assert spans[0][0] == 0 and spans[-1][1] == total
            assert all(spans[i][1] == spans[i + 1][0] for i in range(ws - 1))
            sizes = [e - s for s, e in spans]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        mdist.shard_range(4, 2, 2)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, ws, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(ws))
    dist.init_process_group("gloo", rank=rank, world_size=ws)
    try:
        assert mdist.world() == (rank, ws)
        # per-rank "timings": the job time is the slowest rank's
        t = mdist.max_over_ranks(10.0 + rank)
        # every rank holds the full synthetic batch and keeps its own pairs
        batch = {"source": np.arange(10 * 3).reshape(10, 3), "weight": np.arange(10)}
        mine = mdist.shard_batch(batch, ws, rank)
        # gradient averaging through one flat buffer
        g = [torch.full((3, 2), float(rank + 1)), None, torch.arange(4, dtype=torch.float32) * (rank + 1)]
        mdist.allreduce_mean_(g)
        out[rank] = (t, mine["weight"].tolist(), g[0].tolist(), g[2].tolist())
    finally:
        dist.destroy_process_group()


def test_gloo_world_size_2():
    ws = 2
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(ws, _free_port(), out), nprocs=ws, join=True)
    assert out[0][0] == out[1][0] == 11.0
    assert out[0][1] == [0, 1, 2, 3, 4] and out[1][1] == [5, 6, 7, 8, 9]
    for r in range(ws):
        assert out[r][2] == [[1.5, 1.5]] * 3
        assert out[r][3] == [0.0, 1.5, 3.0, 4.5]


def test_single_process_is_a_no_op():
    assert mdist.world() == (0, 1)
    assert mdist.max_over_ranks(3.5) == 3.5
    g = [torch.ones(2)]
    mdist.allreduce_mean_(g)
    assert g[0].tolist() == [1.0, 1.0]


def _policy_worker(rank, ws, port, out):
    """each rank runs the UNet on its half of the batch; the input normalisation must use the
    min/max of the WHOLE batch (`icp_weight_policy.py:150-155`)"""
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(ws))
    dist.init_process_group("gloo", rank=rank, world_size=ws)
    try:
        from mm_masking_b200.icp_weight_policy import LearnICPWeightPolicy
        torch.manual_seed(3)
        model = LearnICPWeightPolicy({"device": "cpu", "range_input": False, "max_iter": 2}).eval()
        g = torch.Generator().manual_seed(5)
        fft = torch.rand((4, 32, 32), generator=g) * torch.tensor([1.0, 5.0, 0.2, 3.0]).view(4, 1, 1)
        cfar = (torch.rand((4, 32, 32), generator=g) > 0.8).float()
        s, e = mdist.shard_range(4, ws, rank)
        scan = {"fft_data": fft[s:e].clone(), "fft_cfar": cfar[s:e].clone(), "raw_pc": torch.zeros(e - s, 2, 3),
                "filtered_pc": torch.zeros(e - s, 2, 3)}
        with torch.no_grad():
            mask = model(scan, {"pc": torch.zeros(e - s, 2, 6)}, torch.eye(4).repeat(e - s, 1, 1), mask_only=True)
        for norm in ("standardize",):
            model.normalize_type = norm
            with torch.no_grad():
                m2 = model(scan, {"pc": torch.zeros(e - s, 2, 6)}, torch.eye(4).repeat(e - s, 1, 1), mask_only=True)
        out[rank] = (mask.numpy(), m2.numpy())
    finally:
        dist.destroy_process_group()


def test_sharded_policy_forward_equals_full_batch():
    ws = 2
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_policy_worker, args=(ws, _free_port(), out), nprocs=ws, join=True)
    from mm_masking_b200.icp_weight_policy import LearnICPWeightPolicy
    torch.manual_seed(3)
    model = LearnICPWeightPolicy({"device": "cpu", "range_input": False, "max_iter": 2}).eval()
    g = torch.Generator().manual_seed(5)
    fft = torch.rand((4, 32, 32), generator=g) * torch.tensor([1.0, 5.0, 0.2, 3.0]).view(4, 1, 1)
    cfar = (torch.rand((4, 32, 32), generator=g) > 0.8).float()
    scan = {"fft_data": fft, "fft_cfar": cfar, "raw_pc": torch.zeros(4, 2, 3), "filtered_pc": torch.zeros(4, 2, 3)}
    args = (scan, {"pc": torch.zeros(4, 2, 6)}, torch.eye(4).repeat(4, 1, 1))
    with torch.no_grad():
        full = model(*args, mask_only=True).numpy()
        model.normalize_type = "standardize"
        full_std = model(*args, mask_only=True).numpy()
    np.testing.assert_allclose(np.concatenate([out[0][0], out[1][0]]), full, atol=1e-6)
    np.testing.assert_allclose(np.concatenate([out[0][1], out[1][1]]), full_std, atol=1e-5)


def _toy(seed=0):
    torch.manual_seed(seed)
    return torch.nn.Sequential(torch.nn.Linear(6, 16), torch.nn.Tanh(), torch.nn.Linear(16, 16), torch.nn.Tanh(),
                               torch.nn.Linear(16, 1), torch.nn.Linear(1, 1))


def _toy_data():
    g = torch.Generator().manual_seed(9)
    return torch.randn((5, 6), generator=g), torch.randn((5, 1), generator=g)


def _uneven_worker(rank, ws, port, out):
    """batch of 5 over 2 ranks (3 + 2): batch-MEAN loss per rank, weighted so that the mean of the gradients
    over ranks is the gradient of the mean over the whole batch; gradients reduced by the bucketed, hook-driven
    reducer; rank 1 leaves the last layer out of its graph (no gradient there on that rank)."""
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(ws))
    dist.init_process_group("gloo", rank=rank, world_size=ws)
    try:
        model = _toy()
        x, y = _toy_data()
        s, e = mdist.shard_range(5, ws, rank)
        red = mdist.BucketedGradReducer(list(model.parameters()), bucket_bytes=64)
        assert len(red.buckets) >= 3
        for step in range(2):                                        # twice: the reducer re-arms itself
            model.zero_grad(set_to_none=True)
            loss = ((model(x[s:e]) - y[s:e]) ** 2).mean()
            (loss * mdist.shard_loss_scale(e - s)).backward()
            red.finish()
        g = torch.cat([p.grad.reshape(-1) for p in model.parameters()])
        out[rank] = (g.numpy(), float(mdist.global_mean(loss.detach(), e - s)))
        # a parameter with no gradient on one rank: the collectives must still line up
        model.zero_grad(set_to_none=True)
        body, head = model[:-1], model[-1]
        pred = body(x[s:e]) if rank == 1 else head(body(x[s:e]))
        ((pred - y[s:e]) ** 2).mean().backward()
        red.finish()
        assert all(p.grad is not None for p in model.parameters())
        out[10 + rank] = head.weight.grad.numpy().copy()
    finally:
        dist.destroy_process_group()


def test_unequal_shards_reproduce_the_full_batch_gradient():
    ws = 2
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_uneven_worker, args=(ws, _free_port(), out), nprocs=ws, join=True)
    model = _toy()
    x, y = _toy_data()
    loss = ((model(x) - y) ** 2).mean()
    loss.backward()
    g = torch.cat([p.grad.reshape(-1) for p in model.parameters()]).numpy()
    for r in range(ws):
        np.testing.assert_allclose(out[r][0], g, rtol=1e-5, atol=1e-7)
        assert abs(out[r][1] - float(loss)) < 1e-6
    np.testing.assert_allclose(out[10], out[11])                     # both ranks hold the same mean
