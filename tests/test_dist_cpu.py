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
