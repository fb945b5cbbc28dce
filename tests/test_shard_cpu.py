"""Host-side sharding logic on CPU: range arithmetic, and the N > 1 gather path
over the gloo backend with world_size 2 and 3 (the GPU path runs the same code over NCCL)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from vit_research_b200 import shard


@pytest.mark.parametrize("n,world", [(0, 1), (1, 8), (7, 8), (8, 8), (108000, 8), (108001, 4), (255, 2)])
def test_frame_ranges_cover_in_order(n, world):
    r = shard.frame_ranges(n, world)
    assert len(r) == world and r[0][0] == 0 and r[-1][1] == n
    for (a, b), (c, d) in zip(r, r[1:]):
        assert b == c and a <= b and c <= d
    sizes = [b - a for a, b in r]
    assert max(sizes) - min(sizes) <= 1 and sum(sizes) == n


def test_batch_ranges_ragged_tail():
    assert list(shard.batch_ranges(10, 10, 4)) == []
    assert list(shard.batch_ranges(3, 14, 4)) == [(3, 7), (7, 11), (11, 14)]
    with pytest.raises(ValueError):
        list(shard.batch_ranges(0, 1, 0))


def test_gather_single_process_is_identity():
    x = torch.arange(12.).reshape(6, 2)
    assert shard.gather_frames(x, 6) is x
    with pytest.raises(ValueError):
        shard.gather_frames(x, 7)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, n_frames, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        a, b = shard.frame_ranges(n_frames, world)[rank]
        # "embedding" of frame i = [i, 2i, 3i], label = i % 3: a pure function of the global index
        idx = torch.arange(a, b, dtype=torch.float32)
        emb = torch.stack([idx, 2 * idx, 3 * idx], 1)
        lab = (torch.arange(a, b) % 3).to(torch.int32)
        g_emb = shard.gather_frames(emb, n_frames)
        g_lab = shard.gather_frames(lab, n_frames)
        want = torch.arange(n_frames, dtype=torch.float32)
        ok = (g_emb.shape == (n_frames, 3) and torch.equal(g_emb[:, 0], want)
              and torch.equal(g_emb[:, 2], 3 * want)
              and torch.equal(g_lab, (torch.arange(n_frames) % 3).to(torch.int32)))
        # a shard of the wrong size must be rejected, not silently mis-ordered
        try:
            shard.gather_frames(emb[:-1] if emb.shape[0] else torch.zeros(1, 3), n_frames)
            ok = False
        except ValueError:
            pass
        q.put((rank, bool(ok)))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,n_frames", [(2, 11), (2, 10), (3, 2), (3, 100)])
def test_gather_over_gloo(world, n_frames):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n_frames, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    assert sorted(res) == [(r, True) for r in range(world)]
