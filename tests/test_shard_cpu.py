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


# ---------------------------------------------------------------------------------------------
# row-sharded store: the exchange step (all-gather queries -> local top-k -> all-gather -> merge)
# ---------------------------------------------------------------------------------------------
def test_merge_topk_orders_by_score_then_row():
    idx = torch.tensor([[7, 3, -1, 9, 1, 5]], dtype=torch.int32)
    sc = torch.tensor([[0.5, 0.9, 0.0, 0.5, 0.5, 0.1]])
    i, s = shard.merge_topk(idx, sc, 4)
    assert i.tolist() == [[3, 1, 7, 9]] and torch.allclose(s, torch.tensor([[0.9, 0.5, 0.5, 0.5]]))
    i, s = shard.merge_topk(idx, -sc, 3, descending=False)          # distances: ascending
    assert i.tolist() == [[3, 1, 7]]
    i, s = shard.merge_topk(idx[:, :2], sc[:, :2], 4)               # fewer candidates than k: -1 padded
    assert i.tolist() == [[3, 7, -1, -1]]


def _store_worker(rank, world, port, q):
    import numpy as np
    from oracle import knn_ref
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        rng = np.random.default_rng(0)
        m, nq, k, d = 1000, 37, 10, 64
        store = knn_ref.normalise(rng.standard_normal((m, d)).astype(np.float32))
        queries = knn_ref.normalise(rng.standard_normal((nq, d)).astype(np.float32))
        ra, rb = shard.frame_ranges(m, world)[rank]       # this rank's store rows
        qa, qb = shard.frame_ranges(nq, world)[rank]      # this rank's queries

        def local_search(qq, kk):  # the checker stands in for BankIndex.search on the CPU
            i, s, _ = knn_ref.cosine_topk(qq.numpy(), store[ra:rb], kk, dtype=np.float32)
            return torch.from_numpy(i.astype(np.int32)), torch.from_numpy(s.astype(np.float32))

        gi, gs = shard.search_sharded_store(local_search, torch.from_numpy(queries[qa:qb]), k, ra, nq)
        wi, ws, _ = knn_ref.cosine_topk(queries[qa:qb], store, k, dtype=np.float32)
        ok = gi.shape == (qb - qa, k) and np.array_equal(gi.numpy(), wi) and \
            np.allclose(gs.numpy(), ws, atol=1e-6)
        q.put((rank, bool(ok)))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_sharded_store_exchange_over_gloo(world):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_store_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=180) for _ in procs]
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    assert sorted(res) == [(r, True) for r in range(world)]


def test_round_robin_chunks_cover_the_stream_in_time_order():
    """shard.round_robin_chunks / time_order_of_round (the sharding of BASELINE configs[2]): every
    frame belongs to exactly one rank, and replaying the gathers in rounds visits the chunks in
    time order with the right ragged tail -- for even, ragged and tiny streams."""
    for n, chunk, world, every in [(108_000, 256, 8, 4), (1000, 256, 3, 2), (5, 4, 4, 3), (0, 8, 2, 1), (256, 256, 1, 4)]:
        per_rank = shard.round_robin_chunks(n, chunk, world)
        seen = sorted((b, a, e) for r in per_rank for b, a, e in r)
        assert [x[0] for x in seen] == list(range(len(seen)))
        assert sum(e - a for _, a, e in seen) == n
        for r, lst in enumerate(per_rank):
            assert all(b % world == r for b, _, _ in lst)
        rounds = ((n + chunk - 1) // chunk + world - 1) // world
        pos, s = 0, 0
        while s < rounds:
            steps = min(every, rounds - s)
            for r, g, valid in shard.time_order_of_round(n, chunk, world, s, steps):
                b, a, e = per_rank[r][s + g]
                assert a == pos and e - a == valid       # contiguous prefix in time
                pos = e
            s += steps
        assert pos == n


def _rr_worker(rank, world, port, n_frames, chunk, every, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        mine = shard.round_robin_chunks(n_frames, chunk, world)[rank]
        rounds = ((n_frames + chunk - 1) // chunk + world - 1) // world
        # "vote counts" of frame i = [i, 2i, 3i]: a pure function of the global frame index
        local = torch.zeros((max(rounds, 1), chunk, 3), dtype=torch.int32)
        for s_, (_b, a, e) in enumerate(mine):
            idx = torch.arange(a, e, dtype=torch.int32)
            local[s_, : e - a] = torch.stack([idx, 2 * idx, 3 * idx], 1)
        got, done = [], 0
        for s_ in range(rounds):
            done = s_ + 1
            if done % every == 0:
                got += shard.gather_round_in_time_order(local, done - every, done, n_frames, chunk, every)
        if done % every:
            got += shard.gather_round_in_time_order(local, done - done % every, done, n_frames, chunk, every)
        allc = torch.cat(got) if got else torch.zeros((0, 3), dtype=torch.int32)
        want = torch.arange(n_frames, dtype=torch.int32)
        q.put((rank, bool(allc.shape == (n_frames, 3) and torch.equal(allc[:, 0], want)
                          and torch.equal(allc[:, 2], 3 * want))))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,n_frames,chunk,every", [(2, 1000, 64, 4), (3, 997, 32, 3), (2, 5, 8, 2)])
def test_round_robin_gathers_rebuild_the_stream_in_time_order_over_gloo(world, n_frames, chunk, every):
    """The N > 1 data flow of BASELINE configs[2] (shard.label_and_smooth_stream) on CPU: chunks dealt
    round-robin, fixed-shape all-gathers every few local steps, segments handed on in time order --
    every rank rebuilds the whole stream, ragged tail and short last round included."""
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_rr_worker, args=(r, world, port, n_frames, chunk, every, q)) for r in range(world)]
    for p_ in procs:
        p_.start()
    res = [q.get(timeout=120) for _ in range(world)]
    for p_ in procs:
        p_.join(60)
    assert sorted(res) == [(r, True) for r in range(world)]
