"""Frame sharding across the GPUs of one box.

Frames are independent once the banks are frozen, so the multi-GPU form of the hot
path is: one process per GPU, weights and banks replicated, each rank embeds + votes a CONTIGUOUS
range of the frame list (contiguous because the downstream HMM, nba_proj/hmm.py, consumes frames
in time order), and one all-gather of the per-frame results at the end.  There is no collective
inside the forward.  The functions here are backend-agnostic: the same code runs over NCCL with
CUDA tensors (bench.py, N > 1) and over gloo with CPU tensors (tests/test_shard_cpu.py).
"""
from __future__ import annotations

from typing import Callable, Iterator, List, Sequence, Tuple

import numpy as np
import torch


def frame_ranges(n_frames: int, world: int) -> List[Tuple[int, int]]:
    """[start, stop) per rank: contiguous, in rank order, sizes differ by at most one frame
    (rank r of R owns [r*N/R, (r+1)*N/R)).  Ranks beyond the frame count get empty
    ranges."""
    if n_frames < 0 or world < 1:
        raise ValueError("need n_frames >= 0 and world >= 1")
    return [((r * n_frames) // world, ((r + 1) * n_frames) // world) for r in range(world)]


def batch_ranges(start: int, stop: int, batch: int) -> Iterator[Tuple[int, int]]:
    """Split one rank's range into [a, b) batches of at most `batch` frames (last one ragged)."""
    if batch < 1:
        raise ValueError("batch must be >= 1")
    a = start
    while a < stop:
        b = min(a + batch, stop)
        yield a, b
        a = b


def gather_frames(local: torch.Tensor, n_frames: int, group=None) -> torch.Tensor:
    """All-gather per-frame rows that were sharded with frame_ranges(): `local` is this rank's
    (n_local, ...) tensor; returns the (n_frames, ...) tensor in frame order on every rank.
    Ranges may differ by one row, so shards are padded to the largest and trimmed after the
    collective (one all_gather, no per-rank sends)."""
    import torch.distributed as dist

    if not (dist.is_available() and dist.is_initialized()):
        if local.shape[0] != n_frames:
            raise ValueError("single process: local shard must hold every frame")
        return local
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    ranges = frame_ranges(n_frames, world)
    a, b = ranges[rank]
    if local.shape[0] != b - a:
        raise ValueError(f"rank {rank}: shard has {local.shape[0]} rows, its range holds {b - a}")
    width = max(e - s for s, e in ranges)
    pad = torch.zeros((width,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    pad[: b - a] = local
    out = torch.empty((world * width,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    dist.all_gather_into_tensor(out, pad, group=group)
    if all(e - s == width for s, e in ranges):
        return out
    return torch.cat([out[r * width: r * width + (e - s)] for r, (s, e) in enumerate(ranges)], 0)


def label_frames_sharded(embedder, index, frames: Sequence, k: int = 25, batch: int | None = None,
                         device: str | None = None):
    """Embed + vote this rank's share of `frames` (an indexable of uint8 (H,W,C) frames or a
    (N,H,W,C) array) through the streaming API and all-gather (embeddings, labels) in frame order.
    Returns (emb (N,hidden) f32, labels (N,) i32) tensors on `device` (default: cuda if the
    process group is NCCL, else cpu)."""
    import torch.distributed as dist

    n = len(frames)
    world = dist.get_world_size() if dist.is_initialized() else 1
    rank = dist.get_rank() if dist.is_initialized() else 0
    a, b = frame_ranges(n, world)[rank]
    batch = batch or embedder.max_batch
    if device is None:
        device = "cuda" if dist.is_initialized() and dist.get_backend() == "nccl" else "cpu"
    embs, labs = [], []
    gen = (np.ascontiguousarray(np.asarray(frames[s:e], dtype=np.uint8)) for s, e in batch_ranges(a, b, batch))
    for emb, label, _idx, _score, _counts in embedder.embed_and_vote_stream(gen, index, k):
        embs.append(torch.from_numpy(emb.copy()))
        labs.append(torch.from_numpy(label.copy()))
    hidden = embedder.geometry.hidden
    emb_t = torch.cat(embs, 0) if embs else torch.zeros((0, hidden), dtype=torch.float32)
    lab_t = torch.cat(labs, 0) if labs else torch.zeros((0,), dtype=torch.int32)
    return gather_frames(emb_t.to(device), n), gather_frames(lab_t.to(device), n)


# ---------------------------------------------------------------------------------------------
# One long game sharded BY CHUNK with the smoother running during the stream (BASELINE configs[2])
# ---------------------------------------------------------------------------------------------
def round_robin_chunks(n_frames: int, chunk: int, world: int) -> List[List[Tuple[int, int, int]]]:
    """Chunks of `chunk` consecutive frames dealt to the ranks in turn: chunk b = frames
    [b*chunk, min((b+1)*chunk, n_frames)) belongs to rank b % world.  Returns, per rank, its
    (chunk index, start, stop) list in time order.  After every rank has finished its s-th chunk the
    frames [0, (s+1)*world*chunk) are complete -- a contiguous prefix in time, which is what lets
    the sequential Viterbi smoother (nba_proj/hmm.py) advance while the stream is still running."""
    if n_frames < 0 or chunk < 1 or world < 1:
        raise ValueError("need n_frames >= 0, chunk >= 1 and world >= 1")
    out: List[List[Tuple[int, int, int]]] = [[] for _ in range(world)]
    n_chunks = (n_frames + chunk - 1) // chunk
    for b in range(n_chunks):
        out[b % world].append((b, b * chunk, min((b + 1) * chunk, n_frames)))
    return out


def time_order_of_round(n_frames: int, chunk: int, world: int, first_step: int, steps: int):
    """For an all-gather of every rank's local steps [first_step, first_step + steps): the
    (rank, local step offset, valid frames) triples in TIME order, skipping chunks past the end."""
    n_chunks = (n_frames + chunk - 1) // chunk
    order = []
    for g in range(steps):
        for r in range(world):
            b = (first_step + g) * world + r
            if b < n_chunks:
                order.append((r, g, min(chunk, n_frames - b * chunk)))
    return order


def gather_round_in_time_order(local: torch.Tensor, s0: int, s1: int, n_frames: int, chunk: int,
                               gather_every: int, scratch: torch.Tensor | None = None, group=None):
    """One gather of the round-robin stream: `local` is this rank's (rounds, chunk, ...) per-chunk
    result buffer; local steps [s0, s1) (at most gather_every of them) of EVERY rank are all-gathered
    with one fixed-shape collective and returned as the list of per-chunk tensors in TIME order
    (ragged last chunk trimmed, chunks past the end skipped).  Single process: slices of `local`."""
    import torch.distributed as dist

    dist_on = dist.is_available() and dist.is_initialized()
    world = dist.get_world_size(group) if dist_on else 1
    n = s1 - s0
    if world == 1:
        src = local[s0:s1].unsqueeze(0)
    else:
        send = local[s0:s0 + gather_every]
        if send.shape[0] < gather_every:          # last, short round: pad to the fixed shape
            pad = torch.zeros((gather_every,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
            pad[: send.shape[0]] = send
            send = pad
        if scratch is None:
            scratch = torch.empty((world, gather_every) + tuple(local.shape[1:]), dtype=local.dtype,
                                  device=local.device)
        dist.all_gather_into_tensor(scratch.view((world * gather_every,) + tuple(local.shape[1:])),
                                    send.contiguous(), group=group)
        src = scratch
    return [src[r, g, :valid] for r, g, valid in time_order_of_round(n_frames, chunk, world, s0, n)]


def label_and_smooth_stream(embedder, index, load_chunk: Callable, n_frames: int, k: int = 25,
                            chunk: int | None = None, gather_every: int = 4, stats: dict | None = None,
                            outs=None):
    """Embed + vote + Viterbi-smooth ONE long frame sequence on all ranks (the per-frame loop of
    generate_clips_hmm.py:367-490 for a whole game).  Chunks of `chunk` frames are dealt to the
    ranks round-robin (round_robin_chunks); each rank streams its chunks through the overlapped
    host API; every `gather_every` local steps the ranks all-gather those steps' (chunk,3) vote
    counts (NCCL / gloo: 3 KB per chunk) and rank 0 feeds them, in time order, to the streamed
    Viterbi (ViterbiStream), which therefore finishes a few milliseconds after the last chunk
    instead of after a serial pass over the whole game.
    load_chunk(start, stop) -> uint8 (stop-start, H, W, C) frames (any size; pinned for overlap).
    `outs`: optional pair of preallocated (pinned) output tuples for embed_and_vote_stream.
    Returns (states CUDA int32 (n_frames,) on rank 0 else None, labels of this rank's frames)."""
    import torch.distributed as dist
    from .hmm import ViterbiStream

    dist_on = dist.is_available() and dist.is_initialized()
    world = dist.get_world_size() if dist_on else 1
    rank = dist.get_rank() if dist_on else 0
    chunk = int(chunk or embedder.max_batch)
    mine = round_robin_chunks(n_frames, chunk, world)[rank]
    rounds = ((n_frames + chunk - 1) // chunk + world - 1) // world      # local steps of the busiest rank
    dev = f"cuda:{embedder.device}"
    counts_dev = torch.zeros((max(rounds, 1), chunk, 3), dtype=torch.int32, device=dev)
    gathered = torch.empty((world, gather_every, chunk, 3), dtype=torch.int32, device=dev) if world > 1 else None
    vit = ViterbiStream(embedder, max(n_frames, 1), k) if rank == 0 else None
    labels = []
    t_wait = 0.0

    def flush(s0, s1):
        """Counts of local steps [s0, s1) of every rank -> rank 0's lattice, in time order."""
        for seg in gather_round_in_time_order(counts_dev, s0, s1, n_frames, chunk, gather_every, gathered):
            if rank == 0:
                vit.append(seg)

    import time as _time
    batches = (load_chunk(a, b) for _, a, b in mine)
    done = 0
    stream_iter = embedder.embed_and_vote_stream(batches, index, k, outs=outs)
    for s in range(rounds):
        if s < len(mine):
            t0 = _time.perf_counter()
            _emb, lab, _idx, _sc, cnt = next(stream_iter)
            t_wait += _time.perf_counter() - t0
            labels.append(lab.copy())
            counts_dev[s, : cnt.shape[0]].copy_(torch.from_numpy(cnt), non_blocking=False)
        done = s + 1
        if done % gather_every == 0:
            flush(done - gather_every, done)
    if done % gather_every:
        flush(done - done % gather_every, done)
    for _ in stream_iter:      # exhausted; closes the generator
        pass
    states = vit.finish(device=dev) if rank == 0 else None
    if stats is not None:
        stats.update(wait_s=t_wait, local_steps=len(mine), rounds=rounds)
    lab_all = np.concatenate(labels) if labels else np.zeros((0,), np.int32)
    return states, lab_all


# ---------------------------------------------------------------------------------------------
# Row-sharded clip store (the "exchange step" alternative for BASELINE configs[4]):
# rank r holds store rows [row_ranges[r]) and its own share of the queries.
# ---------------------------------------------------------------------------------------------
def merge_topk(idx: torch.Tensor, score: torch.Tensor, k: int, descending: bool = True):
    """Merge per-shard candidate lists: idx/score (Q, C) with GLOBAL row ids (-1 = empty slot) ->
    (Q, k) best in (score, then lower row id) order — the same total order the kernels use.
    CUDA tensors go through vitk_op_merge_topk (C <= 4096); CPU tensors through torch."""
    q, c = idx.shape
    if idx.is_cuda:
        # the product path: one warp-sort kernel next to the re-rank kernels (csrc/knn.cuh,
        # knn_merge_topk_kernel) -- no library sort / gather on the device
        from . import _lib
        lib = _lib.load()
        idx_c = idx.to(torch.int32).contiguous()
        score_c = score.to(torch.float32).contiguous()
        out_idx = torch.empty((q, k), dtype=torch.int32, device=idx.device)
        out_score = torch.empty((q, k), dtype=torch.float32, device=idx.device)
        rc = lib.vitk_op_merge_topk(_lib.ptr(idx_c), _lib.ptr(score_c), q, c, int(k), int(bool(descending)),
                                    _lib.ptr(out_idx), _lib.ptr(out_score), _lib.stream_ptr(None))
        _lib.check(rc, None, "vitk_op_merge_topk")
        return out_idx.to(idx.dtype), out_score.to(score.dtype)
    # host tensors (the gloo tests of the sharding logic): plain torch
    big = torch.iinfo(torch.int64).max
    rid = torch.where(idx < 0, torch.full_like(idx, big, dtype=torch.int64), idx.to(torch.int64))
    key = score.to(torch.float64)
    key = torch.where(idx < 0, torch.full_like(key, float("-inf") if descending else float("inf")), key)
    # stable two-pass sort: by row id, then by score
    o1 = torch.argsort(rid, dim=1, stable=True)
    key1 = torch.gather(key, 1, o1)
    o2 = torch.argsort(key1, dim=1, descending=descending, stable=True)
    order = torch.gather(o1, 1, o2)[:, :k]
    out_idx = torch.gather(idx, 1, order)
    out_score = torch.gather(score, 1, order)
    if c < k:
        pad = k - c
        out_idx = torch.cat([out_idx, torch.full((q, pad), -1, dtype=idx.dtype, device=idx.device)], 1)
        out_score = torch.cat([out_score, torch.zeros((q, pad), dtype=score.dtype, device=score.device)], 1)
    return out_idx, out_score


def search_sharded_store(local_search: Callable, q_local: torch.Tensor, k: int, row_offset: int,
                         n_queries: int, descending: bool = True, group=None):
    """Top-k of this rank's queries against a store whose rows are sharded over the ranks.

    local_search(q (Q,d) tensor, k) -> (idx (Q,k) int32 LOCAL rows, score (Q,k) f32): the rank's
    BankIndex.search on its shard (`lambda q, k: index.search(q, k)[:2]`).  One exchange step:
    all-gather the queries (n_queries rows in total, sharded with frame_ranges), search the local
    shard for ALL of them, all-gather the (global row, score) lists — 8·k bytes per query per rank —
    and merge world·k -> k for the rank's own queries.  Returns (idx (Q_local,k) global rows, score).
    """
    import torch.distributed as dist

    if not (dist.is_available() and dist.is_initialized()):
        idx, score = local_search(q_local, k)
        idx = torch.where(idx >= 0, idx + row_offset, idx)
        return merge_topk(idx, score, k, descending)
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    q_all = gather_frames(q_local, n_queries, group)
    idx, score = local_search(q_all, k)
    idx = torch.where(idx >= 0, idx + row_offset, idx).to(torch.int32).contiguous()
    score = score.to(torch.float32).contiguous()
    nq_all = idx.shape[0]
    all_idx = torch.empty((world * nq_all, k), dtype=idx.dtype, device=idx.device)
    all_score = torch.empty((world * nq_all, k), dtype=score.dtype, device=score.device)
    dist.all_gather_into_tensor(all_idx, idx, group=group)      # rank-major concatenation
    dist.all_gather_into_tensor(all_score, score, group=group)
    all_idx = all_idx.view(world, nq_all, k)
    all_score = all_score.view(world, nq_all, k)
    a, b = frame_ranges(n_queries, world)[rank]
    mine_idx = all_idx[:, a:b].permute(1, 0, 2).reshape(b - a, world * k)
    mine_score = all_score[:, a:b].permute(1, 0, 2).reshape(b - a, world * k)
    return merge_topk(mine_idx, mine_score, k, descending)
