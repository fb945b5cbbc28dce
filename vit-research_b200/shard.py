"""Frame sharding across the GPUs of one box.

Frames are independent once the banks are frozen, so the multi-GPU form of the hot
path is: one process per GPU, weights and banks replicated, each rank embeds + votes a CONTIGUOUS
range of the frame list (contiguous because the downstream HMM, nba_proj/hmm.py, consumes frames
in time order), and one all-gather of the per-frame results at the end.  There is no collective
inside the forward.  The functions here are backend-agnostic: the same code runs over NCCL with
CUDA tensors (bench.py, N > 1) and over gloo with CPU tensors (tests/test_shard_cpu.py).
"""
from __future__ import annotations

from typing import Iterator, List, Sequence, Tuple

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
