"""On-disk formats either side of the hot path, written so the reference's downstream scripts read
our output unchanged.

  frame store   {store}_emb.dat  float32 (n, 768) row-major memmap
                {store}_paths.npy  object array of frame paths
                {store}_meta.npz   n_frames, emb_dim
                writer  nba_proj/ratt_write_clips_to_ragdb.py:165-254 (build_frame_store)
                readers nba_proj/ratt_write_clips_to_ragdb.py:236-254, train/train_chunk_encoder.py:36-58
  label banks   {left,right,none}_embeddings.npz: embeddings (N,1,768) float32, frame_ids <U
                writer  nba_proj/write_embeddings.py:187-188,240-242
                reader  nba_proj/clustering.py:26-36 (knn.load_label_banks here)

Embeddings come from the streaming host API (ViTEmbedder.embed_and_vote_stream): uploads, the
forward and downloads of consecutive batches overlap, and each batch lands directly in the memmap.
"""
from __future__ import annotations

import os
from concurrent.futures import ThreadPoolExecutor

import numpy as np

from ._lib import VitkError

CACHE_DIR = "./frame_cache"  # ratt_write_clips_to_ragdb.py:31


def get_store_paths(store_name, cache_dir=None):
    """ratt_write_clips_to_ragdb.py:165-170."""
    d = CACHE_DIR if cache_dir is None else cache_dir
    return {
        "emb": os.path.join(d, f"{store_name}_emb.dat"),
        "paths": os.path.join(d, f"{store_name}_paths.npy"),
        "meta": os.path.join(d, f"{store_name}_meta.npz"),
    }


def frame_store_exists(store_name, cache_dir=None):
    """ratt_write_clips_to_ragdb.py:173-179."""
    p = get_store_paths(store_name, cache_dir)
    return all(os.path.exists(p[k]) for k in ("emb", "paths", "meta"))


def load_one_image(path):
    """ratt_write_clips_to_ragdb.py:122-124 (PIL, RGB, uint8)."""
    from PIL import Image
    return np.array(Image.open(path).convert("RGB"), dtype=np.uint8)


def load_frame_batch_parallel(frame_paths, executor, loader=load_one_image):
    """ratt_write_clips_to_ragdb.py:127-129."""
    return np.stack(list(executor.map(loader, frame_paths)), axis=0)


def write_frame_store(store_name, embeddings, frame_paths, cache_dir=None):
    """Write an (n, d) float32 embedding matrix (numpy, or a CUDA/CPU torch tensor) plus its frame
    paths in the reference's three-file layout."""
    p = get_store_paths(store_name, cache_dir)
    os.makedirs(os.path.dirname(p["emb"]) or ".", exist_ok=True)
    if hasattr(embeddings, "detach"):
        embeddings = embeddings.detach().float().cpu().numpy()
    e = np.ascontiguousarray(embeddings, dtype=np.float32)
    if e.ndim != 2 or e.shape[0] != len(frame_paths):
        raise VitkError(f"embeddings {e.shape} do not match {len(frame_paths)} frame paths")
    n, d = e.shape
    mm = np.memmap(p["emb"], dtype="float32", mode="w+", shape=(max(n, 1), d))
    mm[:n] = e
    mm.flush()
    del mm
    if n == 0:
        open(p["emb"], "wb").close()
    np.save(p["paths"], np.array(list(frame_paths), dtype=object))
    np.savez_compressed(p["meta"], n_frames=n, emb_dim=d)
    return p


def build_frame_store(frame_paths, batch_size=512, num_workers=16, store_name="train_val_frames",
                      embedder=None, cache_dir=None, loader=load_one_image):
    """Drop-in for build_frame_store (ratt_write_clips_to_ragdb.py:182-233): load frames with a
    thread pool, embed (L2-normalised rows, like hf_vit_embed_batch), write the memmap batch by
    batch.  `embedder` defaults to the module-global installed by vit_research_b200.install().
    Frames arrive at the VIDEO's resolution (PIL .convert("RGB"), :197-206): each batch is
    uploaded raw and the processor's resize runs on the GPU inside the overlapped streaming
    pipeline (vitk_stream_submit_raw), so real frame folders take the same path as 224x224 ones.
    (Like the reference's np.stack, a batch must hold frames of one size.)"""
    from . import embed as _embed
    emb = embedder if embedder is not None else _embed._default
    if emb is None:
        raise VitkError("call vit_research_b200.install(ViTEmbedder...) first")
    p = get_store_paths(store_name, cache_dir)
    os.makedirs(os.path.dirname(p["emb"]) or ".", exist_ok=True)
    n, d = len(frame_paths), emb.geometry.hidden
    step = min(int(batch_size), emb.max_batch)
    mm = np.memmap(p["emb"], dtype="float32", mode="w+", shape=(max(n, 1), d))
    with ThreadPoolExecutor(max_workers=num_workers) as ex:
        batches = (load_frame_batch_parallel(frame_paths[a:a + step], ex, loader)
                   for a in range(0, n, step))
        a = 0
        for i, (e, *_rest) in enumerate(emb.embed_and_vote_stream(batches)):
            mm[a:a + e.shape[0]] = e
            a += e.shape[0]
            if (i + 1) % 10 == 0:
                mm.flush()
    mm.flush()
    del mm
    if n == 0:
        open(p["emb"], "wb").close()
    np.save(p["paths"], np.array(list(frame_paths), dtype=object))
    np.savez_compressed(p["meta"], n_frames=n, emb_dim=d)
    return p


def load_frame_store(store_name="train_val_frames", cache_dir=None):
    """ratt_write_clips_to_ragdb.py:236-254: (memmap (n,d) float32, frame_paths, path_to_idx)."""
    p = get_store_paths(store_name, cache_dir)
    meta = np.load(p["meta"])
    n, d = int(meta["n_frames"]), int(meta["emb_dim"])
    mm = np.memmap(p["emb"], dtype="float32", mode="r", shape=(n, d)) if n else np.zeros((0, d), np.float32)
    frame_paths = np.load(p["paths"], allow_pickle=True).tolist()
    return mm, frame_paths, {q: i for i, q in enumerate(frame_paths)}


def write_label_bank(path, embeddings, frame_ids):
    """One `*_embeddings.npz` in the reference's layout (write_embeddings.py:187-188,240-242):
    embeddings (N,1,768) float32 — the TF model's pre_logits reshaped — and frame_ids as a
    unicode array."""
    if hasattr(embeddings, "detach"):
        embeddings = embeddings.detach().float().cpu().numpy()
    e = np.ascontiguousarray(embeddings, dtype=np.float32)
    e = e.reshape(e.shape[0], 1, -1)
    ids = np.array([str(x) for x in frame_ids])
    if len(ids) != e.shape[0]:
        raise VitkError(f"{len(ids)} frame ids for {e.shape[0]} embeddings")
    np.savez(path, embeddings=e, frame_ids=ids)
    return path


def write_label_banks(directory, embeddings, labels, frame_ids, names=("left", "right", "none")):
    """Split (N,d) embeddings by integer label into the three bank files clustering.py reads."""
    os.makedirs(directory, exist_ok=True)
    if hasattr(embeddings, "detach"):
        embeddings = embeddings.detach().float().cpu().numpy()
    labels = np.asarray(labels)
    out = []
    for li, name in enumerate(names):
        sel = np.nonzero(labels == li)[0]
        out.append(write_label_bank(os.path.join(directory, f"{name}_embeddings.npz"),
                                    np.asarray(embeddings)[sel], [frame_ids[i] for i in sel]))
    return out
