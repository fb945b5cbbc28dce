"""k-NN / vote entry points with the reference's call shapes, backed by libvitk (sm_100a).

Mirrors:
  np.load('*_embeddings.npz')  bank fixtures              nba_proj/clustering.py:26-36
  collection.query(query_embeddings=, n_results=)          nba_proj/chroma.py:211-214
  determine_class_pass1 / determine_class_pass2            nba_proj/chroma.py:36-71,102-134
  RattChunkRetriever(...)(chunk_embs) exact cosine top-k   nba_proj/retrieval/ratt_chunk_retriever.py:62-159
The banks are static (north_star; the reference's self-training upsert loop is out of scope, D4).
"""
from __future__ import annotations

import ctypes as C
from pathlib import Path

import numpy as np

from . import _lib
from ._lib import VitkError, check, ptr, stream_ptr

LABEL_NAMES = ("left", "right", "none")  # np.argmax tie priority, chroma.py:128


def load_label_banks(directory, names=LABEL_NAMES):
    """Read {left,right,none}_embeddings.npz -> (rows (M,768) f32, labels (M,) i32, ids list).
    Accepts the reference's (N,1,768) layout (write_embeddings.py:187-188,240-242) and (N,768)."""
    rows, labels, ids = [], [], []
    for li, name in enumerate(names):
        z = np.load(Path(directory) / f"{name}_embeddings.npz")
        e = np.asarray(z["embeddings"], dtype=np.float32)
        e = e.reshape(e.shape[0], -1)
        rows.append(e)
        labels.append(np.full(e.shape[0], li, np.int32))
        fid = z["frame_ids"] if "frame_ids" in z.files else np.arange(e.shape[0])
        ids += [f"{name}:{x}" for x in fid]
    return np.concatenate(rows), np.concatenate(labels), ids


def temp_smax(x, temperature=1.0):
    """chroma.py:28-34 — float64 temperature softmax of the class counts."""
    x = np.array(x, dtype=np.float64) / temperature
    e = np.exp(x - np.max(x))
    return e / e.sum()


class BankIndex:
    """A replicated bank on one GPU: top-k + vote in one call."""

    def __init__(self, embedder_or_handle, rows, labels=None, n_classes=3, metric="l2",
                 normalise_rows=False, ids=None, metadatas=None):
        self.lib = _lib.load()
        self._owner = embedder_or_handle
        self.handle = getattr(embedder_or_handle, "handle", embedder_or_handle)
        self.metric = {"l2": _lib.METRIC_L2, "cosine": _lib.METRIC_COSINE}[metric]
        self.metric_name = metric
        self.n_classes = int(n_classes) if labels is not None else 0
        self.metadatas = metadatas
        if hasattr(rows, "is_cuda"):
            # a store that already lives on the GPU (e.g. embeddings this library produced, or the
            # 1 M-row synthetic clip store of BASELINE configs[4]): no host round trip
            import torch
            if not rows.is_cuda or rows.dtype != torch.float32:
                raise VitkError("a tensor bank must be CUDA float32 (numpy arrays go through the host path)")
            rows = rows.reshape(rows.shape[0], -1).contiguous()
            self.m, self.d = rows.shape
            lab = None
            if labels is not None:
                lab = torch.as_tensor(labels, dtype=torch.int32, device=rows.device).contiguous()
                if lab.numel() and (int(lab.min()) < 0 or int(lab.max()) >= max(self.n_classes, 1)):
                    raise VitkError("vitk_bank_set_device failed: label outside [0,n_classes)")
            self.labels = None if lab is None else lab.cpu().numpy()
            check(self.lib.vitk_bank_set_device(self.handle, ptr(rows), self.m, self.d, ptr(lab),
                                                self.n_classes, int(normalise_rows), stream_ptr(None)),
                  self.handle, "vitk_bank_set_device")
        else:
            rows = np.ascontiguousarray(rows, dtype=np.float32)
            rows = rows.reshape(rows.shape[0], -1)
            self.m, self.d = rows.shape
            self.labels = None if labels is None else np.ascontiguousarray(labels, dtype=np.int32)
            check(self.lib.vitk_bank_set(self.handle, ptr(rows), self.m, self.d, ptr(self.labels),
                                         self.n_classes, int(normalise_rows)), self.handle,
                  "vitk_bank_set")
        self.ids = list(ids) if ids is not None else None

    def search(self, q, k, normalise_queries=False, stream=None):
        """q: CUDA float32 (n,d).  Returns CUDA tensors (idx i32 (n,k), score f32 (n,k),
        label i32 (n,), counts i32 (n,n_classes))."""
        import torch

        if not q.is_cuda or q.dtype != torch.float32:
            raise VitkError("search() wants a CUDA float32 tensor")
        q = q.contiguous()
        n = q.shape[0]
        idx = torch.empty((n, k), dtype=torch.int32, device=q.device)
        score = torch.empty((n, k), dtype=torch.float32, device=q.device)
        label = torch.empty((n,), dtype=torch.int32, device=q.device)
        counts = torch.zeros((n, max(self.n_classes, 1)), dtype=torch.int32, device=q.device)
        check(self.lib.vitk_knn_vote(self.handle, ptr(q), n, int(k), self.metric,
                                     int(normalise_queries), ptr(idx), ptr(score), ptr(label),
                                     ptr(counts), stream_ptr(stream)), self.handle,
              "vitk_knn_vote")
        return idx, score, label, counts


class BankCollection:
    """Chroma `Collection` look-alike over a static bank (query / get / count only).

    `hnsw:space` 'l2' returns SQUARED L2 ascending, 'cosine' returns 1 - cos ascending, like
    chromadb; results are exact brute force, not HNSW-approximate."""

    def __init__(self, embedder, rows, ids, metadatas, space="l2", label_key="label",
                 label_names=LABEL_NAMES):
        self.space = space
        self.label_names = tuple(label_names)
        labels = np.array([self.label_names.index(m[label_key]) if m and label_key in m and
                           m[label_key] in self.label_names else len(self.label_names) - 1
                           for m in metadatas], np.int32)
        self.rows = np.ascontiguousarray(rows, np.float32).reshape(len(ids), -1)
        self.index = BankIndex(embedder, self.rows, labels, len(self.label_names),
                               metric="l2" if space == "l2" else "cosine",
                               normalise_rows=(space == "cosine"), ids=ids, metadatas=metadatas)
        self.ids, self.metadatas = list(ids), list(metadatas)

    def count(self):
        return len(self.ids)

    def get(self, ids=None, where=None, include=("metadatas",)):
        sel = range(len(self.ids)) if ids is None else [self.ids.index(i) for i in ids]
        if where:
            sel = [i for i in sel if all(self.metadatas[i].get(k) == v for k, v in where.items())]
        out = {"ids": [self.ids[i] for i in sel]}
        if "metadatas" in include:
            out["metadatas"] = [self.metadatas[i] for i in sel]
        if "embeddings" in include:
            out["embeddings"] = self.rows[list(sel)]
        return out

    def query(self, query_embeddings, n_results=10, where=None, include=("metadatas", "distances")):
        import torch

        if where:
            raise VitkError("where= filters are not supported on the static GPU bank")
        q = torch.as_tensor(np.asarray(query_embeddings, dtype=np.float32)).reshape(-1, self.index.d)
        q = q.cuda(self.index._owner.device if hasattr(self.index._owner, "device") else None)
        idx, score, _, _ = self.index.search(q, n_results, normalise_queries=(self.space == "cosine"))
        idx, score = idx.cpu().numpy(), score.cpu().numpy()
        dist = score if self.space == "l2" else 1.0 - score
        out = {"ids": [[self.ids[j] for j in r if j >= 0] for r in idx]}
        if "distances" in include:
            out["distances"] = [[float(x) for x, j in zip(dr, r) if j >= 0] for dr, r in zip(dist, idx)]
        if "metadatas" in include:
            out["metadatas"] = [[self.metadatas[j] for j in r if j >= 0] for r in idx]
        if "embeddings" in include:
            out["embeddings"] = [self.rows[[j for j in r if j >= 0]] for r in idx]
        return out


def determine_class_pass2(counts):
    """chroma.py:102-134 on the GPU's class counts: plain majority, first max wins."""
    a = int(np.argmax(counts))
    return {"dir": LABEL_NAMES[a], "probs": temp_smax(list(counts), 7.0)}


def determine_class_pass1(counts, k=25):
    """chroma.py:36-71: accept only if the top class has >= 20 of 25 votes (0.8 k)."""
    if max(counts) >= int(round(0.8 * k)):
        return determine_class_pass2(counts)
    return {"dir": "pass2"}


class RattChunkRetriever:
    """Exact cosine top-k against a static candidate store (ratt_chunk_retriever.py:62-159
    without the per-query metadata mask, which needs the Chroma metadata store — out of scope).
    Returns (B, top_k, D) float32, zero padded."""

    def __init__(self, embedder, cand_embs, top_k=10, search_k=200):
        self.top_k, self.search_k = top_k, search_k
        self.index = BankIndex(embedder, cand_embs, None, 0, metric="cosine", normalise_rows=True)
        c = np.asarray(cand_embs, np.float32).reshape(self.index.m, -1)
        self.cand = c / (np.linalg.norm(c, axis=1, keepdims=True) + 1e-8)

    def __call__(self, chunk_embs, metadata=None):
        import torch

        q = chunk_embs.numpy() if hasattr(chunk_embs, "numpy") and not hasattr(chunk_embs, "is_cuda") \
            else chunk_embs
        qt = torch.as_tensor(np.asarray(q, np.float32)).cuda() if not hasattr(q, "is_cuda") else q
        idx, _, _, _ = self.index.search(qt, self.top_k, normalise_queries=True)
        idx = idx.cpu().numpy()
        out = np.zeros((idx.shape[0], self.top_k, self.index.d), np.float32)
        for i, r in enumerate(idx):
            r = r[r >= 0]
            out[i, : len(r)] = self.cand[r]
        return out
