"""k-NN / vote entry points with the reference's call shapes, backed by libvitk (sm_100a).

Mirrors:
  np.load('*_embeddings.npz')  bank fixtures              nba_proj/clustering.py:26-36
  collection.query(query_embeddings=, n_results=)          nba_proj/chroma.py:211-214
  determine_class_pass1 / determine_class_pass2            nba_proj/chroma.py:36-71,102-134
  RattChunkRetriever(...)(chunk_embs) exact cosine top-k   nba_proj/retrieval/ratt_chunk_retriever.py:62-159
  collection.upsert / delete / where= filters              nba_proj/generate_clips_hmm.py:456-470,
                                                           retrieval/frame_retriever.py:41-53
  determine_class (mean stored probs, confident upserts)   nba_proj/generate_clips_hmm.py:179-310
The hot path of north_star uses static banks; the mutable bank (confident upserts in frame
order) is `classify_stream_with_upserts`.
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
        self.normalise_rows = bool(normalise_rows)

    def write(self, at_row, rows, labels=None):
        """Write rows at [at_row, at_row+len(rows)): at_row == self.m appends, smaller overwrites
        (vitk_bank_write).  rows: numpy (host) or CUDA float32 tensor."""
        if hasattr(rows, "is_cuda"):
            import torch
            if not rows.is_cuda or rows.dtype != torch.float32:
                raise VitkError("a tensor of bank rows must be CUDA float32")
            rows = rows.reshape(-1, self.d).contiguous()
            lab = None if labels is None else torch.as_tensor(labels, dtype=torch.int32,
                                                              device=rows.device).contiguous()
            on_device, n_new = 1, rows.shape[0]
            lab_host = None if lab is None else lab.cpu().numpy()
        else:
            rows = np.ascontiguousarray(rows, dtype=np.float32).reshape(-1, self.d)
            lab = None if labels is None else np.ascontiguousarray(labels, dtype=np.int32)
            on_device, n_new, lab_host = 0, rows.shape[0], lab
        if (lab is None) != (self.n_classes == 0):
            raise VitkError("vitk_bank_write failed: labels must be given exactly when the bank has labels")
        if lab_host is not None and len(lab_host) != n_new:
            raise VitkError("vitk_bank_write failed: one label per row")
        if lab_host is not None and n_new and (lab_host.min() < 0 or lab_host.max() >= self.n_classes):
            raise VitkError("vitk_bank_write failed: label outside [0,n_classes)")
        check(self.lib.vitk_bank_write(self.handle, int(at_row), ptr(rows), n_new, ptr(lab), on_device,
                                       int(self.normalise_rows), stream_ptr(None)), self.handle,
              "vitk_bank_write")
        self.m = int(self.lib.vitk_bank_rows(self.handle))
        if self.labels is not None and lab_host is not None:
            grown = np.zeros(self.m, np.int32)
            grown[: len(self.labels)] = self.labels
            grown[at_row: at_row + n_new] = lab_host
            self.labels = grown

    def append(self, rows, labels=None):
        self.write(self.m, rows, labels)

    def search(self, q, k, normalise_queries=False, stream=None, row_mask=None):
        """q: CUDA float32 (n,d).  Returns CUDA tensors (idx i32 (n,k), score f32 (n,k),
        label i32 (n,), counts i32 (n,n_classes)).  row_mask: optional (m,) bool/uint8 (numpy or
        CUDA tensor), 1 = row in play (a Chroma where= filter evaluated by the caller)."""
        import torch

        if not q.is_cuda or q.dtype != torch.float32:
            raise VitkError("search() wants a CUDA float32 tensor")
        q = q.contiguous()
        n = q.shape[0]
        idx = torch.empty((n, k), dtype=torch.int32, device=q.device)
        score = torch.empty((n, k), dtype=torch.float32, device=q.device)
        label = torch.empty((n,), dtype=torch.int32, device=q.device)
        counts = torch.zeros((n, max(self.n_classes, 1)), dtype=torch.int32, device=q.device)
        if row_mask is not None:
            mask = torch.as_tensor(np.asarray(row_mask) if not hasattr(row_mask, "is_cuda") else row_mask)
            mask = mask.to(device=q.device, dtype=torch.uint8).contiguous()
            if mask.numel() != self.m:
                raise VitkError(f"row_mask has {mask.numel()} entries for a bank of {self.m} rows")
            check(self.lib.vitk_knn_vote_masked(self.handle, ptr(q), n, int(k), self.metric,
                                                int(normalise_queries), ptr(mask), ptr(idx), ptr(score),
                                                ptr(label), ptr(counts), stream_ptr(stream)), self.handle,
                  "vitk_knn_vote_masked")
            torch.cuda.synchronize(q.device)  # `mask` is released on return
            return idx, score, label, counts
        check(self.lib.vitk_knn_vote(self.handle, ptr(q), n, int(k), self.metric,
                                     int(normalise_queries), ptr(idx), ptr(score), ptr(label),
                                     ptr(counts), stream_ptr(stream)), self.handle,
              "vitk_knn_vote")
        return idx, score, label, counts


_WHERE_OPS = {
    "$eq": lambda a, b: a == b,
    "$ne": lambda a, b: a != b,
    "$gt": lambda a, b: a is not None and a > b,
    "$gte": lambda a, b: a is not None and a >= b,
    "$lt": lambda a, b: a is not None and a < b,
    "$lte": lambda a, b: a is not None and a <= b,
    "$in": lambda a, b: a in b,
    "$nin": lambda a, b: a not in b,
}


def match_where(meta, where):
    """Chroma where-clause on one metadata dict: {"key": value}, {"key": {"$op": value}},
    {"$and": [...]}, {"$or": [...]} (frame_retriever.py:44-51, ratt_chunk_retriever.py:99-105,
    db_rebuild.py:121).  A missing key never matches, $ne / $nin included (chromadb releases differ
    on that point; the reference pins none)."""
    if not where:
        return True
    meta = meta or {}
    for key, cond in where.items():
        if key == "$and":
            if not all(match_where(meta, w) for w in cond):
                return False
        elif key == "$or":
            if not any(match_where(meta, w) for w in cond):
                return False
        elif key.startswith("$"):
            raise VitkError(f"unsupported where operator {key}")
        else:
            if key not in meta:
                return False
            if isinstance(cond, dict):
                for op, val in cond.items():
                    if op not in _WHERE_OPS:
                        raise VitkError(f"unsupported where operator {op}")
                    if not _WHERE_OPS[op](meta[key], val):
                        return False
            elif meta[key] != cond:
                return False
    return True


class BankCollection:
    """Chroma `Collection` look-alike over a GPU bank: query (with where= filters) / get / count /
    upsert / delete.

    `hnsw:space` 'l2' returns SQUARED L2 ascending, 'cosine' returns 1 - cos ascending, like
    chromadb; results are exact brute force, not HNSW-approximate.  upsert of new ids appends to
    the device bank in place (vitk_bank_write: only the new rows are prepared), upsert of an
    existing id overwrites its row; delete rebuilds the bank from the surviving rows."""

    def __init__(self, embedder, rows, ids, metadatas, space="l2", label_key="label",
                 label_names=LABEL_NAMES):
        self.space = space
        self.label_key = label_key
        self.label_names = tuple(label_names)
        self._embedder = embedder
        self.ids, self.metadatas = list(ids), [dict(m) if m else {} for m in metadatas]
        self._row_of = {i: r for r, i in enumerate(self.ids)}
        if len(self._row_of) != len(self.ids):
            raise VitkError("BankCollection: duplicate ids")
        rows = np.ascontiguousarray(rows, np.float32).reshape(len(self.ids), -1)
        self._rows = rows.copy()           # host mirror, capacity grown geometrically
        self._n = len(self.ids)
        self._build()

    def _label_of(self, m):
        k = self.label_key
        return self.label_names.index(m[k]) if m and k in m and m[k] in self.label_names \
            else len(self.label_names) - 1

    def _build(self):
        labels = np.array([self._label_of(m) for m in self.metadatas], np.int32)
        self.index = BankIndex(self._embedder, self._rows[: self._n], labels, len(self.label_names),
                               metric="l2" if self.space == "l2" else "cosine",
                               normalise_rows=(self.space == "cosine"), ids=self.ids,
                               metadatas=self.metadatas)

    @property
    def rows(self):
        return self._rows[: self._n]

    def count(self):
        return self._n

    def get(self, ids=None, where=None, include=("metadatas",)):
        sel = range(self._n) if ids is None else [self._row_of[i] for i in ids if i in self._row_of]
        if where:
            sel = [i for i in sel if match_where(self.metadatas[i], where)]
        sel = list(sel)
        out = {"ids": [self.ids[i] for i in sel]}
        if "metadatas" in include:
            out["metadatas"] = [self.metadatas[i] for i in sel]
        if "embeddings" in include:
            out["embeddings"] = self.rows[sel]
        return out

    def upsert(self, embeddings, ids, metadatas=None):
        """collection.upsert(embeddings=, ids=, metadatas=) (generate_clips_hmm.py:464-468)."""
        if isinstance(ids, str):
            ids = [ids]
        emb = np.ascontiguousarray(np.asarray(embeddings, np.float32)).reshape(len(ids), -1)
        if emb.shape[1] != self._rows.shape[1]:
            raise VitkError(f"upsert: embedding width {emb.shape[1]} != {self._rows.shape[1]}")
        if metadatas is None:
            metadatas = [{}] * len(ids)
        elif isinstance(metadatas, dict):
            metadatas = [metadatas]
        if len(metadatas) != len(ids):
            raise VitkError("upsert: one metadata dict per id")
        last = {i: p for p, i in enumerate(ids)}          # a repeated id: the last one wins
        new = [p for p, i in enumerate(ids) if i not in self._row_of and last[i] == p]
        for p, i in enumerate(ids):                        # existing ids: overwrite in place
            if i in self._row_of and last[i] == p:
                r = self._row_of[i]
                self._rows[r] = emb[p]
                self.metadatas[r] = dict(metadatas[p])
                self.index.write(r, emb[p: p + 1], np.array([self._label_of(metadatas[p])], np.int32))
        if new:
            need = self._n + len(new)
            if need > self._rows.shape[0]:
                grown = np.empty((max(need, 2 * self._rows.shape[0]), self._rows.shape[1]), np.float32)
                grown[: self._n] = self._rows[: self._n]
                self._rows = grown
            block = emb[new]
            self._rows[self._n: need] = block
            labels = np.array([self._label_of(metadatas[p]) for p in new], np.int32)
            self.index.append(block, labels)
            for p in new:
                self._row_of[ids[p]] = len(self.ids)
                self.ids.append(ids[p])
                self.metadatas.append(dict(metadatas[p]))
            self._n = need

    def delete(self, ids=None, where=None):
        """collection.delete(where={...}) (db_rebuild.py:121): drops the matching rows."""
        drop = set(range(self._n)) if ids is None else {self._row_of[i] for i in ids if i in self._row_of}
        if where:
            drop = {i for i in drop if match_where(self.metadatas[i], where)}
        if not drop:
            return
        keep = [i for i in range(self._n) if i not in drop]
        if not keep:
            raise VitkError("delete would empty the collection: a GPU bank needs at least one row")
        self._rows = np.ascontiguousarray(self._rows[keep])
        self.ids = [self.ids[i] for i in keep]
        self.metadatas = [self.metadatas[i] for i in keep]
        self._row_of = {i: r for r, i in enumerate(self.ids)}
        self._n = len(keep)
        self._build()

    def search_device(self, q, n_results, where=None):
        """Device-level query: q CUDA float32 (n,d) -> (idx, dist) CUDA tensors, Chroma order."""
        mask = None
        if where:
            mask = np.fromiter((match_where(m, where) for m in self.metadatas), np.uint8, self._n)
        k = min(int(n_results), 64)
        idx, score, _, _ = self.index.search(q, k, normalise_queries=(self.space == "cosine"),
                                             row_mask=mask)
        return idx, (score if self.space == "l2" else 1.0 - score)

    def query(self, query_embeddings, n_results=10, where=None, include=("metadatas", "distances")):
        import torch

        q = torch.as_tensor(np.asarray(query_embeddings, dtype=np.float32)).reshape(-1, self.index.d)
        q = q.cuda(self.index._owner.device if hasattr(self.index._owner, "device") else None)
        idx, dist = self.search_device(q, n_results, where)
        idx, dist = idx.cpu().numpy(), dist.cpu().numpy()
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


def one_hot_metadata(label, video=None):
    """Metadata of a hand-labelled bank row (write_per_vid_embeddings_chroma.py:248-265): the
    stored class probabilities are one-hot."""
    m = {"label": label, "left_prob": 0.0, "right_prob": 0.0, "none_prob": 0.0}
    m[f"{label}_prob"] = 1.0
    if video is not None:
        m["video"] = video
    return m


def determine_class(ids, metadatas, distances, add_first=False, *, top_n_closest, cur_vid=None,
                    hmm_matrix=None, confidence_threshold=0.7):
    """generate_clips_hmm.py:179-310 for one frame's neighbours (without the prints).

    counts by stored label -> raw fractions; mean of the STORED class probabilities
    (lmean, rmean, nmean) -> the HMM column; class = first argmax of (mean + fraction) / 2.
    Returns [dir, prob] or, when the winning mean >= 0.7, [dir, prob, metadata-to-upsert]
    (0.999998 / 1e-6 / 1e-6 when all top_n_closest neighbours agree).  The reference's module
    globals top_n_closest / cur_vid / hmm_matrix are keyword arguments here."""
    n = len(ids)
    num = [0, 0, 0]
    for m in metadatas:
        lab = m["label"]
        num[0 if lab == "left" else 1 if lab == "right" else 2] += 1
    raw = [c / n for c in num]
    means = [np.mean(np.array([m[f"{side}_prob"] for m in metadatas])) for side in LABEL_NAMES]
    lmean, rmean, nmean = means
    col = {"left": lmean, "right": rmean, "none": nmean}
    if hmm_matrix is not None:
        if add_first:
            hmm_matrix.add_first(col)
        else:
            hmm_matrix.add_col_to_lattice(col)
    a = int(np.array([(means[i] + raw[i]) / 2 for i in range(3)]).argmax())
    side = LABEL_NAMES[a]
    if means[a] >= confidence_threshold:
        if num[a] == top_n_closest:
            md = {"label": side, "video": cur_vid, "left_prob": 0.000001, "right_prob": 0.000001,
                  "none_prob": 0.000001}
            md[f"{side}_prob"] = 0.999998
            return [side, 0.999998, md]
        return [side, means[a], {"label": side, "video": cur_vid, "left_prob": lmean,
                                 "right_prob": rmean, "none_prob": nmean}]
    return [side, means[a]]


def classify_stream_with_upserts(collection, embeddings, frame_names, top_n_closest=50, upsert_batch=16,
                                 cur_vid=None, hmm_matrix=None, block=256):
    """The mutable-bank loop of generate_clips_hmm.py:384-470 in frame order, batched on the GPU.

    Reference: for every frame, query the collection (top_n_closest neighbours), determine_class,
    remember the frame when it is confident, and every `upsert_batch` confident frames upsert
    them -- so each later frame sees a larger bank.  Here a block of frames is queried at once;
    the walk over the block stops at the frame whose confidence completes an upsert batch, the
    rows are appended on the device (vitk_bank_write) and only the REST of the block is queried
    again, so every frame is classified against exactly the bank the sequential loop shows it.
    embeddings: (N,d) float32 numpy or CUDA tensor.  Returns {"sides": [[dir, prob(, meta)]...],
    "upserts": number of upsert calls, "left_over": confident frames still waiting for a batch}."""
    import torch

    dev = getattr(collection.index._owner, "device", None)
    emb = embeddings if hasattr(embeddings, "is_cuda") else torch.as_tensor(
        np.ascontiguousarray(embeddings, dtype=np.float32))
    emb = emb.reshape(len(frame_names), -1).to(device=dev if dev is not None else "cuda",
                                               dtype=torch.float32)
    emb_host = emb.cpu().numpy()
    sides, pend_rows, pend_ids, pend_meta, upserts = [], [], [], [], 0
    n, pos = len(frame_names), 0
    while pos < n:
        end = min(pos + block, n)
        idx, dist = collection.search_device(emb[pos:end], top_n_closest)
        idx, dist = idx.cpu().numpy(), dist.cpu().numpy()
        i = pos
        while i < end:
            r = idx[i - pos]
            keep = r >= 0
            ids = [collection.ids[j] for j in r[keep]]
            metas = [collection.metadatas[j] for j in r[keep]]
            side = determine_class(ids, metas, dist[i - pos][keep], add_first=(i == 0),
                                   top_n_closest=top_n_closest, cur_vid=cur_vid, hmm_matrix=hmm_matrix)
            sides.append(side)
            i += 1
            if len(side) == 3:
                pend_rows.append(emb_host[i - 1])
                pend_ids.append(frame_names[i - 1])
                pend_meta.append(side[2])
                if len(pend_rows) == upsert_batch:
                    collection.upsert(np.stack(pend_rows), pend_ids, pend_meta)
                    pend_rows, pend_ids, pend_meta = [], [], []
                    upserts += 1
                    break  # the rest of the block must see the new rows
        pos = i
    return {"sides": sides, "upserts": upserts, "left_over": len(pend_rows)}


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
