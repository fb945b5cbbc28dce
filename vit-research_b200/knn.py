"""k-NN / vote entry points with the reference's call shapes, backed by libvitk (sm_100a).

Mirrors:
  np.load('*_embeddings.npz')  bank fixtures              nba_proj/clustering.py:26-36
  collection.query(query_embeddings=, n_results=)          nba_proj/chroma.py:211-214
  determine_class_pass1 / determine_class_pass2            nba_proj/chroma.py:36-71,102-134
  RattChunkRetriever(collection, top_k, search_k)(chunk_embs, metadata)
                                                           nba_proj/retrieval/ratt_chunk_retriever.py:62-159
  FrameRetriever(collection, top_k, search_k)(chunk_embs, metadata)
                                                           nba_proj/retrieval/frame_retriever.py:3-72
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
    """One bank on one GPU: top-k + vote in one call.

    Every BankIndex owns its OWN device bank on the embedder's handle (vitk_bank_new); each call
    selects it first, so several indexes -- the left/right/none label bank next to a clip store,
    as the reference keeps several Chroma collections -- live side by side without touching each
    other.  They share the handle's k-NN workspace: calls are ordered on the device."""

    def __init__(self, embedder_or_handle, rows, labels=None, n_classes=3, metric="l2",
                 normalise_rows=False, ids=None, metadatas=None):
        self.lib = _lib.load()
        self._owner = embedder_or_handle
        self.handle = getattr(embedder_or_handle, "handle", embedder_or_handle)
        self.metric = {"l2": _lib.METRIC_L2, "cosine": _lib.METRIC_COSINE}[metric]
        self.metric_name = metric
        self.n_classes = int(n_classes) if labels is not None else 0
        self.metadatas = metadatas
        self.bank_id = int(self.lib.vitk_bank_new(self.handle))
        if self.bank_id < 0:
            raise VitkError("vitk_bank_new failed")
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

    def select(self):
        """Make this index's bank the handle's current one (cheap: a pointer swap)."""
        if self.bank_id is None:
            raise VitkError("this BankIndex was closed")
        check(self.lib.vitk_bank_select(self.handle, self.bank_id), self.handle, "vitk_bank_select")

    def close(self):
        if getattr(self, "bank_id", None) is not None and getattr(self.handle, "value", self.handle):
            try:
                self.lib.vitk_bank_free(self.handle, self.bank_id)
            except Exception:  # pragma: no cover - interpreter shutdown
                pass
        self.bank_id = None

    def __del__(self):  # pragma: no cover
        try:
            if getattr(self._owner, "handle", None) is not None and self._owner.handle.value:
                self.close()
        except Exception:
            pass

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
        self.select()
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

    def search(self, q, k, normalise_queries=False, stream=None, row_mask=None, row_filter=None):
        """q: CUDA float32 (n,d).  Returns CUDA tensors (idx i32 (n,k), score f32 (n,k),
        label i32 (n,), counts i32 (n,n_classes)); 1 <= k <= 1024.
        row_mask: optional (m,) bool/uint8 (numpy or CUDA tensor), 1 = row in play (a Chroma
        where= filter evaluated by the caller).
        row_filter: optional per-query predicate evaluated inside the kernels
        (vitk_knn_vote_filtered): dict with any of
            "ne":  (col int32 CUDA (m,),   q int32 CUDA (n,))    row passes iff col[row] != q[query]
            "eq":  (col int32 CUDA (m,),   q int32 CUDA (n,))    ... col[row] == q[query]
            "rng": (col float64 CUDA (m,), lo, hi float64 CUDA (n,))   lo <= col[row] <= hi
            "rng2": (lo, hi) a second window on the same column, un-rounded
            "rng_round_f32": bool -- compare float32(col[row]) in "rng" (the reference's float32 arrays)."""
        import torch

        if not q.is_cuda or q.dtype != torch.float32:
            raise VitkError("search() wants a CUDA float32 tensor")
        k = int(k)
        if k < 1 or k > _lib.KNN_MAX_K:
            raise VitkError(f"search(): k={k} outside [1,{_lib.KNN_MAX_K}] (no silent truncation)")
        q = q.contiguous()
        n = q.shape[0]
        idx = torch.empty((n, k), dtype=torch.int32, device=q.device)
        score = torch.empty((n, k), dtype=torch.float32, device=q.device)
        label = torch.empty((n,), dtype=torch.int32, device=q.device)
        counts = torch.zeros((n, max(self.n_classes, 1)), dtype=torch.int32, device=q.device)
        self.select()
        if row_mask is None and row_filter is None:
            check(self.lib.vitk_knn_vote(self.handle, ptr(q), n, k, self.metric,
                                         int(normalise_queries), ptr(idx), ptr(score), ptr(label),
                                         ptr(counts), stream_ptr(stream)), self.handle,
                  "vitk_knn_vote")
            return idx, score, label, counts
        mask = None
        if row_mask is not None:
            mask = torch.as_tensor(np.asarray(row_mask) if not hasattr(row_mask, "is_cuda") else row_mask)
            mask = mask.to(device=q.device, dtype=torch.uint8).contiguous()
            if mask.numel() != self.m:
                raise VitkError(f"row_mask has {mask.numel()} entries for a bank of {self.m} rows")
        filt, keep = None, []
        if row_filter is not None:
            filt = _lib.VitkRowFilter()

            def dev(t, dtype, count, what):
                t = t.to(device=q.device, dtype=dtype).contiguous()
                if t.numel() != count:
                    raise VitkError(f"row_filter {what}: {t.numel()} entries, expected {count}")
                keep.append(t)
                return t.data_ptr()
            if "ne" in row_filter:
                filt.ne_col = dev(row_filter["ne"][0], torch.int32, self.m, "ne column")
                filt.q_ne = dev(row_filter["ne"][1], torch.int32, n, "ne operand")
            if "eq" in row_filter:
                filt.eq_col = dev(row_filter["eq"][0], torch.int32, self.m, "eq column")
                filt.q_eq = dev(row_filter["eq"][1], torch.int32, n, "eq operand")
            if "rng" in row_filter:
                filt.rng_col = dev(row_filter["rng"][0], torch.float64, self.m, "range column")
                filt.q_lo = dev(row_filter["rng"][1], torch.float64, n, "range low")
                filt.q_hi = dev(row_filter["rng"][2], torch.float64, n, "range high")
                if "rng2" in row_filter:
                    filt.q_lo2 = dev(row_filter["rng2"][0], torch.float64, n, "range-2 low")
                    filt.q_hi2 = dev(row_filter["rng2"][1], torch.float64, n, "range-2 high")
                filt.rng_round_f32 = int(bool(row_filter.get("rng_round_f32", False)))
        check(self.lib.vitk_knn_vote_filtered(self.handle, ptr(q), n, k, self.metric,
                                              int(normalise_queries), ptr(mask),
                                              C.byref(filt) if filt is not None else None, ptr(idx),
                                              ptr(score), ptr(label), ptr(counts), stream_ptr(stream)),
              self.handle, "vitk_knn_vote_filtered")
        torch.cuda.synchronize(q.device)  # `mask` / the per-query operands are released on return
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


def _where_columns(where, out=None):
    """Metadata keys a where-clause reads (None if it uses an operator this module lacks)."""
    out = set() if out is None else out
    for key, cond in (where or {}).items():
        if key in ("$and", "$or"):
            for w in cond:
                if _where_columns(w, out) is None:
                    return None
        elif key.startswith("$"):
            return None
        else:
            if isinstance(cond, dict) and any(op not in _WHERE_OPS for op in cond):
                return None
            out.add(key)
    return out


class BankCollection:
    """Chroma `Collection` look-alike over a GPU bank: query (with where= filters) / get / count /
    upsert / delete.

    `hnsw:space` 'l2' returns SQUARED L2 ascending, 'cosine' returns 1 - cos ascending, like
    chromadb; results are exact brute force, not HNSW-approximate.  upsert of new ids appends to
    the device bank in place (vitk_bank_write: only the new rows are prepared), upsert of an
    existing id overwrites its row; delete rebuilds the bank from the surviving rows.
    n_results up to 1024 per query (the reference asks for 100-750, frame_retriever.py:41-43,
    config.py:10-11); like chromadb, asking for more rows than the collection holds returns what
    there is; asking for more than 1024 raises (never a silently shortened list).
    A collection may start empty (rows=None): the device bank appears with the first upsert."""

    def __init__(self, embedder, rows=None, ids=(), metadatas=None, space="l2", label_key="label",
                 label_names=LABEL_NAMES, dim=None, name=None):
        if space not in ("l2", "cosine"):
            raise VitkError(f"hnsw:space {space!r} unsupported (l2 or cosine)")
        self.space = space
        self.name = name
        self.label_key = label_key
        self.label_names = tuple(label_names)
        self._embedder = embedder
        ids = list(ids)
        metadatas = list(metadatas) if metadatas is not None else [{}] * len(ids)
        self.ids, self.metadatas = ids, [dict(m) if m else {} for m in metadatas]
        self._row_of = {i: r for r, i in enumerate(self.ids)}
        if len(self._row_of) != len(self.ids):
            raise VitkError("BankCollection: duplicate ids")
        if len(self.metadatas) != len(self.ids):
            raise VitkError("BankCollection: one metadata dict per id")
        if rows is None or len(ids) == 0:
            d = int(dim if dim is not None else getattr(getattr(embedder, "geometry", None), "hidden", 768))
            self._rows = np.zeros((0, d), np.float32)
        else:
            rows = np.ascontiguousarray(rows, np.float32).reshape(len(self.ids), -1)
            self._rows = rows.copy()           # host mirror, capacity grown geometrically
        self._n = len(self.ids)
        self.index = None
        self._cols, self._dev_cols, self._codes = {}, {}, {}
        if self._n:
            self._build()

    def _label_of(self, m):
        k = self.label_key
        return self.label_names.index(m[k]) if m and k in m and m[k] in self.label_names \
            else len(self.label_names) - 1

    def _build(self):
        if self.index is not None:
            self.index.close()
        labels = np.array([self._label_of(m) for m in self.metadatas], np.int32)
        self.index = BankIndex(self._embedder, self._rows[: self._n], labels, len(self.label_names),
                               metric="l2" if self.space == "l2" else "cosine",
                               normalise_rows=(self.space == "cosine"), ids=self.ids,
                               metadatas=self.metadatas)
        self._cols, self._dev_cols = {}, {}

    @property
    def rows(self):
        return self._rows[: self._n]

    @property
    def device(self):
        return getattr(self._embedder, "device", None)

    def count(self):
        return self._n

    # ---- metadata as columns (vectorised where-clauses, device columns for row filters) ----
    def column(self, key):
        """(values object array (n,), present bool array (n,)) of one metadata key."""
        c = self._cols.get(key)
        if c is None or len(c[0]) != self._n:
            vals = np.empty(self._n, object)
            present = np.zeros(self._n, bool)
            for i, m in enumerate(self.metadatas):
                if key in m:
                    vals[i], present[i] = m[key], True
            pv = [v for v, p in zip(vals, present) if p]
            num = None  # numeric columns compare vectorised; mixed / string ones stay object arrays
            if pv and all(isinstance(v, (int, float, np.integer, np.floating)) and not isinstance(v, bool)
                          for v in pv):
                num = np.array([float(v) if p else 0.0 for v, p in zip(vals, present)], dtype=np.float64)
            c = (vals, present, num)
            self._cols[key] = c
        return c

    def where_mask(self, where):
        """uint8 (n,) mask of the rows a Chroma where-clause keeps (vectorised per column)."""
        if not where:
            return None
        if _where_columns(where) is None:
            return np.fromiter((match_where(m, where) for m in self.metadatas), np.uint8, self._n)

        def ev(w):
            out = np.ones(self._n, bool)
            for key, cond in w.items():
                if key == "$and":
                    for x in cond:
                        out &= ev(x)
                elif key == "$or":
                    acc = np.zeros(self._n, bool)
                    for x in cond:
                        acc |= ev(x)
                    out &= acc
                else:
                    vals, present, num = self.column(key)
                    conds = cond.items() if isinstance(cond, dict) else [("$eq", cond)]
                    for op, val in conds:
                        numeric = num is not None and isinstance(val, (int, float, np.integer, np.floating)) \
                            and not isinstance(val, bool)
                        if numeric and op in ("$eq", "$ne", "$gt", "$gte", "$lt", "$lte"):
                            r = {"$eq": num == val, "$ne": num != val, "$gt": num > val, "$gte": num >= val,
                                 "$lt": num < val, "$lte": num <= val}[op]
                        else:
                            f = _WHERE_OPS[op]
                            r = np.fromiter((bool(p and f(v, val)) for v, p in zip(vals, present)), bool, self._n)
                        out &= r & present          # a missing key never matches (see match_where)
            return out
        return ev(where).astype(np.uint8)

    def device_column(self, key, kind):
        """Metadata column as a CUDA tensor for BankIndex.search(row_filter=...): kind "int"
        (int32; strings are coded through a per-key vocabulary, missing -> -2**31) or "float"
        (float64, missing -> NaN, which fails every range test)."""
        import torch
        ck = (key, kind)
        t = self._dev_cols.get(ck)
        if t is not None and t.numel() == self._n:
            return t
        vals, present, num = self.column(key)
        if kind == "float":
            if num is None:
                raise VitkError(f"metadata key {key!r} is not numeric")
            a = np.where(present, num, np.nan)
            t = torch.from_numpy(a).to(device=f"cuda:{self.device or 0}")
        else:
            a = np.full(self._n, -2 ** 31, np.int64)
            for i in range(self._n):
                if present[i]:
                    a[i] = self.code(key, vals[i])
            t = torch.from_numpy(a.astype(np.int32)).to(device=f"cuda:{self.device or 0}")
        self._dev_cols[ck] = t
        return t

    def code(self, key, value):
        """int32 code of a metadata value: integers are themselves, anything else (the 'side'
        strings) gets a per-key vocabulary index >= 2**30."""
        if isinstance(value, (int, np.integer)) and not isinstance(value, bool) and -2 ** 30 < int(value) < 2 ** 30:
            return int(value)
        if isinstance(value, (float, np.floating)) and float(value).is_integer() and abs(value) < 2 ** 30:
            return int(value)
        vocab = self._codes.setdefault(key, {})
        if value not in vocab:
            vocab[value] = 2 ** 30 + len(vocab)
        return vocab[value]

    def get(self, ids=None, where=None, include=("metadatas",)):
        sel = np.arange(self._n) if ids is None else \
            np.array([self._row_of[i] for i in ids if i in self._row_of], np.int64)
        if where and self._n:
            mask = self.where_mask(where)
            sel = sel[mask[sel].astype(bool)]
        sel = [int(i) for i in sel]
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
        ids = list(ids)
        emb = np.ascontiguousarray(np.asarray(embeddings, np.float32)).reshape(len(ids), -1)
        if emb.shape[1] != self._rows.shape[1]:
            if self._n == 0:
                self._rows = np.zeros((0, emb.shape[1]), np.float32)
            else:
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
            for p in new:
                self._row_of[ids[p]] = len(self.ids)
                self.ids.append(ids[p])
                self.metadatas.append(dict(metadatas[p]))
            first = self.index is None
            self._n = need
            if first:
                self._build()
            else:
                self.index.append(block, labels)
        self._cols, self._dev_cols = {}, {}

    add = upsert     # collection.add(...) of fresh ids (write_per_vid_embeddings_chroma.py:248-265)

    def delete(self, ids=None, where=None):
        """collection.delete(where={...}) (db_rebuild.py:121): drops the matching rows."""
        if self._n == 0:
            return
        drop = np.ones(self._n, bool) if ids is None else np.zeros(self._n, bool)
        if ids is not None:
            for i in ids:
                if i in self._row_of:
                    drop[self._row_of[i]] = True
        if where:
            drop &= self.where_mask(where).astype(bool)
        if not drop.any():
            return
        keep = np.nonzero(~drop)[0]
        self._rows = np.ascontiguousarray(self._rows[: self._n][keep])
        self.ids = [self.ids[i] for i in keep]
        self.metadatas = [self.metadatas[i] for i in keep]
        self._row_of = {i: r for r, i in enumerate(self.ids)}
        self._n = len(keep)
        if self._n:
            self._build()
        else:
            if self.index is not None:
                self.index.close()
            self.index = None
            self._cols, self._dev_cols = {}, {}

    def search_device(self, q, n_results, where=None, row_filter=None):
        """Device-level query: q CUDA float32 (n,d) -> (idx, dist) CUDA tensors, Chroma order
        (ascending distance; idx -1 pads queries with fewer rows in play)."""
        if self.index is None:
            raise VitkError("query on an empty collection")
        n_results = int(n_results)
        if n_results < 1:
            raise VitkError("n_results must be >= 1")
        k = min(n_results, self._n)          # chromadb: "n_results > elements in index, updating"
        if k > _lib.KNN_MAX_K:
            raise VitkError(f"n_results={n_results} exceeds the {_lib.KNN_MAX_K}-neighbour limit of one query")
        mask = self.where_mask(where) if where else None
        idx, score, _, _ = self.index.search(q, k, normalise_queries=(self.space == "cosine"),
                                             row_mask=mask, row_filter=row_filter)
        return idx, (score if self.space == "l2" else 1.0 - score)

    def query(self, query_embeddings, n_results=10, where=None, include=("metadatas", "distances")):
        import torch

        q = np.asarray(query_embeddings, dtype=np.float32)
        q = q.reshape(-1, self._rows.shape[1])
        if self.index is None:
            empty = {"ids": [[] for _ in q]}
            for key in ("distances", "metadatas", "embeddings"):
                if key in include:
                    empty[key] = [[] for _ in q]
            return empty
        qt = torch.from_numpy(np.ascontiguousarray(q)).cuda(self.device)
        idx, dist = self.search_device(qt, n_results, where)
        idx, dist = idx.cpu().numpy(), dist.cpu().numpy()
        out = {"ids": [[self.ids[j] for j in r if j >= 0] for r in idx]}
        if "distances" in include:
            out["distances"] = [[float(x) for x, j in zip(dr, r) if j >= 0] for dr, r in zip(dist, idx)]
        if "metadatas" in include:
            out["metadatas"] = [[self.metadatas[j] for j in r if j >= 0] for r in idx]
        if "embeddings" in include:
            out["embeddings"] = [self.rows[[j for j in r if j >= 0]] for r in idx]
        return out


class BankClient:
    """`chromadb.PersistentClient(path=...)` look-alike holding GPU collections on one embedder:
    get_or_create_collection(name, metadata={"hnsw:space": "cosine"}) as the reference scripts
    call it (chroma.py:143-146, ratt_write_clips_to_ragdb.py:302-310).  Nothing is persisted."""

    def __init__(self, embedder, path=None):
        self._embedder, self.path, self._collections = embedder, path, {}

    def get_or_create_collection(self, name, metadata=None, **_):
        if name not in self._collections:
            space = (metadata or {}).get("hnsw:space", "l2")
            self._collections[name] = BankCollection(self._embedder, space=space, name=name)
        return self._collections[name]

    create_collection = get_or_create_collection

    def get_collection(self, name, **_):
        if name not in self._collections:
            raise ValueError(f"Collection {name} does not exist.")
        return self._collections[name]

    def delete_collection(self, name):
        c = self._collections.pop(name, None)
        if c is not None and c.index is not None:
            c.index.close()

    def list_collections(self):
        return list(self._collections.values())


def _label_counts(metadatas):
    """[left, right, none] counts of the neighbours' stored labels; anything that is neither
    'left' nor 'right' counts as none (chroma.py:46-52, 112-118)."""
    nums = [0, 0, 0]
    for m in metadatas:
        lab = m["label"]
        nums[0 if lab == "left" else 1 if lab == "right" else 2] += 1
    return nums


def vote_from_counts(counts, threshold=None):
    """The decision of determine_class_pass1 / pass2 on class counts the GPU vote already produced
    (vitk_knn_vote's counts_out): first max wins; with `threshold`, {'dir': 'pass2'} unless the top
    class reaches it."""
    nums = [int(c) for c in counts]
    if threshold is not None and max(nums) < threshold:
        return {"dir": "pass2"}
    return {"dir": LABEL_NAMES[int(np.array(nums).argmax())], "probs": temp_smax(nums, 7.0)}


def determine_class_pass1(ids, metadatas, distances):
    """chroma.py:36-71, same call: ids / metadatas / distances of ONE query as returned by
    collection.query(...)['ids'][0] etc.  Accepts only if the top class has >= 20 votes (the
    reference's literal threshold, written for top_n_closest = 25, chroma.py:27,62), else
    {'dir': 'pass2'}.  Without the prints."""
    nums = _label_counts(metadatas[: len(ids)])
    return vote_from_counts(nums, threshold=20)


def determine_class_pass2(ids, metadatas, distances):
    """chroma.py:102-134, same call: plain majority of the neighbours' labels, first max wins
    (np.argmax: left > right > none on ties), probs = temp_smax(counts, 7.0)."""
    return vote_from_counts(_label_counts(metadatas[: len(ids)]))


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


def _meta_value(x):
    """One entry of the retrievers' `metadata` dict: a tf tensor (`.numpy()`), numpy / torch scalar
    or plain Python value; bytes are decoded like `metadata["side"][i].numpy().decode()`."""
    if hasattr(x, "numpy"):
        x = x.numpy()
    if isinstance(x, np.ndarray) and x.ndim == 0:
        x = x.item()
    if isinstance(x, (bytes, np.bytes_)):
        x = x.decode()
    return x


class RattChunkRetriever:
    """Drop-in for retrieval/ratt_chunk_retriever.py:62-159: same constructor
    `(collection, top_k=10, search_k=200)`, same call `(chunk_embs, metadata)` with
    metadata = {"vid", "side", "t_center", "t_width"} (tf tensors or arrays), same output
    `(B, top_k, D)` float32 of the L2-normalised candidate vectors, best first, zero padded.

    Reference, per side: collection.get(side == s, t_center within the side's widest window) pulls
    every candidate to the host, sims = Q @ C.T, and per query a mask
    `(cand_vid != vid) & (q_start <= t_center <= q_end)` (float32 compares) picks the rows before
    argpartition / argsort.  Here the whole batch is ONE filtered search on the device bank: the
    predicate (other video, same side, float32 window, plus the side-wide window on the stored
    double) is evaluated inside the k-NN kernels from three metadata columns, so no candidate
    matrix, no (queries x candidates) mask and no host GEMM exist.  Like the reference, float32
    query arrays are normalised IN PLACE.  `search_k` is stored and, as in the reference, unused.
    `collection` is a BankCollection in 'cosine' space with 'vid_num', 'side', 't_center' metadata."""

    def __init__(self, collection, top_k=10, search_k=200):
        self.collection = collection
        self.top_k = top_k
        self.search_k = search_k

    def __call__(self, chunk_embs, metadata):
        import torch

        q_np = chunk_embs.numpy() if hasattr(chunk_embs, "numpy") and not hasattr(chunk_embs, "is_cuda") \
            else chunk_embs
        if hasattr(q_np, "is_cuda"):
            q_np = q_np.detach().cpu().numpy()
        q_np = np.asarray(q_np, dtype=np.float32)
        q_np /= (np.linalg.norm(q_np, axis=1, keepdims=True) + 1e-8)     # ratt_chunk_retriever.py:74-75
        B, D = q_np.shape
        out = np.zeros((B, self.top_k, D), dtype=np.float32)
        col = self.collection
        if B == 0 or col.count() == 0:
            return out
        vids = np.array([int(_meta_value(metadata["vid"][i])) for i in range(B)], dtype=np.int32)
        sides = [_meta_value(metadata["side"][i]) for i in range(B)]
        t_centers = np.array([float(_meta_value(metadata["t_center"][i])) for i in range(B)], dtype=np.float32)
        t_widths = np.array([float(_meta_value(metadata["t_width"][i])) for i in range(B)], dtype=np.float32)
        q_starts = t_centers - t_widths / 2.0          # float32, like the reference (:91-92,126-127)
        q_ends = t_centers + t_widths / 2.0
        # the side-wide window of collection.get (:94-105): float(np.min / np.max) over the side's queries
        g_lo, g_hi = np.zeros(B, np.float64), np.zeros(B, np.float64)
        for s_ in set(sides):
            sel = [i for i in range(B) if sides[i] == s_]
            g_lo[sel] = float(np.min(q_starts[sel]))
            g_hi[sel] = float(np.max(q_ends[sel]))
        dev = f"cuda:{col.device or 0}"
        filt = {
            "ne": (col.device_column("vid_num", "int"), torch.from_numpy(vids).to(dev)),
            "eq": (col.device_column("side", "int"),
                   torch.tensor([col.code("side", s_) for s_ in sides], dtype=torch.int32, device=dev)),
            "rng": (col.device_column("t_center", "float"),
                    torch.from_numpy(q_starts.astype(np.float64)).to(dev),
                    torch.from_numpy(q_ends.astype(np.float64)).to(dev)),
            "rng2": (torch.from_numpy(g_lo).to(dev), torch.from_numpy(g_hi).to(dev)),
            "rng_round_f32": True,
        }
        k = min(int(self.top_k), col.count())
        idx, _ = col.search_device(torch.from_numpy(q_np).to(dev), k, row_filter=filt)
        idx = idx.cpu().numpy()
        rows = col.rows
        for i, r in enumerate(idx):
            r = r[r >= 0]
            if len(r):
                c = rows[r].astype(np.float32)
                out[i, : len(r)] = c / (np.linalg.norm(c, axis=1, keepdims=True) + 1e-8)   # :115-116
        return out


class FrameRetriever:
    """Drop-in for retrieval/frame_retriever.py:3-72: per query
    `collection.query(n_results=search_k, where={$and: [vid_num $ne vid, side, t_norm $gte/$lte]},
    include=["embeddings"])`, keep the first top_k stored embeddings, zero-pad, normalise.  The
    whole batch is one filtered device search; exact search makes the search_k over-fetch
    unnecessary (the first top_k of an exact top-search_k ARE the exact top_k)."""

    def __init__(self, collection, top_k=10, search_k=200):
        self.collection = collection
        self.top_k = top_k
        self.search_k = search_k

    def __call__(self, chunk_embs, metadata):
        import torch

        q_np = chunk_embs.numpy() if hasattr(chunk_embs, "numpy") and not hasattr(chunk_embs, "is_cuda") \
            else chunk_embs
        if hasattr(q_np, "is_cuda"):
            q_np = q_np.detach().cpu().numpy()
        q_np = np.ascontiguousarray(np.asarray(q_np), dtype=np.float32)
        B, D = q_np.shape
        out = np.zeros((B, self.top_k, D), dtype=np.float32)
        col = self.collection
        if B == 0 or col.count() == 0:
            return out
        vids = np.array([int(_meta_value(metadata["vid"][i])) for i in range(B)], dtype=np.int32)
        sides = [_meta_value(metadata["side"][i]) for i in range(B)]
        t_c = [float(_meta_value(metadata["t_center"][i])) for i in range(B)]
        t_w = [float(_meta_value(metadata["t_width"][i])) for i in range(B)]
        t_min = np.array([c - (w / 2) for c, w in zip(t_c, t_w)], np.float64)     # python floats (:35-36)
        t_max = np.array([c + (w / 2) for c, w in zip(t_c, t_w)], np.float64)
        dev = f"cuda:{col.device or 0}"
        filt = {
            "ne": (col.device_column("vid_num", "int"), torch.from_numpy(vids).to(dev)),
            "eq": (col.device_column("side", "int"),
                   torch.tensor([col.code("side", s_) for s_ in sides], dtype=torch.int32, device=dev)),
            "rng": (col.device_column("t_norm", "float"), torch.from_numpy(t_min).to(dev),
                    torch.from_numpy(t_max).to(dev)),
        }
        k = min(int(self.top_k), int(self.search_k), col.count())
        idx, _ = col.search_device(torch.from_numpy(q_np).to(dev), k, row_filter=filt)
        idx = idx.cpu().numpy()
        rows = col.rows
        for i, r in enumerate(idx):
            r = r[r >= 0]
            if len(r):
                out[i, : len(r)] = rows[r]
        out /= (np.linalg.norm(out, axis=2, keepdims=True) + 1e-8)      # :68-69 (zero pads stay zero)
        return out
