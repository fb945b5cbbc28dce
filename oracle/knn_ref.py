"""ORACLE (test infrastructure, never shipped, never on the product path).

numpy restatement of the reference's nearest-neighbour + vote half:
  exact cosine top-k   nba_proj/retrieval/ratt_chunk_retriever.py:74-75,115-116,123,144-148
  Chroma 'l2' query    nba_proj/chroma.py:145,211-214  (hnsw:space="l2": squared L2, ascending).
                       chromadb/hnswlib are absent from /root/reference and this image (unpinned by
                       the reference); HNSW's approximation is NOT reproduced — exact brute force
                       is the oracle.
  vote                 nba_proj/chroma.py:36-71 (pass 1), :102-134 (pass 2), :28-34 (temp softmax)
  HMM-variant score    nba_proj/generate_clips_hmm.py:193-245
Pinned against the reference's only fixtures, the three *_embeddings.npz banks
(tests/golden/banks, sha256 in tests/golden/MANIFEST.json): leave-one-out own-label vote rates
99.66 % / 97.01 % / 93.68 % at k = 5 / 25 / 50.
Only tests/, __graft_entry__.smoke() and bench.py's CPU baseline / reference arm may import this.
"""
from __future__ import annotations

import numpy as np

LABELS = ("left", "right", "none")


def normalise(x: np.ndarray) -> np.ndarray:
    """ratt_chunk_retriever.py:75,116 — float32 in place division by (norm + 1e-8)."""
    x = np.asarray(x, dtype=np.float32).copy()
    x /= (np.linalg.norm(x, axis=1, keepdims=True) + 1e-8)
    return x


def _stable_topk(key: np.ndarray, k: int) -> np.ndarray:
    """Indices of the k largest keys per row, ordered (key desc, index asc) — the deterministic
    completion of argpartition+argsort (ratt_chunk_retriever.py:144-148), whose order among exact
    ties is unspecified."""
    idx = np.argsort(-key, axis=1, kind="stable")
    return idx[:, :k]


def cosine_topk(q: np.ndarray, bank: np.ndarray, k: int, exclude_self: bool = False,
                dtype=np.float64):
    """sims = normalise(q) @ normalise(bank).T (ratt_chunk_retriever.py:123), top-k descending.
    Returns (idx (Q,k) int64, sims (Q,k) float64, full sims)."""
    qn, bn = normalise(q).astype(dtype), normalise(bank).astype(dtype)
    sims = qn @ bn.T
    if exclude_self:
        np.fill_diagonal(sims, -np.inf)
    idx = _stable_topk(sims, k)
    return idx, np.take_along_axis(sims, idx, 1), sims


def l2_topk(q: np.ndarray, bank: np.ndarray, k: int, exclude_self: bool = False, dtype=np.float64):
    """Squared-L2 brute force, ascending (what Chroma's 'l2' space returns as `distances`)."""
    q, bank = np.asarray(q, dtype), np.asarray(bank, dtype)
    d2 = (q * q).sum(1)[:, None] + (bank * bank).sum(1)[None, :] - 2.0 * (q @ bank.T)
    if dtype == np.float64:
        pass
    if exclude_self:
        np.fill_diagonal(d2, np.inf)
    idx = _stable_topk(-d2, k)
    return idx, np.take_along_axis(d2, idx, 1), d2


def l2_topk_exact(q: np.ndarray, bank: np.ndarray, k: int):
    """Same as l2_topk but with the distances of the winners recomputed as sum((q-b)^2) in
    float64 (no cancellation)."""
    idx, _, d2 = l2_topk(q, bank, k)
    q64, b64 = np.asarray(q, np.float64), np.asarray(bank, np.float64)
    d = ((q64[:, None, :] - b64[idx]) ** 2).sum(-1)
    return idx, d, d2


def class_counts(idx: np.ndarray, labels: np.ndarray, n_classes: int = 3) -> np.ndarray:
    """chroma.py:112-121 — count labels among the neighbours."""
    lab = labels[idx]
    return np.stack([(lab == c).sum(1) for c in range(n_classes)], 1).astype(np.int32)


def vote_pass2(counts: np.ndarray) -> np.ndarray:
    """chroma.py:126-128 — np.argmax: first maximum wins => left > right > none on ties."""
    return np.argmax(counts, axis=1).astype(np.int32)


def vote_pass1(counts: np.ndarray, threshold: int = 20) -> np.ndarray:
    """chroma.py:62 — accept only if max(count) >= 20 (of 25); -1 == 'pass2' (deferred)."""
    out = vote_pass2(counts)
    out[counts.max(1) < threshold] = -1
    return out


def temp_smax(x, temperature=1.0):
    """chroma.py:28-34."""
    x = np.array(x, dtype=np.float64) / temperature
    e = np.exp(x - np.max(x))
    return e / e.sum()


def hmm_vote_scores(counts: np.ndarray, k: int) -> np.ndarray:
    """generate_clips_hmm.py:204-238 with one-hot stored probabilities (static banks): the mean
    stored prob of class c equals count_c/k, so score_c = (count_c/k + count_c/k)/2 = count_c/k."""
    return counts.astype(np.float64) / float(k)


def margins(sorted_vals: np.ndarray, k: int) -> np.ndarray:
    """|value_k - value_{k+1}| per row, given values sorted in result order with >= k+1 columns:
    the reference's similarity margin the index-exactness bar is stated on."""
    return np.abs(sorted_vals[:, k - 1] - sorted_vals[:, k])


def count_margin(counts: np.ndarray) -> np.ndarray:
    """top class count minus runner-up count (label robustness)."""
    s = np.sort(counts, axis=1)
    return s[:, -1] - s[:, -2]


def retriever_restated(q: np.ndarray, cand: np.ndarray, top_k: int = 10, valid_mask=None):
    """ratt_chunk_retriever.py:113-157 for one side group: normalise, GEMM, optional per-query
    candidate mask, top_k, gather, zero-pad -> (B, top_k, D) float32."""
    qn, cn = normalise(q), normalise(cand)
    sims = qn @ cn.T
    B, D = qn.shape
    out = np.zeros((B, top_k, D), np.float32)
    for i in range(B):
        valid = np.arange(cn.shape[0]) if valid_mask is None else np.where(valid_mask[i])[0]
        if len(valid) == 0:
            continue
        vs = sims[i, valid]
        order = np.argsort(-vs, kind="stable")[:top_k]
        chosen = valid[order]
        out[i, : len(chosen)] = cn[chosen]
    return out


def load_banks(directory):
    """clustering.py:26-36 — (rows (1472,768) f32, labels i32); accepts (N,1,768)."""
    rows, labels = [], []
    for li, name in enumerate(LABELS):
        e = np.load(f"{directory}/{name}_embeddings.npz")["embeddings"].astype(np.float32)
        e = e.reshape(e.shape[0], -1)
        rows.append(e)
        labels.append(np.full(len(e), li, np.int32))
    return np.concatenate(rows), np.concatenate(labels)


# ---------------------------------------------------------------------------------------------
# Retrievers as CLASSES (B6): RattChunkRetriever(collection, top_k, search_k)(chunk_embs, metadata)
# restated from retrieval/ratt_chunk_retriever.py:62-159 and FrameRetriever from
# retrieval/frame_retriever.py:3-72, over a numpy stand-in for the Chroma collection.
# PINNED: tools/make_golden_retriever.py imports the reference's OWN two classes from
# /root/reference (they are pure numpy) and runs them on the seeded cases below through the same
# stand-in collection; tests/golden/retriever_golden.npz holds their outputs, and
# tests/test_oracle_cpu.py checks these restatements against them.
# ---------------------------------------------------------------------------------------------
class ScalarBox:
    """Stands in for one tf scalar tensor: `.numpy()` returns the value (bytes for strings)."""

    def __init__(self, v):
        self.v = v

    def numpy(self):
        return self.v.encode() if isinstance(self.v, str) else self.v


def _match(meta, where):
    for key, cond in where.items():
        if key == "$and":
            if not all(_match(meta, w) for w in cond):
                return False
        elif isinstance(cond, dict):
            for op, val in cond.items():
                x = meta[key]
                ok = {"$gte": x >= val, "$lte": x <= val, "$ne": x != val, "$eq": x == val}[op]
                if not ok:
                    return False
        elif meta[key] != cond:
            return False
    return True


class NumpyCollection:
    """What the two retrievers need of a chromadb Collection ('cosine' space): get(where, include)
    and query(query_embeddings, n_results, where, include) by float64 brute force."""

    def __init__(self, rows, metadatas):
        self.rows = np.asarray(rows, np.float32)
        self.metadatas = list(metadatas)

    def get(self, where=None, include=()):
        sel = [i for i, m in enumerate(self.metadatas) if where is None or _match(m, where)]
        return {"embeddings": self.rows[sel], "metadatas": [self.metadatas[i] for i in sel]}

    def query(self, query_embeddings, n_results=10, where=None, include=()):
        out = []
        r64 = self.rows.astype(np.float64)
        rn = r64 / np.linalg.norm(r64, axis=1, keepdims=True)
        for q in np.asarray(query_embeddings, np.float64):
            sel = np.array([i for i, m in enumerate(self.metadatas) if where is None or _match(m, where)], np.int64)
            if len(sel) == 0:
                out.append(np.zeros((0, self.rows.shape[1]), np.float32))
                continue
            d = 1.0 - (rn[sel] @ (q / np.linalg.norm(q)))
            order = np.argsort(d, kind="stable")[:n_results]
            out.append(self.rows[sel[order]])
        return {"embeddings": out}


RETRIEVER_CASES = {   # name -> (rows, d, queries, top_k, seed)
    "small_d64": (600, 64, 24, 5, 11),
    "clips_d768": (3000, 768, 32, 10, 12),
    "sparse_d64": (400, 64, 16, 10, 13),     # narrow windows: most queries find < top_k rows
}


def retriever_case(name):
    """Seeded store + query batch of one golden case: (rows f32 (m,d), metadatas, queries f32
    (B,d), metadata dict of ScalarBox lists).  Metadata as db_rebuild_chunk.py writes it: vid_num
    int, side 'left'/'right', t_center / t_norm floats in [0,1]."""
    m, d, B, top_k, seed = RETRIEVER_CASES[name]
    rng = np.random.default_rng(seed)
    centres = rng.standard_normal((8, d)).astype(np.float32)
    rows = (centres[rng.integers(0, 8, m)] + 0.9 * rng.standard_normal((m, d))).astype(np.float32)
    metas = [{"vid_num": int(rng.integers(1, 8)), "side": "left" if rng.random() < 0.5 else "right",
              "t_center": float(rng.random()), "t_norm": float(rng.random()), "clip": f"c{i}"}
             for i in range(m)]
    q = (centres[rng.integers(0, 8, B)] + 0.9 * rng.standard_normal((B, d))).astype(np.float32)
    width = 0.04 if name.startswith("sparse") else 0.5
    md = {"vid": [ScalarBox(int(rng.integers(1, 8))) for _ in range(B)],
          "side": [ScalarBox("left" if rng.random() < 0.5 else "right") for _ in range(B)],
          "t_center": [ScalarBox(float(rng.random())) for _ in range(B)],
          "t_width": [ScalarBox(float(width * (0.5 + rng.random()))) for _ in range(B)]}
    return rows, metas, q, md, top_k


def ratt_chunk_retriever_restated(collection, chunk_embs, metadata, top_k=10):
    """retrieval/ratt_chunk_retriever.py:68-159 -> (out (B,top_k,D) f32, idx (B,top_k) i64 row of
    the collection or -1, gap (B,) f32: the smallest similarity gap that decided the selection)."""
    q = np.asarray(chunk_embs, np.float32).copy()
    q /= (np.linalg.norm(q, axis=1, keepdims=True) + 1e-8)
    B, D = q.shape
    out = np.zeros((B, top_k, D), np.float32)
    idx_out = np.full((B, top_k), -1, np.int64)
    gap = np.full(B, np.inf, np.float32)
    vids = np.array([int(metadata["vid"][i].numpy()) for i in range(B)], np.int32)
    sides = [metadata["side"][i].numpy().decode() for i in range(B)]
    tc = np.array([float(metadata["t_center"][i].numpy()) for i in range(B)], np.float32)
    tw = np.array([float(metadata["t_width"][i].numpy()) for i in range(B)], np.float32)
    for side in sorted(set(sides)):
        idxs = [i for i in range(B) if sides[i] == side]
        qs, qe = tc[idxs] - tw[idxs] / 2.0, tc[idxs] + tw[idxs] / 2.0
        g0, g1 = float(np.min(qs)), float(np.max(qe))
        sel = [i for i, m in enumerate(collection.metadatas)
               if m["side"] == side and m["t_center"] >= g0 and m["t_center"] <= g1]
        if not sel:
            continue
        cand = collection.rows[sel].astype(np.float32)
        cand /= (np.linalg.norm(cand, axis=1, keepdims=True) + 1e-8)
        cvid = np.array([int(collection.metadatas[i]["vid_num"]) for i in sel], np.int32)
        ct = np.array([float(collection.metadatas[i]["t_center"]) for i in sel], np.float32)
        sims = q[idxs] @ cand.T
        for lj, i in enumerate(idxs):
            valid = np.where((cvid != vids[i]) & (ct >= tc[i] - tw[i] / 2.0) & (ct <= tc[i] + tw[i] / 2.0))[0]
            if len(valid) == 0:
                continue
            vs = sims[lj, valid]
            order = np.argsort(-vs, kind="stable")
            chosen = valid[order[:top_k]]
            out[i, : len(chosen)] = cand[chosen]
            idx_out[i, : len(chosen)] = np.array(sel)[chosen]
            sv = vs[order[: top_k + 1]]
            if len(sv) > 1:
                gap[i] = np.min(sv[:-1] - sv[1:])
    return out, idx_out, gap
