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
