"""Golden vectors for the retriever drop-ins (B6 / a10), produced by the reference's OWN classes:
retrieval/ratt_chunk_retriever.py and retrieval/frame_retriever.py are pure numpy, so they are
imported from /root/reference as they stand and run on seeded stores (oracle/knn_ref.retriever_case)
through a numpy stand-in for the Chroma collection (oracle/knn_ref.NumpyCollection: get / query
with the where-clauses the two classes send).  Stored: the (B, top_k, D) outputs for the 64-wide
cases in full, and for every case the chosen collection rows (recovered by matching the returned
vectors back to the normalised store) plus the similarity gap that decided each selection.
Nothing of the reference is copied into the repo.

Run in the build container (needs /root/reference):  python tools/make_golden_retriever.py
"""
import importlib.util
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import knn_ref  # noqa: E402

REF = "/root/reference/nba_proj/retrieval"


def _load(name):
    spec = importlib.util.spec_from_file_location(name, os.path.join(REF, name + ".py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def _rows_of(out, store_n):
    """(B, top_k) collection row of every returned vector (-1 for the zero padding)."""
    B, K, _ = out.shape
    idx = np.full((B, K), -1, np.int64)
    for i in range(B):
        for j in range(K):
            if not out[i, j].any():
                continue
            hit = np.nonzero((store_n == out[i, j]).all(1))[0]
            assert len(hit) >= 1, "returned vector is not a normalised store row"
            idx[i, j] = hit[0]
    return idx


def main():
    ratt = _load("ratt_chunk_retriever").RattChunkRetriever
    frame = _load("frame_retriever").FrameRetriever
    out = {"cases": np.array(list(knn_ref.RETRIEVER_CASES))}
    for name in knn_ref.RETRIEVER_CASES:
        rows, metas, q, md, top_k = knn_ref.retriever_case(name)
        col = knn_ref.NumpyCollection(rows, metas)
        store_n = rows / (np.linalg.norm(rows, axis=1, keepdims=True) + 1e-8)
        r = ratt(col, top_k=top_k, search_k=200)(q.copy(), md)
        _, idx_restated, gap = knn_ref.ratt_chunk_retriever_restated(col, q, md, top_k)
        idx = _rows_of(r, store_n.astype(np.float32))
        out[name + "/ratt_idx"] = idx.astype(np.int32)
        out[name + "/ratt_gap"] = gap
        out[name + "/ratt_sum"] = np.float64(r.astype(np.float64).sum())
        if rows.shape[1] == 64:
            out[name + "/ratt_out"] = r
        f = frame(col, top_k=top_k, search_k=200)(q.copy(), md)
        out[name + "/frame_sum"] = np.float64(f.astype(np.float64).sum())
        out[name + "/frame_idx"] = _rows_of(f, store_n.astype(np.float32)).astype(np.int32)
        if rows.shape[1] == 64:
            out[name + "/frame_out"] = f
        print(name, "ratt valid rows per query:", (idx >= 0).sum(1).tolist()[:12], "restatement agrees:",
              bool(np.array_equal(idx, idx_restated)))
    path = os.path.join(ROOT, "tests", "golden", "retriever_golden.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
