"""GPU parity of the retrieval boundary: n_results up to 1024,
where-clauses, per-query row predicates evaluated inside the k-NN kernels, and the two retriever
classes as drop-ins -- against numpy brute force and against golden outputs of the reference's OWN
RattChunkRetriever / FrameRetriever (tests/golden/retriever_golden.npz, produced by importing
them from /root/reference: tools/make_golden_retriever.py).

Bars: row lists exact wherever the deciding similarity gap exceeds the stated epsilon; distances
to 2e-5 (cosine) / 2e-4 relative (squared L2); padding identical.
"""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def v():
    import vit_research_b200 as v
    assert torch.cuda.is_available()
    v.load()
    return v


@pytest.fixture(scope="module")
def emb(v):
    from oracle import vit_ref
    model = vit_ref.build_hf_model("vit_b16_l2", 0, randomise_ln=True)
    e = v.ViTEmbedder.from_hf(model, device=0, max_batch=8)
    yield e
    e.close()


def _unit(x):
    return (x / (np.linalg.norm(x, axis=1, keepdims=True) + 1e-8)).astype(np.float32)


def _brute(q, store, k, metric, valid=None):
    """float64 brute force: (idx (Q,k) with -1 padding, value (Q,k+1) sorted in result order)."""
    q64, s64 = q.astype(np.float64), store.astype(np.float64)
    if metric == "cosine":
        key = q64 @ s64.T
    else:
        key = -((q64 * q64).sum(1)[:, None] + (s64 * s64).sum(1)[None, :] - 2.0 * (q64 @ s64.T))
    if valid is not None:
        key = np.where(valid, key, -np.inf)
    order = np.argsort(-key, axis=1, kind="stable")[:, : k + 1]
    val = np.take_along_axis(key, order, 1)
    idx = np.where(np.isfinite(val), order, -1)
    _brute.key = key          # full key matrix of the last call (larger == better)
    return idx[:, :k], val


@pytest.mark.parametrize("metric", ["cosine", "l2"])
@pytest.mark.parametrize("m,k", [(1472, 100), (2048, 750), (1500, 1024), (30000, 100), (120000, 750), (60000, 1024)])
def test_n_results_beyond_64_match_bruteforce(v, emb, m, k, metric):
    """The clip-store queries ask for 100-750 neighbours (frame_retriever.py:41-43, config.py:10-11):
    the ordered lists equal float64 brute force, on the score-matrix path (m <= 2048) and the pool
    path, for both metrics."""
    rng = np.random.default_rng(m + k)
    store = _unit(rng.standard_normal((m, 256)) + 0.6 * rng.standard_normal((1, 256)))
    q = _unit(store[rng.integers(0, m, 12)] + 0.3 * rng.standard_normal((12, 256)))
    index = v.BankIndex(emb, store, None, 0, metric=metric)
    idx, score, _, _ = index.search(torch.from_numpy(q).cuda(), k)
    idx, score = idx.cpu().numpy(), score.cpu().numpy()
    oi, ov = _brute(q, store, k, metric)
    want = ov[:, :k] if metric == "cosine" else -ov[:, :k]
    np.testing.assert_allclose(score, want, atol=3e-5 if metric == "cosine" else 3e-4)
    assert (idx >= 0).all() and all(len(set(r)) == k for r in idx)
    # the SET is exact where the k-th / (k+1)-th gap is comfortable; the ORDER wherever every
    # adjacent gap inside the list is above fp32 rounding
    # every returned row belongs: its true key is not below the k-th best by more than rounding
    tol = 3e-6 if metric == "cosine" else 3e-5
    true_key = np.take_along_axis(_brute.key, idx.astype(np.int64), 1)
    assert (true_key >= ov[:, k - 1:k] - tol).all()
    gap_k = np.abs(ov[:, k - 1] - ov[:, k])
    safe = gap_k > (2e-5 if metric == "cosine" else 1e-4)
    assert np.array_equal(np.sort(idx[safe], 1), np.sort(oi[safe], 1))
    for r in range(len(q)):
        ok = np.abs(np.diff(ov[r, :k])) > (2e-6 if metric == "cosine" else 2e-5)
        strict = np.concatenate([[True], ok]) & np.concatenate([ok, [True]])
        if safe[r]:
            assert np.array_equal(idx[r][strict], oi[r][strict])


def test_query_750_with_where_clause_is_the_reference_call_shape(v, emb):
    """frame_retriever.py:41-53: collection.query(n_results=search_k, where={$and: [vid_num $ne,
    side, t_norm $gte, t_norm $lte]}, include=["embeddings"]) -> 750 ordered rows identical to
    brute force over the rows the clause keeps; nothing is silently shortened."""
    rng = np.random.default_rng(8)
    m = 9000
    rows = rng.standard_normal((m, 128)).astype(np.float32)
    metas = [{"vid_num": int(rng.integers(1, 5)), "side": "left" if rng.random() < 0.6 else "right",
              "t_norm": float(rng.random())} for _ in range(m)]
    ids = [f"c{i}" for i in range(m)]
    col = v.BankCollection(emb, rows, ids, metas, space="cosine")
    q = rng.standard_normal((3, 128)).astype(np.float32)
    where = {"$and": [{"vid_num": {"$ne": 2}}, {"side": "left"}, {"t_norm": {"$gte": 0.1}}, {"t_norm": {"$lte": 0.9}}]}
    res = col.query(query_embeddings=[x.tolist() for x in q], n_results=750, where=where,
                    include=["embeddings", "distances", "metadatas"])
    keep = np.array([mm["vid_num"] != 2 and mm["side"] == "left" and 0.1 <= mm["t_norm"] <= 0.9 for mm in metas])
    assert keep.sum() > 750
    oi, ov = _brute(_unit(q), _unit(rows), 750, "cosine", valid=np.broadcast_to(keep, (3, m)))
    for r in range(3):
        assert len(res["ids"][r]) == 750 and len(res["embeddings"][r]) == 750
        got = [int(i[1:]) for i in res["ids"][r]]
        assert all(keep[j] for j in got)
        d = np.array(res["distances"][r])
        np.testing.assert_allclose(d, 1.0 - ov[r, :750], atol=3e-5)
        ok = np.abs(np.diff(ov[r, :751])) > 2e-6
        strict = (np.concatenate([[True], ok[:-1]]) & ok)
        assert np.array_equal(np.array(got)[strict], oi[r][strict])
        assert np.array_equal(res["embeddings"][r][0], rows[got[0]])
    # more than the collection holds: like chromadb, what there is; beyond the limit: an error
    small = v.BankCollection(emb, rows[:40], ids[:40], metas[:40], space="l2")
    assert len(small.query(q[:1], n_results=100)["ids"][0]) == 40
    with pytest.raises(v.VitkError):
        col.query(q[:1], n_results=2000)
    with pytest.raises(v.VitkError):
        col.index.search(torch.from_numpy(q).cuda(), 1025)


@pytest.mark.parametrize("m", [1800, 40000])
@pytest.mark.parametrize("metric", ["cosine", "l2"])
def test_per_query_predicate_equals_bruteforce_over_each_querys_rows(v, emb, m, metric):
    """vitk_knn_vote_filtered: every query has its OWN set of rows in play -- (col != q_ne) &
    (col == q_eq) & (lo <= x <= hi), float32-rounded and raw windows -- evaluated inside the
    kernels.  Selective windows leave some queries with fewer than k rows (padded with -1)."""
    rng = np.random.default_rng(m)
    d, nq, k = 128, 40, 20
    store = _unit(rng.standard_normal((m, d)))
    vid = rng.integers(1, 6, m).astype(np.int32)
    side = rng.integers(0, 2, m).astype(np.int32)
    t = rng.random(m)
    q = _unit(rng.standard_normal((nq, d)))
    q_vid = rng.integers(1, 6, nq).astype(np.int32)
    q_side = rng.integers(0, 2, nq).astype(np.int32)
    centre = rng.random(nq).astype(np.float32)
    width = np.where(np.arange(nq) % 4 == 0, 20.0 / m, 0.5).astype(np.float32)  # every 4th: a sliver (~8 rows)
    lo, hi = centre - width / 2.0, centre + width / 2.0                        # float32 like the reference
    g_lo, g_hi = float(lo.min()), float(hi.max())
    t32 = t.astype(np.float32)
    valid = ((vid[None, :] != q_vid[:, None]) & (side[None, :] == q_side[:, None]) &
             (t32[None, :] >= lo[:, None]) & (t32[None, :] <= hi[:, None]) &
             (t[None, :] >= g_lo) & (t[None, :] <= g_hi))
    index = v.BankIndex(emb, store, None, 0, metric=metric)
    dev = "cuda"
    filt = {"ne": (torch.from_numpy(vid).to(dev), torch.from_numpy(q_vid).to(dev)),
            "eq": (torch.from_numpy(side).to(dev), torch.from_numpy(q_side).to(dev)),
            "rng": (torch.from_numpy(t).to(dev), torch.from_numpy(lo.astype(np.float64)).to(dev),
                    torch.from_numpy(hi.astype(np.float64)).to(dev)),
            "rng2": (torch.full((nq,), g_lo, dtype=torch.float64, device=dev),
                     torch.full((nq,), g_hi, dtype=torch.float64, device=dev)),
            "rng_round_f32": True}
    idx, score, _, _ = index.search(torch.from_numpy(q).cuda(), k, row_filter=filt)
    idx, score = idx.cpu().numpy(), score.cpu().numpy()
    oi, ov = _brute(q, store, k, metric, valid=valid)
    n_valid = valid.sum(1)
    assert (n_valid < k).any() and (n_valid >= k).any()
    for r in range(nq):
        nv = min(int(n_valid[r]), k)
        assert (idx[r, nv:] == -1).all() and (idx[r, :nv] >= 0).all()
        assert all(valid[r, j] for j in idx[r, :nv])
        want = ov[r, :nv] if metric == "cosine" else -ov[r, :nv]
        np.testing.assert_allclose(score[r, :nv], want, atol=3e-5 if metric == "cosine" else 3e-4)
        assert set(idx[r, :nv]) == set(oi[r, :nv]) or np.abs(ov[r, nv - 1] - ov[r, nv]) < 1e-4
    # a single mask on top of the predicate composes (both must hold)
    mask = (np.arange(m) % 3 != 0).astype(np.uint8)
    idx2, _, _, _ = index.search(torch.from_numpy(q).cuda(), k, row_mask=mask, row_filter=filt)
    idx2 = idx2.cpu().numpy()
    oi2, _ = _brute(q, store, k, metric, valid=valid & (mask[None, :] != 0))
    for r in range(nq):
        assert set(idx2[r][idx2[r] >= 0]) <= set(np.nonzero(valid[r] & (mask != 0))[0])
        assert (idx2[r] >= 0).sum() == (oi2[r] >= 0).sum()


def _collection_of_case(v, emb, name):
    from oracle import knn_ref
    rows, metas, q, md, top_k = knn_ref.retriever_case(name)
    col = v.BankCollection(emb, rows, [f"r{i}" for i in range(len(rows))], metas, space="cosine")
    return col, rows, q, md, top_k


@pytest.mark.parametrize("name", ["small_d64", "clips_d768", "sparse_d64"])
def test_ratt_chunk_retriever_equals_the_reference_class(v, emb, golden_dir, name):
    """B6: RattChunkRetriever(collection, top_k, search_k)(chunk_embs, metadata) -- the reference's
    constructor and call -- against the outputs of the reference's own class: chosen rows, order,
    zero padding, and the full (B, top_k, D) float32 tensor."""
    z = np.load(os.path.join(golden_dir, "retriever_golden.npz"))
    col, rows, q, md, top_k = _collection_of_case(v, emb, name)
    r = v.RattChunkRetriever(col, top_k=top_k, search_k=200)
    got = r(q.copy(), md)
    assert got.shape == (len(q), top_k, rows.shape[1]) and got.dtype == np.float32
    want_idx, gap = z[name + "/ratt_idx"], z[name + "/ratt_gap"]
    store_n = (rows / (np.linalg.norm(rows, axis=1, keepdims=True) + 1e-8)).astype(np.float32)
    want = np.where((want_idx >= 0)[..., None], store_n[np.maximum(want_idx, 0)], 0.0).astype(np.float32)
    if name + "/ratt_out" in z.files:
        assert np.array_equal(want, z[name + "/ratt_out"])        # the reconstruction IS the reference output
    safe = gap > 1e-5            # selections decided by a similarity gap above fp32 rounding
    assert safe.mean() > 0.8
    assert np.array_equal(got[safe], want[safe])
    # padding identical everywhere
    assert np.array_equal((np.abs(got).sum(2) == 0), (want_idx < 0))
    # like the reference, a float32 ndarray of queries is normalised in place
    q2 = q.copy()
    r(q2, md)
    np.testing.assert_allclose(np.linalg.norm(q2, axis=1), 1.0, atol=1e-5)


@pytest.mark.parametrize("name", ["small_d64", "sparse_d64"])
def test_frame_retriever_equals_the_reference_class(v, emb, golden_dir, name):
    z = np.load(os.path.join(golden_dir, "retriever_golden.npz"))
    col, rows, q, md, top_k = _collection_of_case(v, emb, name)
    got = v.FrameRetriever(col, top_k=top_k, search_k=200)(q.copy(), md)
    want = z[name + "/frame_out"]
    assert got.shape == want.shape
    same = np.array([np.array_equal(g, w) for g, w in zip(got, want)])
    assert same.mean() > 0.8          # the rest differ only by near-tied neighbours swapping places
    assert np.array_equal((np.abs(got).sum(2) == 0), (np.abs(want).sum(2) == 0))
    for g, w in zip(got[~same], want[~same]):     # at most the boundary neighbour differs
        assert len({x.tobytes() for x in g} & {x.tobytes() for x in w}) >= len({x.tobytes() for x in w}) - 1


def test_two_indexes_on_one_embedder_are_independent(v, emb):
    """Each BankIndex owns its bank on the handle: building a second one must not change what the
    first returns (the label bank next to a clip store, as the reference keeps several Chroma
    collections)."""
    rng = np.random.default_rng(0)
    a_rows = rng.standard_normal((300, 768)).astype(np.float32)
    b_rows = rng.standard_normal((5000, 768)).astype(np.float32)
    qa = torch.from_numpy(a_rows[:9] + 0.01).cuda()
    a = v.BankIndex(emb, a_rows, rng.integers(0, 3, 300).astype(np.int32), 3, metric="l2")
    first = [t.cpu().numpy() for t in a.search(qa, 5)]
    b = v.BankIndex(emb, b_rows, None, 0, metric="cosine", normalise_rows=True)
    ib = b.search(torch.from_numpy(b_rows[100:104]).cuda(), 3, normalise_queries=True)[0].cpu().numpy()
    assert list(ib[:, 0]) == [100, 101, 102, 103]
    again = [t.cpu().numpy() for t in a.search(qa, 5)]
    for x, y in zip(first, again):
        assert np.array_equal(x, y)
    assert list(again[0][:, 0]) == list(range(9)) and a.m == 300 and b.m == 5000
    b.close()
    with pytest.raises(v.VitkError):
        b.search(qa, 3)
    assert np.array_equal(a.search(qa, 5)[0].cpu().numpy(), first[0])


def test_bank_client_and_empty_collections(v, emb):
    """chromadb.PersistentClient(...).get_or_create_collection(name, metadata={"hnsw:space": ...})
    as the scripts call it (chroma.py:143-146): starts empty, fills by upsert, queries like Chroma."""
    client = v.BankClient(emb, path="./unused")
    col = client.get_or_create_collection(name="test_vid", metadata={"hnsw:space": "l2"})
    assert col.count() == 0 and col.query([[0.0] * 768], n_results=5)["ids"] == [[]]
    rng = np.random.default_rng(1)
    rows = rng.standard_normal((40, 768)).astype(np.float32)
    col.upsert(embeddings=rows, ids=[f"f{i}" for i in range(40)],
               metadatas=[v.one_hot_metadata(["left", "right", "none"][i % 3]) for i in range(40)])
    assert client.get_or_create_collection(name="test_vid") is col and col.count() == 40
    res = col.query(query_embeddings=rows[7:8], n_results=25)
    assert res["ids"][0][0] == "f7" and abs(res["distances"][0][0]) < 1e-3
    assert v.determine_class_pass2(res["ids"][0], res["metadatas"][0], res["distances"][0])["dir"] in v.LABEL_NAMES
    col.delete(where={"label": "none"})
    assert col.count() == 40 - 13
    client.delete_collection("test_vid")
    with pytest.raises(ValueError):
        client.get_collection("test_vid")


def test_merge_kernel_equals_the_host_merge(v):
    """shard.merge_topk on CUDA tensors (knn_merge_topk_kernel) against the torch host form the
    gloo tests pin: (score, then lower row id) order, -1 slots, fewer candidates than k."""
    from vit_research_b200 import shard
    g = torch.Generator().manual_seed(4)
    for c, k, desc in [(80, 10, True), (16, 10, True), (8, 10, False), (600, 64, False), (4096, 750, True)]:
        idx = torch.stack([torch.randperm(100000, generator=g)[:c] for _ in range(7)]).to(torch.int32)
        score = torch.randn((7, c), generator=g)
        score[:, ::5] = score[:, :1]                    # ties: the lower row id wins
        idx[:, 3::7] = -1
        wi, ws = shard.merge_topk(idx, score, k, desc)
        gi, gs = shard.merge_topk(idx.cuda(), score.cuda(), k, desc)
        wi, ws, gi, gs = wi.numpy(), ws.numpy(), gi.cpu().numpy(), gs.cpu().numpy()
        assert np.array_equal(gi, wi)
        live = wi >= 0
        assert np.array_equal(gs[live], ws[live])
