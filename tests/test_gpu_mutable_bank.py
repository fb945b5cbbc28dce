"""GPU tests of the mutable bank and filtered queries:
vitk_bank_write (append / overwrite, growth across the direct -> large-store boundary) against a
bank set in one piece, vitk_knn_vote_masked (Chroma where= filters) against a bank holding only the
rows in play, the Chroma-style collection surface, and the ordered confident-upsert stream
(generate_clips_hmm.py:384-470) against the reference's frame-at-a-time loop."""
import numpy as np
import pytest
import torch

from oracle import selftrain_ref as ref

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


def _search(index, q, k):
    idx, score, label, counts = index.search(q, k, normalise_queries=(index.metric_name == "cosine"))
    torch.cuda.synchronize()
    return idx.cpu().numpy(), score.cpu().numpy(), label.cpu().numpy(), counts.cpu().numpy()


@pytest.mark.parametrize("metric", ["l2", "cosine"])
@pytest.mark.parametrize("m0,m1", [(300, 1472), (1900, 2300), (3000, 9000)])
def test_bank_write_equals_bank_set(v, emb, m0, m1, metric):
    """Appending in ragged pieces (host and device rows) and overwriting rows gives bit for bit the
    results of a bank set in one piece -- below, across and above the 2048-row direct limit."""
    rng = np.random.default_rng(m0 + m1)
    d = 768
    rows = rng.standard_normal((m1, d)).astype(np.float32)
    labels = rng.integers(0, 3, m1).astype(np.int32)
    q = torch.from_numpy(rng.standard_normal((37, d)).astype(np.float32)).cuda()
    norm = metric == "cosine"
    stale = rows[:m0].copy()
    stale[5] = -stale[5]                       # row 5 starts wrong and is overwritten below
    index = v.BankIndex(emb, stale, labels[:m0], 3, metric=metric, normalise_rows=norm)
    at, piece = m0, 0
    while at < m1:
        n = min([16, 1, 333, 16, 2048][piece % 5], m1 - at)
        if piece % 2:
            index.append(torch.from_numpy(rows[at:at + n]).cuda(), labels[at:at + n])
        else:
            index.append(rows[at:at + n], labels[at:at + n])
        at += n
        piece += 1
    index.write(5, rows[5:6], labels[5:6])
    assert index.m == m1 and index.lib.vitk_bank_rows(index.handle) == m1
    got = _search(index, q, 25)
    fresh = v.BankIndex(emb, rows, labels, 3, metric=metric, normalise_rows=norm)
    want = _search(fresh, q, 25)
    for g, w in zip(got, want):
        assert np.array_equal(g, w)


def test_bank_write_rejects_bad_arguments(v, emb):
    rows = np.random.default_rng(0).standard_normal((100, 768)).astype(np.float32)
    index = v.BankIndex(emb, rows, np.zeros(100, np.int32), 3)
    with pytest.raises(v.VitkError):
        index.write(101, rows[:1], np.zeros(1, np.int32))      # hole
    with pytest.raises(v.VitkError):
        index.append(rows[:1])                                  # labelled bank, no label
    with pytest.raises(v.VitkError):
        index.append(rows[:1], np.array([3], np.int32))         # label out of range
    index.append(rows[:0], np.zeros(0, np.int32))               # empty upsert: no-op
    assert index.m == 100


@pytest.mark.parametrize("metric", ["l2", "cosine"])
@pytest.mark.parametrize("m,keep", [(1472, 0.5), (1472, 0.01), (6000, 0.3), (6000, 0.002)])
def test_masked_query_equals_query_of_the_subset(v, emb, m, keep, metric):
    """where= filter: the rows in play behave exactly like a bank of their own (same scores, same
    order, same vote); with fewer than k rows in play the tail is padded with -1."""
    rng = np.random.default_rng(int(m * keep) + (metric == "l2"))
    rows = rng.standard_normal((m, 768)).astype(np.float32)
    labels = rng.integers(0, 3, m).astype(np.int32)
    mask = rng.random(m) < keep
    mask[rng.integers(0, m)] = True
    sel = np.flatnonzero(mask)
    q = torch.from_numpy(rng.standard_normal((19, 768)).astype(np.float32)).cuda()
    norm, k = metric == "cosine", 25
    index = v.BankIndex(emb, rows, labels, 3, metric=metric, normalise_rows=norm)
    idx, score, label, counts = index.search(q, k, normalise_queries=norm, row_mask=mask)
    idx, score, label, counts = idx.cpu().numpy(), score.cpu().numpy(), label.cpu().numpy(), counts.cpu().numpy()
    sub = v.BankIndex(emb, rows[sel], labels[sel], 3, metric=metric, normalise_rows=norm)
    sidx, sscore, slabel, scounts = _search(sub, q, k)
    kk = min(k, len(sel))
    assert (idx[:, kk:] == -1).all() and (sidx[:, kk:] == -1).all()
    assert np.array_equal(idx[:, :kk], sel[sidx[:, :kk]])
    assert np.array_equal(score[:, :kk], sscore[:, :kk])
    assert np.array_equal(counts, scounts) and np.array_equal(label, slabel)
    assert mask[idx[:, :kk]].all()


def _collection(v, emb, rng, n_per=60, d=768, spread=0.05, soft=False):
    centres = rng.standard_normal((3, d)).astype(np.float32)
    rows = np.concatenate([centres[c] + spread * rng.standard_normal((n_per, d)) for c in range(3)]).astype(np.float32)
    metas = []
    for i in range(3 * n_per):
        c = i // n_per
        md = v.knn.one_hot_metadata(ref.NAMES[c], "vid0")
        if soft and i % 3 == 0:
            p = rng.dirichlet([8.0 if s == c else 1.0 for s in range(3)])
            md.update(left_prob=float(p[0]), right_prob=float(p[1]), none_prob=float(p[2]))
        md["t_norm"] = float(i % n_per) / n_per
        metas.append(md)
    ids = ["b%d" % i for i in range(3 * n_per)]
    return centres, rows, ids, metas


def test_collection_upsert_delete_where(v, emb):
    rng = np.random.default_rng(11)
    centres, rows, ids, metas = _collection(v, emb, rng)
    col = v.BankCollection(emb, rows, ids, metas, space="l2")
    assert col.count() == 180
    # a query under the frame_retriever filter only returns matching rows, nearest first
    w = {"$and": [{"label": "right"}, {"t_norm": {"$gte": 0.25}}, {"t_norm": {"$lte": 0.5}}]}
    r = col.query(centres[0:1], n_results=10, where=w, include=["metadatas", "distances", "embeddings"])
    assert len(r["ids"][0]) == 10 and all(v.knn.match_where(m, w) for m in r["metadatas"][0])
    assert r["distances"][0] == sorted(r["distances"][0])
    want = sorted(float(((rows[i] - centres[0]) ** 2).sum()) for i in range(180) if v.knn.match_where(metas[i], w))
    np.testing.assert_allclose(r["distances"][0], want[:10], rtol=1e-4)
    # upsert: a new id becomes the nearest neighbour of its own embedding; an existing id moves
    new = (centres[2] * 3.0).astype(np.float32)
    col.upsert(new[None], ["f_new"], [v.knn.one_hot_metadata("none", "vid9")])
    assert col.count() == 181 and col.query(new[None], n_results=1)["ids"][0] == ["f_new"]
    col.upsert((centres[1] * 5.0)[None], ["b0"], [v.knn.one_hot_metadata("right", "vid0")])
    assert col.count() == 181
    got = col.query((centres[1] * 5.0)[None], n_results=1, include=["metadatas"])
    assert got["ids"][0] == ["b0"] and got["metadatas"][0][0]["label"] == "right"
    assert col.get(ids=["b0"], include=["embeddings"])["embeddings"].shape == (1, 768)
    # delete(where=) like db_rebuild.py:121
    col.delete(where={"video": {"$ne": "vid0"}})
    assert col.count() == 180 and "f_new" not in col.ids
    assert col.query(new[None], n_results=1)["ids"][0] != ["f_new"]


@pytest.mark.parametrize("soft,spread", [(False, 0.05), (True, 0.6)])
def test_confident_upsert_stream_equals_frame_at_a_time_loop(v, emb, soft, spread):
    """The blocked GPU stream classifies every frame against exactly the bank the sequential loop
    shows it: same sides, same probabilities, same upserted rows -- with unanimous neighbourhoods
    and with mixed ones (soft stored probabilities, overlapping clusters, unconfident frames)."""
    rng = np.random.default_rng(5 + soft)
    centres, rows, ids, metas = _collection(v, emb, rng, spread=spread, soft=soft)
    seq = rng.integers(0, 3, 300)
    frames = (centres[seq] + spread * rng.standard_normal((300, 768))).astype(np.float32)
    if soft:   # every fifth frame sits between two classes: mixed neighbourhoods, not confident
        other = (seq + 1 + rng.integers(0, 2, 300)) % 3
        mid = (0.5 * (centres[seq] + centres[other]) + spread * rng.standard_normal((300, 768))).astype(np.float32)
        frames[::5] = mid[::5]
    names = ["vid7_frame_%d.jpg" % i for i in range(300)]
    k = 25

    col_a = v.BankCollection(emb, rows, ids, metas, space="l2")

    def query_one(e, n_results):
        r = col_a.query(e[None], n_results=n_results, include=["metadatas", "distances"])
        return r["ids"][0], r["metadatas"][0], r["distances"][0]

    want = ref.sequential_loop(query_one, col_a.upsert, frames, names, k, "vid7")
    ids_a, metas_a = list(col_a.ids), [dict(m) for m in col_a.metadatas]

    col_b = v.BankCollection(emb, rows, ids, metas, space="l2")
    got = v.knn.classify_stream_with_upserts(col_b, frames, names, top_n_closest=k, cur_vid="vid7", block=64)
    assert got["upserts"] == want["upserts"] and got["left_over"] == want["left_over"]
    assert got["upserts"] >= 3
    assert len(got["sides"]) == 300
    for a, b in zip(got["sides"], want["sides"]):
        assert a[0] == b[0] and float(a[1]) == float(b[1]) and len(a) == len(b)
        if len(a) == 3:
            assert a[2] == b[2]
    assert col_b.ids == ids_a and col_b.metadatas == metas_a
    if not soft:
        assert [s[0] for s in got["sides"]] == [ref.NAMES[c] for c in seq]
    else:
        assert any(len(s) == 2 for s in got["sides"])          # some frames were not confident


def test_confident_upsert_stream_matches_numpy_oracle(v, emb):
    """Against the float64 brute-force collection on well separated clusters (every decision has
    margin): same labels, same confident set, same upsert order."""
    rng = np.random.default_rng(21)
    centres, rows, ids, metas = _collection(v, emb, rng, n_per=40)
    seq = rng.integers(0, 3, 150)
    frames = (centres[seq] + 0.05 * rng.standard_normal((150, 768))).astype(np.float32)
    names = ["f%d" % i for i in range(150)]
    ncol = ref.NumpyCollection(rows, ids, metas)
    want = ref.sequential_loop(ncol.query, ncol.upsert, frames, names, 25, "v")
    col = v.BankCollection(emb, rows, ids, metas, space="l2")
    got = v.knn.classify_stream_with_upserts(col, torch.from_numpy(frames).cuda(), names, top_n_closest=25,
                                             cur_vid="v")
    assert [s[0] for s in got["sides"]] == [s[0] for s in want["sides"]]
    assert [len(s) for s in got["sides"]] == [len(s) for s in want["sides"]]
    assert got["upserts"] == want["upserts"] and col.ids == ncol.ids
