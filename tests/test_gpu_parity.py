"""GPU parity tests proper: the CUDA path, through the C ABI, against the oracle (HF ViTModel on
CPU fp32 + numpy brute-force k-NN) on the same seeded inputs, against the committed golden
fixtures, and — at BASELINE.json's full sizes — through size-independent properties.

Bars (BASELINE.json north_star): embeddings cosine >= 0.999 per frame (bf16 vs the reference's
fp32); labels and top-k indices bit-exact wherever the reference's similarity margin > 1e-3.
"""
import json
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

COS_BAR = 0.999


def _cos(a, b):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return (a * b).sum(1) / (np.linalg.norm(a, axis=1) * np.linalg.norm(b, axis=1))


@pytest.fixture(scope="module")
def v():
    import vit_research_b200 as v
    assert torch.cuda.is_available()
    v.load()
    return v


@pytest.fixture(scope="module")
def b16(v):
    """ViT-B/16 seed-0 oracle model + its vitk twin."""
    from oracle import vit_ref
    torch.set_num_threads(max(1, os.cpu_count() or 1))
    model = vit_ref.build_hf_model("vit_b16", 0, randomise_ln=True)
    emb = v.ViTEmbedder.from_hf(model, device=0, max_batch=64)
    yield model, emb
    emb.close()


def test_embed_matches_oracle_f32_pixels(v, b16):
    from oracle import vit_ref
    model, emb = b16
    px = vit_ref.synthetic_pixels(8, seed=1234)
    want = vit_ref.hf_embed(model, px, normalise=True)
    got = emb.embed(px.cuda(), l2_normalise=True).cpu().numpy()
    c = _cos(got, want)
    assert c.min() >= COS_BAR, c
    np.testing.assert_allclose(np.linalg.norm(got, axis=1), 1.0, atol=1e-5)
    # bf16 pixel input (the bench's device-resident format) meets the same bar
    got_b = emb.embed(px.cuda().bfloat16(), l2_normalise=True).cpu().numpy()
    assert _cos(got_b, want).min() >= COS_BAR


def test_embed_matches_golden_vectors(v, b16, golden_dir):
    """Committed HF outputs (tools/make_golden.py) — no oracle execution needed."""
    from oracle import vit_ref
    _, emb = b16
    z = np.load(os.path.join(golden_dir, "vit_b16_seed0.npz"))
    model, _ = b16
    if abs(vit_ref.weights_fingerprint(model) - float(z["fingerprint"])) > 1e-6 * float(z["fingerprint"]):
        pytest.skip("torch RNG stream differs from the golden generator's")
    got = emb.embed(vit_ref.synthetic_pixels(8, 1234).cuda(), l2_normalise=False).cpu().numpy()
    assert _cos(got, z["cls"]).min() >= COS_BAR
    # un-normalised CLS rows are LayerNorm outputs: compare magnitudes too
    assert np.abs(got - z["cls"]).max() < 0.25 and np.abs(got - z["cls"]).mean() < 0.03
    got8 = emb.embed_frames(vit_ref.synthetic_frames_u8(8, 4321), l2_normalise=False)
    assert _cos(got8, z["cls_u8"]).min() >= COS_BAR


def test_hf_vit_embed_batch_signature_and_parity(v, b16):
    """B1: (N,H,W,3) uint8 or float32 in -> (N,768) float32 C-contiguous unit rows out."""
    from oracle import vit_ref
    model, emb = b16
    v.install(emb)
    frames = vit_ref.synthetic_frames_u8(5, seed=77)
    out = v.hf_vit_embed_batch(frames)
    assert out.shape == (5, 768) and out.dtype == np.float32 and out.flags["C_CONTIGUOUS"]
    want = vit_ref.hf_embed(model, vit_ref.process_u8(frames), normalise=True)
    assert _cos(out, want).min() >= COS_BAR
    # float input is clipped to [0,255] and cast (ratt_write_clips_to_ragdb.py:140-141)
    out_f = v.hf_vit_embed_batch(frames.astype(np.float32) + 0.25)
    np.testing.assert_array_equal(out_f, out)
    # Keras-style and HF-style adaptors expose the same numbers
    pre = emb.predict(frames)["pre_logits"]
    assert pre.shape == (5, 1, 1, 768)
    hs = emb(pixel_values=vit_ref.process_u8(frames).cuda()).last_hidden_state[:, 0, :].cpu().numpy()
    np.testing.assert_allclose(pre.reshape(5, 768), hs, atol=1e-6)
    assert _cos(v.hf_vit_embed_batch(frames[:0]), want[:0]).shape == (0,)  # empty batch


def test_wrong_image_size_raises_like_hf(v, b16):
    _, emb = b16
    with pytest.raises(ValueError):
        emb.embed(torch.zeros(1, 3, 200, 224, device="cuda"))
    with pytest.raises(ValueError):
        emb.embed_frames(np.zeros((1, 100, 100, 4), np.uint8))   # wrong channel count
    # other frame SIZES no longer raise: like hf_vit_embed_batch, the processor's resize runs first
    assert emb.embed_frames(np.zeros((1, 100, 100, 3), np.uint8)).shape == (1, 768)


def test_chunking_is_invisible(v, b16):
    """n > max_batch is processed in chunks; ragged tails give identical rows."""
    from oracle import vit_ref
    _, emb = b16
    px = vit_ref.synthetic_pixels(70, seed=3).cuda()  # max_batch = 64 -> 64 + 6
    all_ = emb.embed(px)
    one = emb.embed(px[64:70])
    assert torch.equal(all_[64:70], one)
    assert torch.equal(emb.embed(px[:3]), all_[:3])


def test_vit_l16_geometry(v):
    from oracle import vit_ref
    model = vit_ref.build_hf_model("vit_l16_l2", 1, randomise_ln=True)
    emb = v.ViTEmbedder.from_hf(model, device=0, max_batch=8)
    px = vit_ref.synthetic_pixels(4, seed=1234)
    want = vit_ref.hf_embed(model, px, normalise=False)
    got = emb.embed(px.cuda(), l2_normalise=False).cpu().numpy()
    assert _cos(got, want).min() >= COS_BAR
    emb.close()


def test_tf_path_geometry_432x768_patch32(v):
    """The reference's TF-path hyper-parameters (generate_clips_hmm.py:48-56): 432x768 frames fed
    un-normalised (raw 0..255), patch 32 'valid' -> 313 tokens, LayerNorm eps 1e-6, tanh-GELU,
    (B,1,1,768) pre_logits.  Oracle: HF ViTModel configured with exactly that geometry (the TF
    Model Garden block is absent here, so TF semantics beyond the geometry stay unpinned)."""
    from oracle import vit_ref
    model = vit_ref.build_hf_model("vit_tf432_l2", 2, randomise_ln=True)
    emb = v.ViTEmbedder.from_hf(model, device=0, max_batch=4)
    assert emb.geometry.tokens == 313 and emb.geometry.gelu_tanh and emb.geometry.ln_eps == 1e-6
    g = torch.Generator().manual_seed(3)
    frames = torch.randint(0, 256, (6, 432, 768, 3), generator=g, dtype=torch.uint8)
    px = frames.permute(0, 3, 1, 2).float().contiguous()          # raw pixel values, NCHW
    want = vit_ref.hf_embed(model, px, normalise=False)
    got = emb.embed(px.cuda(), l2_normalise=False).cpu().numpy()  # 6 frames > max_batch 4: chunked
    assert got.shape == (6, 768)
    assert _cos(got, want).min() >= COS_BAR, _cos(got, want)
    # the CLS-only last layer and the full last layer agree bit for bit on this geometry too
    hs = emb(pixel_values=px[:2].cuda()).last_hidden_state
    assert hs.shape == (2, 1, 768) and torch.equal(hs[:, 0].cpu(), torch.from_numpy(got[:2]))
    emb.close()
    # Keras-style call on raw uint8 frames, as the TF scripts do (write_embeddings.py:185-188):
    # geometry.u8_mean/u8_std = 0/1 feeds the pixel values unchanged -> identical embeddings
    import dataclasses
    g_tf = dataclasses.replace(v.ViTGeometry.from_hf_config(model.config), u8_mean=0.0, u8_std=1.0)
    assert g_tf == dataclasses.replace(v.ViTGeometry.tf_432x768(), layers=2)
    emb2 = v.ViTEmbedder(g_tf, model.state_dict(), device=0, max_batch=4)
    out = emb2.predict(frames.numpy(), batch_size=1024, verbose=0)["pre_logits"]
    assert out.shape == (6, 1, 1, 768) and out.dtype == np.float32
    assert np.array_equal(out.reshape(6, 768), got)
    emb2.close()


# ------------------------------------------------------------------------------------------
# k-NN + vote
# ------------------------------------------------------------------------------------------
@pytest.fixture(scope="module")
def bank_index(v, b16, banks):
    _, emb = b16
    rows, labels = banks
    return rows, labels


@pytest.mark.parametrize("metric", ["l2", "cosine"])
@pytest.mark.parametrize("k", [5, 25, 50])
def test_knn_on_reference_banks_matches_oracle(v, b16, banks, golden_dir, metric, k):
    """Queries = bank rows + small noise (so self-matches do not dominate); oracle = fp64 brute
    force.  Indices must match wherever margin(k, k+1) > 1e-3 (cosine units); labels wherever the
    class-count margin >= 1 ... and in practice everywhere thanks to the fp32 re-rank."""
    from oracle import knn_ref
    _, emb = b16
    rows, labels = banks
    rng = np.random.default_rng(k)
    q = (rows[rng.permutation(len(rows))[:256]] + 0.8 * rng.standard_normal((256, 768))).astype(np.float32)
    index = v.BankIndex(emb, rows, labels, 3, metric=metric, normalise_rows=(metric == "cosine"))
    idx, score, label, counts = index.search(torch.from_numpy(q).cuda(), k,
                                             normalise_queries=(metric == "cosine"))
    idx, score, label, counts = (t.cpu().numpy() for t in (idx, score, label, counts))
    if metric == "cosine":
        oi, ov, _ = knn_ref.cosine_topk(q, rows, k + 1)
        margin = knn_ref.margins(ov, k)
    else:
        oi, ov, _ = knn_ref.l2_topk_exact(q, rows, k + 1)
        # squared-L2 margin expressed in cosine units: d2 = 2|b|^2(1-cos) for |q|~|b|
        margin = knn_ref.margins(ov, k) / (2 * 768.0)
    ocounts = knn_ref.class_counts(oi[:, :k], labels)
    olabel = knn_ref.vote_pass2(ocounts)
    safe = margin > 1e-3
    assert safe.sum() > 0
    assert np.array_equal(np.sort(idx[safe], 1), np.sort(oi[safe, :k], 1)), "top-k set differs where margin > 1e-3"
    exact_rows = (idx == oi[:, :k]).all(1).mean()
    assert exact_rows > 0.98, f"ordered index lists equal on only {exact_rows:.3f} of rows"
    np.testing.assert_allclose(score, ov[:, :k], rtol=2e-5, atol=2e-5 * (1 if metric == "cosine" else 1500))
    robust = knn_ref.count_margin(ocounts) >= 1
    assert np.array_equal(label[robust & safe], olabel[robust & safe])
    assert (label == olabel).mean() > 0.995
    assert np.array_equal(counts[safe], ocounts[safe])


def test_knn_leave_one_out_known_answers(v, b16, banks, golden_dir):
    """The reference fixtures' known answers through the GPU path: each bank row
    queries the bank; its own row must come first with distance ~0, and the vote over the next k
    reproduces the golden own-label rates."""
    _, emb = b16
    rows, labels = banks
    z = np.load(os.path.join(golden_dir, "banks_loo_topk.npz"))
    index = v.BankIndex(emb, rows, labels, 3, metric="l2")
    idx, score, _, _ = index.search(torch.from_numpy(rows).cuda(), 51)
    idx, score = idx.cpu().numpy(), score.cpu().numpy()
    assert np.array_equal(idx[:, 0], np.arange(len(rows))) and np.all(score[:, 0] < 1e-3)
    man = json.load(open(os.path.join(golden_dir, "MANIFEST.json")))["loo_own_label_rate_l2"]
    for k in (5, 25, 50):
        nb = idx[:, 1:k + 1]
        cnt = np.stack([(labels[nb] == c).sum(1) for c in range(3)], 1)
        rate = (np.argmax(cnt, 1) == labels).mean()
        assert abs(rate - man[str(k)]) < 2e-3, (k, rate)
        gi = z["l2_idx"][:, :k].astype(np.int64)
        d = z["l2_dist"]
        safe = (d[:, k] - d[:, k - 1]) / (2 * 768.0) > 1e-3
        assert np.array_equal(np.sort(nb[safe], 1), np.sort(gi[safe], 1))


def test_chroma_style_collection(v, b16, banks):
    from oracle import knn_ref
    _, emb = b16
    rows, labels = banks
    ids = [f"id{i}" for i in range(len(rows))]
    metas = [{"label": v.LABEL_NAMES[l], "left_prob": float(l == 0), "right_prob": float(l == 1),
              "none_prob": float(l == 2)} for l in labels]
    col = v.BankCollection(emb, rows, ids, metas, space="l2")
    assert col.count() == 1472
    q = rows[[3, 700, 1400]] + 0.01
    res = col.query(query_embeddings=q, n_results=25)
    assert set(res) >= {"ids", "distances", "metadatas"} and len(res["ids"]) == 3 and len(res["ids"][0]) == 25
    assert [r[0] for r in res["ids"]] == ["id3", "id700", "id1400"]
    for d in res["distances"]:
        assert all(d[i] <= d[i + 1] for i in range(24))  # ascending squared L2
    oi, od, _ = knn_ref.l2_topk_exact(q, rows, 25)
    np.testing.assert_allclose(np.array(res["distances"]), od, rtol=1e-4, atol=1e-2)
    # the reference's own call shape: ids / metadatas / distances of one query (chroma.py:211-216)
    r2 = v.determine_class_pass2(res["ids"][0], res["metadatas"][0], res["distances"][0])
    assert r2["dir"] == "left" and abs(r2["probs"].sum() - 1) < 1e-12
    assert v.determine_class_pass1(res["ids"][0], res["metadatas"][0], res["distances"][0])["dir"] in ("left", "pass2")


def test_retriever_top10_cosine(v, b16):
    """Exact cosine top-10 through a cosine-space collection (no metadata filter): the plain
    `sims = Q @ C.T; argpartition; argsort` of ratt_chunk_retriever.py:123,144-148."""
    from oracle import knn_ref
    _, emb = b16
    rng = np.random.default_rng(5)
    cand = rng.standard_normal((5000, 768)).astype(np.float32)
    q = (cand[:40] + 2.0 * rng.standard_normal((40, 768))).astype(np.float32)
    col = v.BankCollection(emb, cand, [f"c{i}" for i in range(5000)], None, space="cosine")
    idx, dist = col.search_device(torch.from_numpy(q).cuda(), 10)
    oi, ov, _ = knn_ref.cosine_topk(q, cand, 11)
    safe = knn_ref.margins(ov, 10) > 1e-5
    assert safe.sum() >= 38
    assert np.array_equal(idx.cpu().numpy()[safe], oi[safe, :10])
    np.testing.assert_allclose(1.0 - dist.cpu().numpy(), ov[:, :10], atol=2e-5)
    want = knn_ref.retriever_restated(q, cand, top_k=10)
    got = knn_ref.normalise(cand)[idx.cpu().numpy()]
    np.testing.assert_allclose(got[safe], want[safe], atol=1e-6)


def test_embed_then_vote_end_to_end(v, b16, banks):
    """C2's shape end to end on a synthetic bank drawn around oracle embeddings: labels from the
    GPU path (bf16 embed -> kNN) equal the oracle's (fp32 embed -> fp64 kNN) wherever the
    oracle's own margin is comfortable."""
    from oracle import knn_ref, vit_ref
    model, emb = b16
    px = vit_ref.synthetic_pixels(48, seed=99)
    want = vit_ref.hf_embed(model, px, normalise=True)
    rng = np.random.default_rng(7)
    centres = want[rng.permutation(48)[:3]]
    bank = np.concatenate([knn_ref.normalise(c + 0.02 * rng.standard_normal((n, 768)).astype(np.float32))
                           for c, n in zip(centres, (633, 749, 90))])
    labels = np.concatenate([np.full(n, i, np.int32) for i, n in enumerate((633, 749, 90))])
    index = v.BankIndex(emb, bank, labels, 3, metric="cosine", normalise_rows=True)
    got = emb.embed(px.cuda(), l2_normalise=True)
    idx, score, label, counts = index.search(got, 25, normalise_queries=True)
    oi, ov, _ = knn_ref.cosine_topk(want, bank, 26)
    ocounts = knn_ref.class_counts(oi[:, :25], labels)
    olabel = knn_ref.vote_pass2(ocounts)
    robust = knn_ref.count_margin(ocounts) >= 5
    assert robust.sum() >= 24
    assert np.array_equal(label.cpu().numpy()[robust], olabel[robust])
    assert _cos(got.cpu().numpy(), want).min() >= COS_BAR


def test_full_size_properties(v, banks):
    """BASELINE.json C2 size (256 frames, ViT-B/16 x 12 layers): properties that need no oracle —
    duplicate frames give bit-identical rows regardless of batch position, rows are unit norm,
    kNN of a bank row returns itself, results are deterministic across calls."""
    from oracle import vit_ref
    model = vit_ref.build_hf_model("vit_b16", 0)
    emb = v.ViTEmbedder.from_hf(model, device=0, max_batch=256)
    base = vit_ref.synthetic_pixels(32, seed=11).cuda().bfloat16()
    px = base.repeat(8, 1, 1, 1)  # 256 frames: frame i == frame i+32
    a = emb.embed(px)
    b = emb.embed(px)
    assert torch.equal(a, b)
    assert torch.equal(a[:32], a[224:256]) and torch.equal(a[:32], a[96:128])
    assert torch.allclose(a.norm(dim=1), torch.ones(256, device="cuda"), atol=1e-5)
    assert torch.isfinite(a).all()
    rows, labels = banks
    index = v.BankIndex(emb, rows, labels, 3, metric="l2")
    idx, score, label, counts = index.search(torch.from_numpy(rows[:256]).cuda(), 25)
    assert torch.equal(idx[:, 0].cpu(), torch.arange(256, dtype=torch.int32))
    assert (counts.sum(1) == 25).all()
    emb.close()


# ---------------------------------------------------------------------------------------------
# code-path variants must agree (they are selected by environment variables read once per process)
# ---------------------------------------------------------------------------------------------
def _fingerprint(env_extra):
    import json
    import subprocess
    import sys
    env = dict(os.environ)
    env.update(env_extra)
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, os.path.join(root, "tools", "embed_fingerprint.py")], env=env,
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    return np.array(json.loads(r.stdout.strip().splitlines()[-1])["emb"])


def test_code_path_variants_agree():
    base = _fingerprint({})
    assert base.shape == (11, 768) and np.isfinite(base).all()
    # last layer on all tokens vs CLS rows only: the same arithmetic for the rows that are read
    full = _fingerprint({"VITK_FULL_LAST_LAYER": "1"})
    assert np.array_equal(base, full)
    # single-CTA tcgen05.mma (cta_group::1) vs the CTA pair: same K order, same accumulators
    cg1 = _fingerprint({"VITK_GEMM_CG": "1"})
    assert np.array_equal(base, cg1)
    # CTA-pair GEMMs with five operand stages / two staging tiles vs six / one: pipeline depth only
    s5 = _fingerprint({"VITK_GEMM_S6": "0"})
    assert np.array_equal(base, s5)
    # every kernel walking its rows upwards vs the "snake" order (each kernel starts on the rows its
    # producer wrote last): tile order only
    up = _fingerprint({"VITK_SNAKE": "0"})
    assert np.array_equal(base, up)
    # attention v2 (exact row max, two passes) vs v3 (shift by the first-32 max): the bf16 P tiles
    # are rounded at a different scale, so the two agree to bf16 noise (either one is 0.99996 from
    # the fp32 oracle; measured 0.999988 between them on B200), not bit for bit
    att2 = _fingerprint({"VITK_ATTN": "2"})
    cos = (base * att2).sum(1) / (np.linalg.norm(base, axis=1) * np.linalg.norm(att2, axis=1))
    assert cos.min() > 0.9999, cos.min()
    # attention v5 (softmax split over a loader and a math warp per scheduler, S through a shared-
    # memory ring): v3's arithmetic in another schedule -- bit for bit (slower: profiles/r02_attn_v5_timeline.txt)
    # (v3's default for 197 keys takes the 32-key chunks in the order 3,4,5 | 0,1,2,tail so that O = P V can
    # start inside the pass: another shift estimate and summation order.  VITK_ATTN_EARLY_PV=0 is the
    # in-order pass, whose arithmetic v5, v6 and the single-issuer form share.)
    inorder = _fingerprint({"VITK_ATTN_EARLY_PV": "0"})
    cos = (base * inorder).sum(1) / (np.linalg.norm(base, axis=1) * np.linalg.norm(inorder, axis=1))
    assert cos.min() > 0.9999, cos.min()
    att5 = _fingerprint({"VITK_ATTN": "5"})
    assert np.array_equal(inorder, att5)
    # attention v6 (P in shared memory, SS-form P.V in two halves): again v3's arithmetic, bit for bit
    # (slower too: profiles/r02_attn_v6_timeline.txt); and v3 with its round-1 single MMA-issuing thread
    assert np.array_equal(inorder, _fingerprint({"VITK_ATTN": "6"}))
    assert np.array_equal(inorder, _fingerprint({"VITK_ATTN_ISSUER": "1"}))
    # the patch GEMM writing the bf16 stream + row sums itself vs fp32 scratch + a cast / statistics
    # pass (one rounding of the same fp32 value either way; the row sums are grouped differently)
    unfused = _fingerprint({"VITK_FUSED_PATCH": "0"})
    cos = (base * unfused).sum(1) / (np.linalg.norm(base, axis=1) * np.linalg.norm(unfused, axis=1))
    assert cos.min() > 0.9999, cos.min()        # measured 0.999989
    # last layer: Q projected for the CLS rows only vs for every token: same arithmetic for the rows read
    assert np.array_equal(base, _fingerprint({"VITK_SPLIT_LAST_QKV": "0"}))
    # programmatic dependent launch of the GEMM / attention chain (default) vs plain stream order:
    # the same kernels on the same data -- bit for bit, or the early launch read something too soon
    assert np.array_equal(base, _fingerprint({"VITK_PDL": "0"}))
    # residual stream kept in bf16 in place (default) vs the fp32 master copy + bf16 copy of
    # round 1: one bf16 rounding per residual add separates them
    rf32 = _fingerprint({"VITK_RESID_F32": "1"})
    cos = (base * rf32).sum(1) / (np.linalg.norm(base, axis=1) * np.linalg.norm(rf32, axis=1))
    assert cos.min() > 0.9999, cos.min()
    assert not np.array_equal(base, rf32)      # the switch really selects another code path


def test_empty_and_ragged_inputs(v, b16, banks):
    """Empty batches return empty results (no launch, no error); a 1-frame batch and a batch one
    past a multiple of max_batch give the same rows as the big batch."""
    from oracle import vit_ref
    _, emb = b16
    rows, labels = banks
    assert emb.embed_frames(np.zeros((0, 224, 224, 3), np.uint8)).shape == (0, 768)
    assert emb.embed(torch.zeros((0, 3, 224, 224), device="cuda")).shape == (0, 768)
    index = v.BankIndex(emb, rows, labels, 3, metric="l2")
    idx, score, label, counts = index.search(torch.zeros((0, 768), device="cuda"), 5)
    assert idx.shape == (0, 5) and label.shape == (0,) and counts.shape == (0, 3)
    assert v.viterbi_from_probs(emb, np.zeros((0, 3))).shape == (0,)
    st = v.viterbi_from_counts(emb, torch.zeros((0, 3), dtype=torch.int32, device="cuda"), 25)
    assert st.shape == (0,)
    frames = vit_ref.synthetic_frames_u8(65, seed=9)           # max_batch is 64: ragged tail of 1
    big = emb.embed_frames(frames)
    assert np.array_equal(emb.embed_frames(frames[64:65]), big[64:65])
    assert np.array_equal(emb.embed_frames(frames[:1]), big[:1])
    # k larger than a tiny bank: missing neighbours come back as -1 like an empty Chroma slot
    tiny = v.BankIndex(emb, rows[:3], labels[:3], 3, metric="cosine", normalise_rows=True)
    idx, score, label, counts = tiny.search(torch.from_numpy(rows[:2]).cuda(), 5, normalise_queries=True)
    assert (idx[:, :3] >= 0).all() and (idx[:, 3:] == -1).all() and (counts.sum(1) == 3).all()


def test_in_step_fc1_timing_hook(v, b16):
    """vitk_profile_fc1: one CUDA-event pair per fc1 launch of the forward (11 of the 12 layers run
    the full-width MLP, the last one only its CLS rows), a positive summed duration, list cleared
    by the read; results are unchanged by the instrumentation."""
    import ctypes
    _, emb = b16
    lib = v.load()
    px = (torch.randn(8, 3, 224, 224, generator=torch.Generator().manual_seed(3)) * 0.5).clamp_(-1, 1).cuda()
    want = emb.embed(px).clone()
    assert lib.vitk_profile_fc1(emb.handle, 1) == 0
    got = emb.embed(px)
    got2 = emb.embed(px)
    tot, cnt = ctypes.c_double(0.0), ctypes.c_int(0)
    assert lib.vitk_profile_fc1(emb.handle, 0) == 0
    assert lib.vitk_profile_fc1_read(emb.handle, ctypes.byref(tot), ctypes.byref(cnt)) == 0
    n_full = emb.geometry.layers - 1
    assert cnt.value == 2 * n_full and tot.value > 0.0
    assert torch.equal(got, want) and torch.equal(got2, want)
    assert lib.vitk_profile_fc1_read(emb.handle, ctypes.byref(tot), ctypes.byref(cnt)) == 0
    assert cnt.value == 0
    emb.embed(px)
    assert lib.vitk_profile_fc1_read(emb.handle, ctypes.byref(tot), ctypes.byref(cnt)) == 0 and cnt.value == 0
