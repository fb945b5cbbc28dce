"""GPU parity at the sizes and depths that are BENCHMARKED (VERDICT r1 "no oracle parity at the
sizes you benchmark"): the 256-frame bench shape of ViT-B/16, the full 24 layers of ViT-L/16, the
TF bank-writing geometry (patch 256 -> 4 tokens, raw pixels, CLS zeros), plus the failure modes the
advisor listed (outlier channels through the folded LayerNorm, calls alternating between the
caller's stream and the library's own, weight files).  Oracle = the reference's own dependency
(HF ViTModel fp32) through committed golden vectors (tools/make_golden.py), so the GPU box needs no
long CPU run.  Bar: cosine >= 0.999 per frame (north_star); the measured minima are printed.
"""
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


def _golden(golden_dir, name, model):
    from oracle import vit_ref
    z = np.load(os.path.join(golden_dir, name))
    if abs(vit_ref.weights_fingerprint(model) - float(z["fingerprint"])) > 1e-6 * float(z["fingerprint"]):
        pytest.skip("torch RNG stream differs from the golden generator's")
    return z


def test_bench_shape_256_frames_matches_oracle_and_small_batch(v, golden_dir):
    """ViT-B/16, 12 layers, max_batch = 256 (M = 50 432 rows, 7.99 waves of CTA pairs, snake order
    over > 77 MB activations): 8 golden frames planted at rows {0,1,63,127,128,200,254,255} of one
    256-frame batch are within the bar of the HF oracle AND bit-equal to the same frames run at
    max_batch = 64."""
    from oracle import vit_ref
    model = vit_ref.build_hf_model("vit_b16", 0, randomise_ln=True)
    z = _golden(golden_dir, "vit_b16_seed0.npz", model)
    gold = vit_ref.synthetic_pixels(8, seed=1234)
    rows = [0, 1, 63, 127, 128, 200, 254, 255]
    px = vit_ref.synthetic_pixels(256, seed=77)
    px[rows] = gold
    big = v.ViTEmbedder.from_hf(model, device=0, max_batch=256)
    small = v.ViTEmbedder.from_hf(model, device=0, max_batch=64)
    try:
        out = big.embed(px.cuda(), l2_normalise=False).cpu().numpy()
        c = _cos(out[rows], z["cls"])
        print(f"\nViT-B/16 x12 @ max_batch 256: min cosine vs HF fp32 = {c.min():.6f}")
        assert c.min() >= COS_BAR, c
        ref64 = small.embed(gold.cuda(), l2_normalise=False).cpu().numpy()
        assert np.array_equal(out[rows], ref64)          # batch size / tile order do not change a row
        # the whole batch through the 64-frame handle (4 chunks) is bit-identical too
        assert np.array_equal(small.embed(px.cuda(), l2_normalise=False).cpu().numpy(), out)
        # bf16 device-resident input, the bench's format
        outb = big.embed(px.cuda().bfloat16(), l2_normalise=True).cpu().numpy()
        assert _cos(outb[rows], z["cls"]).min() >= COS_BAR
    finally:
        big.close()
        small.close()


def test_vit_l16_all_24_layers_match_oracle(v, golden_dir):
    """BASELINE configs[3]: ViT-L/16 at its full depth (24 layers of bf16 GEMMs + a bf16 residual
    stream are where the margin goes), 4 golden frames vs HF fp32."""
    from oracle import vit_ref
    model = vit_ref.build_hf_model("vit_l16", 1, randomise_ln=True)
    z = _golden(golden_dir, "vit_l16_seed1.npz", model)
    emb = v.ViTEmbedder.from_hf(model, device=0, max_batch=8)
    try:
        got = emb.embed(vit_ref.synthetic_pixels(4, seed=1234).cuda(), l2_normalise=False).cpu().numpy()
        c = _cos(got, z["cls"])
        print(f"\nViT-L/16 x24: min cosine vs HF fp32 = {c.min():.6f}")
        assert c.min() >= COS_BAR, c
        assert np.abs(got - z["cls"]).mean() < 0.05
    finally:
        emb.close()


def test_tf_bank_writing_geometry_patch256_four_tokens(v, golden_dir):
    """write_embeddings.py:101-109 / chroma.py:148-156: 432x768 RAW uint8 frames, patch 256 'valid'
    (1 x 3 patches + CLS = 4 tokens, K = 196 608 in the patch GEMM), eps 1e-6, tanh-GELU, CLS token
    initialised to zeros (models/vit_backbone.py:93-104) -- the whole 12-layer model against HF
    configured identically, through predict(...)['pre_logits'] like the script calls it."""
    from oracle import vit_ref
    model = vit_ref.build_hf_model("vit_tf432_p256", 2, randomise_ln=True, cls_zeros=True)
    z = _golden(golden_dir, "vit_tf432_p256_seed2.npz", model)
    geom = v.ViTGeometry.tf_432x768(patch=256)
    assert geom.tokens == 4
    emb = v.ViTEmbedder(geom, model.state_dict(), device=0, max_batch=4)
    try:
        u8 = vit_ref.synthetic_frames_u8(4, seed=4321, size=(432, 768))
        out = emb.predict(u8, batch_size=1024, verbose=1)["pre_logits"]
        assert out.shape == (4, 1, 1, 768) and out.dtype == np.float32
        c = _cos(out.reshape(4, 768), z["cls_u8"])
        print(f"\nTF geometry patch 256 (4 tokens): min cosine vs HF fp32 = {c.min():.6f}")
        assert c.min() >= COS_BAR, c
        # random_init_state_dict(cls_zeros=True) builds that state without transformers
        sd = v.random_init_state_dict(geom, 3, cls_zeros=True)
        assert float(sd["embeddings.cls_token"].abs().sum()) == 0.0
    finally:
        emb.close()


def test_layernorm_fold_survives_outlier_channels_and_large_means(v):
    """The folded LayerNorm consumes the RAW residual row in bf16 and a variance from E[x^2]-mean^2
    (advisor): real checkpoints have outlier channels and rows whose |mean| / std is large.  Two
    stress models -- three position-embedding channels x500 (outlier channels ~40 sigma), and a
    constant +1 on every channel of the embedded tokens (|mean|/std ~ 4) -- stay within the bar of
    HF fp32.  (A bf16 residual stream, like any bf16 inference path, resolves a row only to
    2^-9 of its magnitude: offsets far beyond the row's spread are outside every bf16 model.)"""
    from oracle import vit_ref
    for tag in ("outlier_channels", "row_offset"):
        model = vit_ref.build_hf_model("vit_b16_l2", 5, randomise_ln=True)
        with torch.no_grad():
            if tag == "outlier_channels":
                model.embeddings.position_embeddings[:, :, [7, 300, 511]] *= 500.0
                model.encoder.layer[0].output.dense.bias[[7, 300]] += 3.0
            else:
                model.embeddings.position_embeddings += 1.0
        emb = v.ViTEmbedder.from_hf(model, device=0, max_batch=8)
        try:
            px = vit_ref.synthetic_pixels(6, seed=21)
            want = vit_ref.hf_embed(model, px, normalise=False)
            got = emb.embed(px.cuda(), l2_normalise=False).cpu().numpy()
            c = _cos(got, want)
            print(f"\nLN fold, {tag}: min cosine {c.min():.6f}")
            assert c.min() >= COS_BAR, (tag, c)
        finally:
            emb.close()


def test_calls_alternating_between_user_and_library_streams_are_ordered(v):
    """vitk_embed runs on the caller's stream, the host-buffer calls on the library's own stream,
    over ONE workspace (advisor): an asynchronous embed() followed at once by embed_frames(), and
    the reverse, must give the rows each gives alone."""
    from oracle import vit_ref
    model = vit_ref.build_hf_model("vit_b16_l2", 0, randomise_ln=True)
    emb = v.ViTEmbedder.from_hf(model, device=0, max_batch=32)
    try:
        px = vit_ref.synthetic_pixels(32, seed=1).cuda()
        frames = vit_ref.synthetic_frames_u8(32, seed=2)
        alone_a = emb.embed(px).clone()
        torch.cuda.synchronize()
        alone_b = emb.embed_frames(frames)
        side = torch.cuda.Stream()
        for _ in range(6):
            with torch.cuda.stream(side):
                a = emb.embed(px, stream=side)           # asynchronous on a non-default stream
            b = emb.embed_frames(frames)                 # library stream, same workspace
            with torch.cuda.stream(side):
                a2 = emb.embed(px, stream=side)
            torch.cuda.synchronize()
            assert torch.equal(a, alone_a) and torch.equal(a2, alone_a)
            assert np.array_equal(b, alone_b)
    finally:
        emb.close()


def test_save_and_load_weights_round_trip(v, tmp_path):
    """model.save_weights(path) / model.load_weights(path) (write_embeddings.py:243, chroma.py:159):
    .npz keyed by HF names; loading other weights changes the output, loading the saved ones back
    restores it bit for bit; HDF5 paths are refused with a reason."""
    from oracle import vit_ref
    geom = v.ViTGeometry(layers=2)
    a = v.ViTEmbedder(geom, v.random_init_state_dict(geom, 0), device=0, max_batch=8)
    try:
        frames = vit_ref.synthetic_frames_u8(5, seed=3)
        before = a.embed_frames(frames)
        path = a.save_weights(str(tmp_path / "vit_random_weights"))
        assert path.endswith(".npz") and os.path.exists(path)
        a.load_weights(v.random_init_state_dict(geom, 1))
        assert not np.array_equal(a.embed_frames(frames), before)
        a.load_weights(path)
        assert np.array_equal(a.embed_frames(frames), before)
        with pytest.raises(v.VitkError):
            a.load_weights(str(tmp_path / "vit_random_weights.h5"))
        with pytest.raises(v.VitkError):
            a.save_weights(str(tmp_path / "x.h5"))
        with pytest.raises(v.VitkError):
            a.load_weights({"layernorm.weight": np.ones(768, np.float32)})
    finally:
        a.close()
