"""GPU parity of the preprocessing front-end through the C ABI: the CUDA
resize kernel is BIT-EXACT against the restated ATen/Pillow fixed-point algorithm (oracle, pinned
on the reference's ViTImageProcessor) and against the golden images produced by that processor;
frames of any size then embed like the reference's hf_vit_embed_batch."""
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
def model_emb(v):
    from oracle import vit_ref
    model = vit_ref.build_hf_model("vit_b16_l2", 0, randomise_ln=True)
    e = v.ViTEmbedder.from_hf(model, device=0, max_batch=8)
    yield model, e
    e.close()


def test_resize_matches_golden_images_of_the_reference_processor(v, model_emb, golden_dir):
    from oracle import preprocess_ref
    _, emb = model_emb
    z = np.load(os.path.join(golden_dir, "preprocess_golden.npz"))
    for c in [str(x) for x in z["cases"]]:
        f = preprocess_ref.synthetic_frames(c)
        got = emb.resize_frames(torch.from_numpy(f).cuda()).cpu().numpy()
        assert np.array_equal(got, z[c + "_out"]), c


@pytest.mark.parametrize("hw", [(432, 768), (720, 1280), (1080, 1920), (224, 224), (224, 500), (500, 224),
                                (17, 23), (225, 223), (2160, 3840)])
def test_resize_is_bit_exact_against_oracle(v, model_emb, hw):
    from oracle import preprocess_ref
    _, emb = model_emb
    n = 3 if hw[0] * hw[1] < 1_000_000 else 1
    f = np.random.default_rng(hw[1]).integers(0, 256, (n, hw[0], hw[1], 3), dtype=np.uint8)
    got = emb.resize_frames(torch.from_numpy(f).cuda()).cpu().numpy()
    assert np.array_equal(got, preprocess_ref.resize_u8(f))


def test_frames_of_any_size_embed_like_the_reference(v, model_emb):
    """hf_vit_embed_batch on 432x768 frames (the reference's TF-path frame size): processor
    (oracle) -> HF ViTModel fp32 vs GPU resize -> bf16 forward; cosine >= 0.999 per frame."""
    from oracle import preprocess_ref, vit_ref
    model, emb = model_emb
    f = np.random.default_rng(5).integers(0, 256, (11, 432, 768, 3), dtype=np.uint8)
    want = vit_ref.hf_embed(model, torch.from_numpy(preprocess_ref.pixel_values(f)), normalise=True)
    v.install(emb)
    got = v.hf_vit_embed_batch(f)           # 11 frames > max_batch 8: chunked inside
    assert got.shape == (11, 768) and got.dtype == np.float32
    cos = (got * want).sum(1) / (np.linalg.norm(got, axis=1) * np.linalg.norm(want, axis=1))
    assert cos.min() >= 0.999, cos
    # identical to resizing explicitly and embedding the 224x224 result
    small = emb.resize_frames(torch.from_numpy(f).cuda()).cpu().numpy()
    assert np.array_equal(emb.embed_frames(small), got)


def test_build_frame_store_from_image_files(v, model_emb, tmp_path):
    """build_frame_store (ratt_write_clips_to_ragdb.py:182-233) end to end: PNG files -> PIL loader
    -> embed -> memmap store the reference's reader opens."""
    from PIL import Image
    _, emb = model_emb
    rng = np.random.default_rng(2)
    frames = rng.integers(0, 256, (19, 224, 224, 3), dtype=np.uint8)
    paths = []
    for i, fr in enumerate(frames):
        p = str(tmp_path / f"frame_{i:04d}.png")
        Image.fromarray(fr).save(p)
        paths.append(p)
    v.store.build_frame_store(paths, batch_size=8, num_workers=4, store_name="t", embedder=emb,
                              cache_dir=str(tmp_path / "cache"))
    mm, fp, idx = v.store.load_frame_store("t", cache_dir=str(tmp_path / "cache"))
    assert mm.shape == (19, 768) and fp == paths and idx[paths[3]] == 3
    assert np.array_equal(np.asarray(mm), emb.embed_frames(frames))


@pytest.mark.parametrize("hw", [(720, 1280), (1080, 1920), (864, 1536), (433, 769), (480, 854), (2160, 3840),
                                (224, 224), (448, 672), (1000, 225)])
def test_inter_area_resize_is_bit_exact_against_restatement(v, model_emb, hw):
    """cv2.resize(frame, (w, h), interpolation=cv2.INTER_AREA) of the TF path (generate_clips_hmm.py:392):
    the CUDA kernel against the restated OpenCV algorithm -- itself pinned byte for byte on the live
    cv2 (tests/test_oracle_cpu.py) -- bit for bit (general tables, the float box form, the integer
    2 x 2 form and identity); growing an image is refused."""
    from oracle import preprocess_ref
    _, emb = model_emb            # 224 x 224 model: any larger frame shrinks to it
    n = 2 if hw[0] * hw[1] < 1_000_000 else 1
    f = np.random.default_rng(hw[0]).integers(0, 256, (n, hw[0], hw[1], 3), dtype=np.uint8)
    got = emb.resize_frames(torch.from_numpy(f).cuda(), interpolation="area").cpu().numpy()
    want = np.stack([preprocess_ref.resize_area_u8(x, 224, 224) for x in f])
    assert np.array_equal(got, want)


def test_inter_area_to_the_tf_frame_size_and_errors(v):
    from oracle import preprocess_ref, vit_ref
    geom = v.ViTGeometry.tf_432x768()
    import dataclasses
    geom = dataclasses.replace(geom, layers=1)
    emb = v.ViTEmbedder(geom, v.random_init_state_dict(geom, 0), device=0, max_batch=4)
    try:
        f = np.random.default_rng(1).integers(0, 256, (2, 720, 1280, 3), dtype=np.uint8)
        got = emb.resize_frames(torch.from_numpy(f).cuda(), interpolation="area").cpu().numpy()
        assert got.shape == (2, 432, 768, 3)
        assert np.array_equal(got, np.stack([preprocess_ref.resize_area_u8(x, 432, 768) for x in f]))
        small = torch.zeros((1, 200, 300, 3), dtype=torch.uint8, device="cuda")
        with pytest.raises(v.VitkError):
            emb.resize_frames(small, interpolation="area")
        with pytest.raises(v.VitkError):
            emb.resize_frames(torch.from_numpy(f).cuda(), interpolation="lanczos")
    finally:
        emb.close()


def _tf_embedder(v, layers=1):
    import dataclasses
    geom = dataclasses.replace(v.ViTGeometry.tf_432x768(), layers=layers)
    return v.ViTEmbedder(geom, v.random_init_state_dict(geom, 0), device=0, max_batch=4)


def test_cv2_resizes_match_bytes_produced_by_opencv(v, model_emb, golden_dir):
    """INTER_AREA and INTER_LINEAR kernels against the committed digests of what cv2.resize itself
    produced (tools/make_golden_preprocess.py, OpenCV 4.13) on seeded frames: 1080p / 720p, exact
    2x (the integer fast path), 3x, integer in one axis, odd sizes; to 432x768 and to 224x224."""
    import hashlib
    from oracle import preprocess_ref as pr
    z = np.load(os.path.join(golden_dir, "preprocess_cv2_golden.npz"))
    _, emb224 = model_emb
    tf = _tf_embedder(v)
    try:
        for name in [str(c) for c in z["cases"]]:
            (h, w), (oh, ow) = pr.CV2_GOLDEN_CASES[name]
            e = tf if (oh, ow) == (432, 768) else emb224
            f = torch.from_numpy(pr.cv2_case_frame(name)[None]).cuda()
            for tag in ("area", "linear"):
                got = e.resize_frames(f, interpolation=tag).cpu().numpy()[0]
                assert got.shape == (oh, ow, 3)
                assert np.array_equal(got[:24, :24], z[f"{name}/{tag}/crop0"]), (name, tag)
                assert np.array_equal(got[-24:, -24:], z[f"{name}/{tag}/crop1"]), (name, tag)
                assert int(got.astype(np.int64).sum()) == int(z[f"{name}/{tag}/sum"]), (name, tag)
                assert hashlib.sha256(got.tobytes()).hexdigest() == str(z[f"{name}/{tag}/sha256"]), (name, tag)
    finally:
        tf.close()


@pytest.mark.parametrize("hw", [(1080, 1920), (720, 1280), (864, 1536), (433, 769), (300, 500), (431, 767),
                                (216, 384), (100, 1000), (432, 768), (2, 2)])
def test_inter_linear_resize_is_bit_exact(v, hw):
    """cv2.resize(im, (768, 432)) -- default INTER_LINEAR, extract_embeddings' resize
    (smarter_generate_clips.py:81): the CUDA kernel against the restatement (pinned on live cv2),
    shrinking and growing, and against cv2 itself when it imports on this box."""
    from oracle import preprocess_ref as pr
    tf = _tf_embedder(v)
    try:
        f = np.random.default_rng(hw[0] + hw[1]).integers(0, 256, (2, hw[0], hw[1], 3), dtype=np.uint8)
        got = tf.resize_frames(torch.from_numpy(f).cuda(), interpolation="linear").cpu().numpy()
        want = np.stack([pr.resize_linear_u8(x, 432, 768) for x in f])
        assert np.array_equal(got, want)
        try:
            import cv2
        except ImportError:
            return
        assert np.array_equal(got[0], cv2.resize(f[0], (768, 432)))
        if hw[0] >= 432 and hw[1] >= 768:
            ga = tf.resize_frames(torch.from_numpy(f).cuda(), interpolation="area").cpu().numpy()
            assert np.array_equal(ga[1], cv2.resize(f[1], (768, 432), interpolation=cv2.INTER_AREA))
    finally:
        tf.close()


def test_build_frame_store_from_720p_image_files(v, model_emb, tmp_path):
    """The reference's build_frame_store feeds frames at the VIDEO's resolution
    (ratt_write_clips_to_ragdb.py:197-206); they go through the streaming path (upload raw ->
    GPU resize -> forward, overlapped) and must equal hf_vit_embed_batch on the same frames and
    the HF oracle (processor + ViTModel fp32)."""
    from PIL import Image
    from oracle import preprocess_ref, vit_ref
    model, emb = model_emb
    rng = np.random.default_rng(12)
    frames = rng.integers(0, 256, (19, 720, 1280, 3), dtype=np.uint8)
    paths = []
    for i, fr in enumerate(frames):
        p = str(tmp_path / f"frame_{i:04d}.png")
        Image.fromarray(fr).save(p)
        paths.append(p)
    v.store.build_frame_store(paths, batch_size=8, num_workers=4, store_name="hd", embedder=emb,
                              cache_dir=str(tmp_path / "cache"))
    mm, fp, _ = v.store.load_frame_store("hd", cache_dir=str(tmp_path / "cache"))
    got = np.asarray(mm)
    assert got.shape == (19, 768) and fp == paths
    v.install(emb)
    assert np.array_equal(got, v.hf_vit_embed_batch(frames))      # 19 > max_batch 8: same streaming path
    small = emb.resize_frames(torch.from_numpy(frames).cuda()).cpu().numpy()
    assert np.array_equal(got, emb.embed_frames(small))            # == explicit resize + model-size frames
    want = vit_ref.hf_embed(model, torch.from_numpy(preprocess_ref.pixel_values(frames)), normalise=True)
    cos = (got * want).sum(1) / (np.linalg.norm(got, axis=1) * np.linalg.norm(want, axis=1))
    assert cos.min() >= 0.999, cos
    # label_frames_sharded (single process) takes raw-size frames too
    rows = rng.standard_normal((64, 768)).astype(np.float32)
    index = v.BankIndex(emb, rows, (np.arange(64) % 3).astype(np.int32), 3, metric="l2")
    e2, lab = v.label_frames_sharded(emb, index, frames, k=5)
    assert np.array_equal(e2.cpu().numpy(), got) and lab.shape == (19,)


def test_extract_embeddings_uses_cv2_default_interpolation(v, tmp_path):
    """smarter_generate_clips.py:69-94: imread + BGR2RGB + cv2.resize(im, (768, 432)) [INTER_LINEAR]
    + predict in batches + vstack.  Default loader + GPU INTER_LINEAR == host cv2.resize + predict,
    bit for bit; mixed sizes fall back to cv2 per frame."""
    cv2 = pytest.importorskip("cv2")
    tf = _tf_embedder(v, layers=2)
    try:
        rng = np.random.default_rng(3)
        paths = []
        for i in range(7):
            fr = rng.integers(0, 256, (720, 1280, 3), dtype=np.uint8)
            p = str(tmp_path / f"f{i}.png")
            cv2.imwrite(p, fr)
            paths.append(p)
        got = v.extract_embeddings(tf, paths, batch_size=4)
        assert got.shape == (7, 768) and got.dtype == np.float32
        host = np.stack([cv2.resize(cv2.cvtColor(cv2.imread(p), cv2.COLOR_BGR2RGB), (768, 432)) for p in paths])
        want = tf.predict(host, batch_size=4, verbose=0)["pre_logits"].reshape(7, 768)
        assert np.array_equal(got, want)
        # a folder with two frame sizes
        cv2.imwrite(str(tmp_path / "odd.png"), rng.integers(0, 256, (500, 900, 3), dtype=np.uint8))
        mixed = v.extract_embeddings(tf, paths[:2] + [str(tmp_path / "odd.png")], batch_size=1024)
        assert np.array_equal(mixed[:2], got[:2]) and mixed.shape == (3, 768)
    finally:
        tf.close()
