"""On-disk formats either side of the path: what we write must be read back
unchanged by the reference's readers, restated here line by line from
ratt_write_clips_to_ragdb.py:236-254, train/train_chunk_encoder.py:36-58 and clustering.py:26-36."""
import os

import numpy as np
import pytest


def _reference_load_frame_store(cache_dir, store_name):
    # ratt_write_clips_to_ragdb.py:236-254
    meta = np.load(os.path.join(cache_dir, f"{store_name}_meta.npz"))
    n_frames, emb_dim = int(meta["n_frames"]), int(meta["emb_dim"])
    emb_mm = np.memmap(os.path.join(cache_dir, f"{store_name}_emb.dat"), dtype="float32", mode="r",
                       shape=(n_frames, emb_dim))
    frame_paths = np.load(os.path.join(cache_dir, f"{store_name}_paths.npy"), allow_pickle=True).tolist()
    return emb_mm, frame_paths, {p: i for i, p in enumerate(frame_paths)}


def test_frame_store_round_trip(tmp_path):
    import vit_research_b200 as v
    rng = np.random.default_rng(0)
    emb = rng.standard_normal((37, 768)).astype(np.float32)
    paths = [f"/data/game1/frame_{i:06d}.jpg" for i in range(37)]
    p = v.store.write_frame_store("unit", emb, paths, cache_dir=str(tmp_path))
    assert os.path.getsize(p["emb"]) == 37 * 768 * 4
    assert v.store.frame_store_exists("unit", cache_dir=str(tmp_path))
    mm, fp, idx = _reference_load_frame_store(str(tmp_path), "unit")
    assert mm.dtype == np.float32 and mm.shape == (37, 768) and np.array_equal(np.asarray(mm), emb)
    assert fp == paths and idx[paths[5]] == 5
    mm2, fp2, idx2 = v.store.load_frame_store("unit", cache_dir=str(tmp_path))
    assert np.array_equal(np.asarray(mm2), emb) and fp2 == paths and idx2 == idx
    # torch tensors are accepted as well (device buffers are copied out once)
    import torch
    v.store.write_frame_store("unit_t", torch.from_numpy(emb), paths, cache_dir=str(tmp_path))
    assert np.array_equal(np.asarray(_reference_load_frame_store(str(tmp_path), "unit_t")[0]), emb)


def test_frame_store_rejects_mismatched_paths(tmp_path):
    import vit_research_b200 as v
    with pytest.raises(v.VitkError):
        v.store.write_frame_store("bad", np.zeros((3, 768), np.float32), ["a", "b"], cache_dir=str(tmp_path))


def test_label_banks_round_trip_through_the_reference_reader(tmp_path, golden_dir):
    import vit_research_b200 as v
    rows, labels, ids = v.load_label_banks(os.path.join(golden_dir, "banks"))
    frame_ids = [i.split(":", 1)[1] for i in ids]
    v.store.write_label_banks(str(tmp_path), rows, labels, frame_ids)
    for name, n in (("left", 633), ("right", 749), ("none", 90)):
        # clustering.py:26-36
        data = np.load(os.path.join(str(tmp_path), f"{name}_embeddings.npz"))
        ref = np.load(os.path.join(golden_dir, "banks", f"{name}_embeddings.npz"))
        assert data["embeddings"].shape == (n, 1, 768) and data["embeddings"].dtype == np.float32
        t = data["embeddings"].reshape(len(data["embeddings"]), 768)
        assert np.array_equal(t, ref["embeddings"].reshape(n, 768))
        assert data["frame_ids"].dtype.kind == "U"
        assert [str(x) for x in data["frame_ids"]] == [str(x) for x in ref["frame_ids"]]
        assert np.mean(data["embeddings"], axis=0)[0].shape == (768,)
    rows2, labels2, ids2 = v.load_label_banks(str(tmp_path))
    assert np.array_equal(rows2, rows) and np.array_equal(labels2, labels) and ids2 == ids
