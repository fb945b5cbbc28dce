"""CPU tests: the C-ABI library builds, loads and exports every symbol include/vitk.h declares;
host-side logic; loud failure without a GPU (no CPU fallback)."""
import ctypes as C
import os
import re

import numpy as np
import pytest
import torch

import vit_research_b200 as v
from vit_research_b200 import _lib


def header_symbols():
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    src = open(os.path.join(root, "include", "vitk.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(vitk_[a-z0-9_]+)\s*\(", src)))


def test_library_builds_and_exports_every_declared_symbol():
    lib = v.load()
    syms = header_symbols()
    assert len(syms) >= 15
    for s in syms:
        assert hasattr(lib, s), f"{s} declared in include/vitk.h but not exported"
        assert s in _lib.SIGNATURES, f"{s} has no ctypes prototype"
    assert sorted(_lib.SIGNATURES) == syms
    assert b"sm_100a" in lib.vitk_version()


def test_sass_contains_blackwell_tensor_and_tma_instructions():
    import shutil, subprocess
    if not shutil.which("cuobjdump"):
        pytest.skip("cuobjdump not on PATH")
    sass = subprocess.run(["cuobjdump", "-sass", str(_lib.LIB_PATH)], capture_output=True, text=True).stdout
    for mnemonic in ("UTCHMMA", "UTMALDG", "UTMASTG", "LDTM"):
        assert mnemonic in sass, f"{mnemonic} missing: not a tcgen05/TMA kernel"
    assert "HMMA." not in sass.replace("UTCHMMA", ""), "legacy mma.sync path present"


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU failure mode")
def test_no_cpu_fallback_without_gpu():
    lib = v.load()
    assert lib.vitk_device_ok(0) == 0
    cfg = _lib.VitkConfig(224, 224, 16, 3, 768, 12, 3072, 12, 1e-12, 0, 8)
    h = C.c_void_p()
    assert lib.vitk_create(C.byref(cfg), 0, C.byref(h)) != 0
    assert b"no CPU fallback" in lib.vitk_last_error(None)
    with pytest.raises(v.VitkError):
        v.ViTEmbedder(v.ViTGeometry.vit_b16(), {}, device=0)
    with pytest.raises(v.VitkError):
        v.hf_vit_embed_batch(np.zeros((1, 224, 224, 3), np.uint8))


def test_geometry_from_hf_config():
    from transformers import ViTConfig
    g = v.ViTGeometry.from_hf_config(ViTConfig())
    assert (g.hidden, g.heads, g.mlp, g.layers, g.patch, g.tokens) == (768, 12, 3072, 12, 16, 197)
    assert g.ln_eps == pytest.approx(1e-12) and not g.gelu_tanh
    gl = v.ViTGeometry.vit_l16()
    assert (gl.hidden, gl.heads, gl.mlp, gl.layers) == (1024, 16, 4096, 24)


def test_load_label_banks_accepts_reference_layout(golden_dir):
    rows, labels, ids = v.load_label_banks(os.path.join(golden_dir, "banks"))
    assert rows.shape == (1472, 768) and labels.shape == (1472,) and len(ids) == 1472
    assert ids[0].startswith("left:") and ids[-1].startswith("none:")


def test_vote_helpers_follow_reference(golden_dir):
    """determine_class_pass1 / pass2 keep the reference's call (ids, metadatas, distances) and
    return what the reference's OWN functions returned (chroma.py:36-71,102-134 cut out with ast
    and executed: tools/make_golden_determine_class.py), literal threshold 20 included."""
    import json
    z = json.load(open(os.path.join(golden_dir, "determine_class_pass12_golden.json")))
    assert len(z["cases"]) == 96
    for c in z["cases"]:
        metas = [{"label": lab} for lab in c["labels"]]
        ids = [f"id{j}" for j in range(len(metas))]
        for fn, key in ((v.determine_class_pass1, "pass1"), (v.determine_class_pass2, "pass2")):
            got = fn(ids, metas, c["distances"])
            assert got["dir"] == c[key]["dir"]
            assert ("probs" in got) == ("probs" in c[key])
            if "probs" in got:
                assert np.array_equal(np.asarray(got["probs"]), np.asarray(c[key]["probs"]))
    assert v.vote_from_counts([10, 10, 5])["dir"] == "left"            # first max wins
    assert v.vote_from_counts([19, 4, 2], threshold=20)["dir"] == "pass2"
    r = v.vote_from_counts([21, 2, 2], threshold=20)
    assert r["dir"] == "left" and abs(r["probs"].sum() - 1) < 1e-12


def test_collection_where_clauses_vectorised_equal_the_per_row_form():
    """BankCollection.where_mask (columnar, what query(where=...) uploads) against match_where
    row by row, on the clause shapes the reference sends (frame_retriever.py:44-51,
    ratt_chunk_retriever.py:99-105, db_rebuild.py:121) -- host logic, no GPU."""
    rng = np.random.default_rng(3)
    metas = []
    for i in range(500):
        m = {"vid_num": int(rng.integers(1, 6)), "side": "left" if rng.random() < 0.5 else "right",
             "t_norm": float(rng.random()), "label": ["left", "right", "none"][i % 3]}
        if i % 11 == 0:
            del m["t_norm"]                      # a missing key never matches
        if i % 13 == 0:
            m["video"] = f"vid{i % 4}"
        metas.append(m)
    col = object.__new__(v.BankCollection)
    col.metadatas, col._n, col._cols, col._codes = metas, len(metas), {}, {}
    clauses = [
        {"side": "left"},
        {"vid_num": {"$ne": 3}},
        {"$and": [{"vid_num": {"$ne": 2}}, {"side": "right"}, {"t_norm": {"$gte": 0.25}}, {"t_norm": {"$lte": 0.75}}]},
        {"$or": [{"vid_num": 1}, {"t_norm": {"$lt": 0.1}}]},
        {"video": "vid1"},
        {"vid_num": {"$in": [1, 4]}},
        {"label": {"$nin": ["none"]}},
        {"$and": [{"side": "left"}, {"$or": [{"vid_num": {"$gt": 3}}, {"t_norm": {"$gte": 0.9}}]}]},
    ]
    for w in clauses:
        want = np.array([v.match_where(m, w) for m in metas], np.uint8)
        assert np.array_equal(col.where_mask(w), want), w
    assert col.where_mask(None) is None
    assert col.code("side", "left") != col.code("side", "right") and col.code("vid_num", 7) == 7


def test_knn_limits_are_stated_not_silent():
    from vit_research_b200 import _lib
    assert _lib.KNN_MAX_K == 1024 and _lib.STREAM_MAX_K == 64
