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


def test_vote_helpers_follow_reference():
    assert v.determine_class_pass2([10, 10, 5])["dir"] == "left"
    assert v.determine_class_pass1([19, 4, 2])["dir"] == "pass2"
    r = v.determine_class_pass1([21, 2, 2])
    assert r["dir"] == "left" and abs(r["probs"].sum() - 1) < 1e-12
