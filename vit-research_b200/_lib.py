"""ctypes binding of libvitk.so (include/vitk.h) + the in-tree nvcc build.

This is the binding a maintainer of the reference would add: the reference is
pure Python, so the FFI is ctypes over a C ABI with plain pointers and sizes.  There is no CPU
fallback: if the shared library is missing or no sm_100 GPU is visible, the compute entry points
raise.
"""
from __future__ import annotations

import ctypes as C
import os
import shutil
import subprocess
from pathlib import Path

PKG_DIR = Path(__file__).resolve().parent
CSRC = PKG_DIR / "csrc"
LIB_PATH = PKG_DIR / "libvitk.so"
INCLUDE = PKG_DIR.parent / "include"

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "--shared", "-Xcompiler", "-fPIC",
]


class VitkError(RuntimeError):
    pass


class VitkConfig(C.Structure):
    _fields_ = [
        ("image_h", C.c_int32), ("image_w", C.c_int32), ("patch", C.c_int32),
        ("channels", C.c_int32), ("hidden", C.c_int32), ("heads", C.c_int32),
        ("mlp", C.c_int32), ("layers", C.c_int32), ("ln_eps", C.c_float),
        ("gelu_tanh", C.c_int32), ("max_batch", C.c_int32),
        ("u8_mean", C.c_float), ("u8_std", C.c_float),
    ]


class VitkRowFilter(C.Structure):
    """vitk_row_filter (include/vitk.h): per-query row predicate of vitk_knn_vote_filtered."""
    _fields_ = [
        ("ne_col", C.c_void_p), ("q_ne", C.c_void_p), ("eq_col", C.c_void_p), ("q_eq", C.c_void_p),
        ("rng_col", C.c_void_p), ("q_lo", C.c_void_p), ("q_hi", C.c_void_p),
        ("q_lo2", C.c_void_p), ("q_hi2", C.c_void_p), ("rng_round_f32", C.c_int32),
    ]


PIX_F32_NCHW, PIX_BF16_NCHW, PIX_U8_NHWC = 0, 1, 2
METRIC_COSINE, METRIC_L2 = 0, 1
RESIZE_BILINEAR_AA, RESIZE_AREA, RESIZE_LINEAR = 0, 1, 2
EPI_BIAS_BF16, EPI_BIAS_GELU_BF16, EPI_BIAS_GELUT_BF16 = 0, 1, 2
EPI_BIAS_RESID_F32, EPI_TABLE_F32, EPI_F32 = 3, 4, 5
EPI_BIAS_RESID_BF16 = 7
KNN_MAX_K = 1024          # vitk_knn_vote*: 1 <= k <= 1024
STREAM_MAX_K = 64         # vitk_stream_submit* / vitk_embed_vote_host_u8 (the vote: k = 5 / 25 / 50)

# name -> (restype, argtypes); every symbol include/vitk.h declares
_vp, _i, _f = C.c_void_p, C.c_int, C.c_float
SIGNATURES = {
    "vitk_version": (C.c_char_p, []),
    "vitk_device_ok": (_i, [_i]),
    "vitk_create": (_i, [C.POINTER(VitkConfig), _i, C.POINTER(_vp)]),
    "vitk_destroy": (_i, [_vp]),
    "vitk_last_error": (C.c_char_p, [_vp]),
    "vitk_set_weight": (_i, [_vp, C.c_char_p, _vp, C.POINTER(C.c_int64), _i]),
    "vitk_finalize_weights": (_i, [_vp]),
    "vitk_embed": (_i, [_vp, _vp, _i, _i, _vp, _vp, _i, _vp]),
    "vitk_embed_host_u8": (_i, [_vp, _vp, _i, _vp, _i]),
    "vitk_embed_vote_host_u8": (_i, [_vp, _vp, _i, _i, _i, _vp, _vp, _vp, _vp, _vp]),
    "vitk_stream_submit": (_i, [_vp, _i, _vp, _i, _i, _i, _vp, _vp, _vp, _vp, _vp]),
    "vitk_stream_wait": (_i, [_vp, _i]),
    "vitk_stream_submit_raw": (_i, [_vp, _i, _vp, _i, _i, _i, _i, _i, _i, _i, _vp, _vp, _vp, _vp, _vp]),
    "vitk_preprocess_resize_linear_u8": (_i, [_vp, _vp, _i, _i, _i, _vp, _vp]),
    "vitk_knn_vote_filtered": (_i, [_vp, _vp, _i, _i, _i, _i, _vp, C.POINTER(VitkRowFilter), _vp, _vp, _vp, _vp, _vp]),
    "vitk_bank_new": (_i, [_vp]),
    "vitk_bank_select": (_i, [_vp, _i]),
    "vitk_bank_current": (_i, [_vp]),
    "vitk_bank_free": (_i, [_vp, _i]),
    "vitk_op_merge_topk": (_i, [_vp, _vp, _i, _i, _i, _i, _vp, _vp, _vp]),
    "vitk_preprocess_resize_u8": (_i, [_vp, _vp, _i, _i, _i, _vp, _vp]),
    "vitk_embed_resize_host_u8": (_i, [_vp, _vp, _i, _i, _i, _vp, _i]),
    "vitk_bank_set": (_i, [_vp, _vp, _i, _i, _vp, _i, _i]),
    "vitk_bank_set_device": (_i, [_vp, _vp, _i, _i, _vp, _i, _i, _vp]),
    "vitk_knn_vote": (_i, [_vp, _vp, _i, _i, _i, _i, _vp, _vp, _vp, _vp, _vp]),
    "vitk_knn_vote_masked": (_i, [_vp, _vp, _i, _i, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp]),
    "vitk_bank_write": (_i, [_vp, _i, _vp, _i, _vp, _i, _i, _vp]),
    "vitk_bank_rows": (_i, [_vp]),
    "vitk_hmm_decode_counts": (_i, [_vp, _vp, _i, _vp, _i, _vp, _vp, _vp]),
    "vitk_hmm_decode_counts_static": (_i, [_vp, _i, _i, _vp, _vp]),
    "vitk_hmm_decode_probs_host": (_i, [_vp, _vp, _vp, _i, _vp, _vp]),
    "vitk_hmm_stream_begin": (_i, [_vp, C.c_longlong]),
    "vitk_hmm_stream_append_counts": (_i, [_vp, _vp, _i, _i, _vp]),
    "vitk_hmm_stream_finish": (_i, [_vp, _vp, C.POINTER(C.c_longlong), _vp]),
    "vitk_op_gemm": (_i, [_vp, _vp, _vp, _i, _i, _i, _i, _vp, _vp, _vp, _i, _vp]),
    "vitk_op_gemm_fused": (_i, [_vp, _vp, _vp, _i, _i, _i, _i, _vp, _vp, _vp, _i, _f, _vp, _vp, _vp, _vp]),
    "vitk_op_rowstats": (_i, [_vp, _vp, _vp, _i, _i, _i, _vp]),
    "vitk_preprocess_resize_area_u8": (_i, [_vp, _vp, _i, _i, _i, _vp, _vp]),
    "vitk_op_attention_small": (_i, [_vp, _vp, _vp, _i, _i, _i, _i, _vp]),
    "vitk_op_layernorm_rows_f32": (_i, [_vp, _vp, _vp, _vp, _i, C.c_longlong, _i, _f, _vp]),
    "vitk_op_chunk_head": (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _vp]),
    "vitk_op_layernorm": (_i, [_vp, _vp, _vp, _vp, _i, _i, _f, _vp]),
    "vitk_op_attention": (_i, [_vp, _vp, _i, _i, _i, _vp]),
    "vitk_op_patchify": (_i, [_vp, _i, _vp, _i, _i, _i, _i, _i, _vp]),
    "vitk_profile_fc1": (_i, [_vp, _i]),
    "vitk_profile_fc1_read": (_i, [_vp, C.POINTER(C.c_double), C.POINTER(_i)]),
    "vitk_launch_count": (C.c_int64, [_vp]),
    "vitk_knn_redo_count": (C.c_int64, [_vp]),
}


def sources() -> list[Path]:
    return sorted(CSRC.glob("*.cu")) + sorted(CSRC.glob("*.cuh")) + [INCLUDE / "vitk.h"]


def needs_build() -> bool:
    if not LIB_PATH.exists():
        return True
    t = LIB_PATH.stat().st_mtime
    return any(s.stat().st_mtime > t for s in sources())


def build(force: bool = False, verbose: bool = False) -> Path:
    """Compile csrc/*.cu for sm_100a into libvitk.so next to this file (nvcc cross-compiles
    without a GPU)."""
    if not force and not needs_build():
        return LIB_PATH
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(nvcc):
        raise VitkError("nvcc not found: cannot build libvitk.so")
    tmp = LIB_PATH.with_suffix(".so.tmp%d" % os.getpid())
    extra = os.environ.get("VITK_NVCC_EXTRA", "").split()     # developer builds, e.g. -DVITK_ATT_POLY_MASK=0x8888u
    cmd = [nvcc, *NVCC_FLAGS, *extra, "-o", str(tmp), str(CSRC / "vitk_api.cu")]
    if verbose:
        cmd.insert(1, "-Xptxas=-v")
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise VitkError("nvcc failed:\n" + r.stdout + r.stderr)
    if verbose:
        print(r.stderr)
    os.replace(tmp, LIB_PATH)
    return LIB_PATH


_lib = None


def load(auto_build: bool = True):
    """dlopen libvitk.so and attach the prototypes.  Raises if it cannot be built/loaded."""
    global _lib
    if _lib is not None:
        return _lib
    global LIB_PATH
    if os.environ.get("VITK_LIB"):              # developer A/B runs against another build of the library
        LIB_PATH = Path(os.environ["VITK_LIB"]).resolve()
        auto_build = False
    if auto_build and needs_build():
        build()
    if not LIB_PATH.exists():
        raise VitkError(f"{LIB_PATH} is missing (run __graft_entry__.build()); no CPU fallback")
    lib = C.CDLL(str(LIB_PATH))
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the symbol is not exported
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(rc: int, handle=None, what: str = "vitk"):
    if rc != 0:
        msg = load().vitk_last_error(handle)
        raise VitkError(f"{what} failed: {msg.decode() if msg else 'unknown error'}")


def ptr(t):
    """Device/host pointer of a torch tensor or numpy array (None -> NULL)."""
    if t is None:
        return None
    if hasattr(t, "data_ptr"):
        return C.c_void_p(t.data_ptr())
    return C.c_void_p(t.ctypes.data)


def stream_ptr(stream=None):
    import torch

    s = stream if stream is not None else torch.cuda.current_stream()
    return C.c_void_p(s.cuda_stream)
