"""ORACLE (test infrastructure only — never imported by the product path).

CPU restatement of the reference's frame preprocessing, i.e. what `processor(images=frames_list,
return_tensors="pt")` does inside hf_vit_embed_batch (nba_proj/ratt_write_clips_to_ragdb.py:143-146)
with `ViTImageProcessor` [3P transformers 5.5.0 models/vit/image_processing_vit.py:20-27,
image_processing_backends.py TorchvisionBackend.resize/_preprocess]:
  uint8 HWC -> CHW -> torchvision resize to 224x224, BILINEAR, antialias=True, ON UINT8
  -> fused rescale+normalise (x - 127.5) / 127.5 -> float32 NCHW.
The uint8 antialiased resize is ATen's separable fixed-point kernel (the Pillow algorithm):
horizontal pass then vertical pass, uint8 in between, int16 weights with a per-dimension
precision (aten/src/ATen/native/cpu/UpSampleKernel.cpp: _compute_indices_min_size_weights_aa,
_compute_index_ranges_int16_weights).  Restated here in numpy.

Pinned: tests/test_oracle_cpu.py checks this restatement bit for bit against the reference's own
processor (transformers.ViTImageProcessor, importable in this image) on seeded frames, and against
tests/golden/preprocess_golden.npz produced by it (tools/make_golden_preprocess.py).
"""
from __future__ import annotations

import math

import numpy as np


def aa_weights(in_size: int, out_size: int):
    """Per output index: (xmin, xsize, int16 weights[max_interp]) and the shared precision."""
    scale = in_size / out_size
    support = 1.0 * scale if scale >= 1.0 else 1.0          # interp_size(2) * 0.5 [* scale]
    max_interp = int(math.ceil(support)) * 2 + 1
    invscale = 1.0 / scale if scale >= 1.0 else 1.0
    xmin = np.zeros(out_size, np.int64)
    xsize = np.zeros(out_size, np.int64)
    w = np.zeros((out_size, max_interp), np.float64)
    wt_max = 0.0
    for i in range(out_size):
        center = scale * (i + 0.5)
        lo = max(int(center - support + 0.5), 0)
        n = min(int(center + support + 0.5), in_size) - lo
        n = min(max(n, 0), max_interp)
        tot = 0.0
        for j in range(n):
            x = abs((j + lo - center + 0.5) * invscale)
            w[i, j] = 1.0 - x if x < 1.0 else 0.0
            tot += w[i, j]
        if tot != 0.0:
            for j in range(n):
                w[i, j] /= tot
                wt_max = max(wt_max, w[i, j])
        xmin[i], xsize[i] = lo, n
    prec = 0
    while prec < 22:
        if int(0.5 + wt_max * (1 << (prec + 1))) >= (1 << 15):
            break
        prec += 1
    wi = np.where(w < 0, (-0.5 + w * (1 << prec)).astype(np.int64), (0.5 + w * (1 << prec)).astype(np.int64))
    return xmin, xsize, wi.astype(np.int64), prec


def _resize_axis_u8(img: np.ndarray, out_size: int, axis: int) -> np.ndarray:
    in_size = img.shape[axis]
    if in_size == out_size:
        return img
    xmin, xsize, wi, prec = aa_weights(in_size, out_size)
    src = np.moveaxis(img, axis, 0).astype(np.int64)
    out = np.empty((out_size,) + src.shape[1:], np.int64)
    for i in range(out_size):
        acc = np.full(src.shape[1:], 1 << (prec - 1), np.int64)
        for j in range(int(xsize[i])):
            acc += wi[i, j] * src[xmin[i] + j]
        out[i] = acc >> prec
    return np.moveaxis(np.clip(out, 0, 255).astype(np.uint8), 0, axis)


def resize_u8(frames: np.ndarray, out_h: int = 224, out_w: int = 224) -> np.ndarray:
    """(N,H,W,C) uint8 -> (N,out_h,out_w,C) uint8: horizontal pass, then vertical."""
    x = _resize_axis_u8(np.asarray(frames, np.uint8), out_w, 2)
    return _resize_axis_u8(x, out_h, 1)


def pixel_values(frames: np.ndarray, out_h: int = 224, out_w: int = 224) -> np.ndarray:
    """The processor's output: float32 (N,C,out_h,out_w); rescale and normalise fused the way
    transformers 5.5 does: (u8 - 127.5) / 127.5 in float32."""
    r = resize_u8(frames, out_h, out_w).astype(np.float32)
    r = (r - np.float32(127.5)) / np.float32(127.5)
    return np.ascontiguousarray(r.transpose(0, 3, 1, 2))


GOLDEN_CASES = {"tf_432x768": (432, 768), "hd_180x320": (180, 320), "up_60x90": (60, 90),
                "tall_300x224": (300, 224), "odd_97x131": (97, 131)}


def synthetic_frames(name: str, n: int = 1) -> np.ndarray:
    """Seeded (n,h,w,3) uint8 inputs of the golden cases, integer arithmetic only (identical on
    every platform): gradients + blocks + noise, so rounding and the antialias support are hit."""
    h, w = GOLDEN_CASES[name]
    rng = np.random.default_rng(sum(map(ord, name)))
    yy, xx = np.mgrid[0:h, 0:w]
    out = []
    for i in range(n):
        base = (xx * (3 + i) + yy * (5 + 2 * i)) % 256
        img = np.stack([base, base[::-1], base[:, ::-1]], -1).astype(np.int64)
        img += rng.integers(-40, 41, (h, w, 3))
        img[h // 4:h // 2, w // 3:w // 2] = rng.integers(0, 256, 3)
        out.append(np.clip(img, 0, 255).astype(np.uint8))
    return np.stack(out)


# ---------------------------------------------------------------------------------------------
# cv2.resize(..., interpolation=cv2.INTER_AREA), shrinking (the TF path's front-end:
# generate_clips_hmm.py:392, chroma.py:204-205, write_embeddings.py:225-226, loader.py:7).
# Restated from OpenCV's imgproc/resize.cpp: computeResizeAreaTab, ResizeArea_Invoker (float
# accumulation, saturate_cast<uchar> = round half to even); when both scale factors are integers
# ResizeAreaFast (integer box sum * (1.f / area)); and at exactly 2 x 2 its vector form
# ResizeAreaFastVec, the INTEGER mean (a + b + c + d + 2) >> 2 (round half up).
# PINNED: tests/test_oracle_cpu.py compares this restatement byte for byte with the live
# cv2.resize (opencv-python-headless 4.13.0 in this image; the reference pins no version) on 18
# geometries, and tests/golden/preprocess_cv2_golden.npz holds cv2-produced bytes
# (tools/make_golden_preprocess.py) that the CUDA kernels are compared with on the GPU box.
# ---------------------------------------------------------------------------------------------
def area_tab(ssize, dsize):
    """computeResizeAreaTab: list per destination index of (source index, float32 weight)."""
    scale = ssize / dsize
    tab = []
    for dx in range(dsize):
        fsx1 = dx * scale
        fsx2 = fsx1 + scale
        cell = min(scale, ssize - fsx1)
        sx1, sx2 = int(np.ceil(fsx1)), int(np.floor(fsx2))
        sx2 = min(sx2, ssize - 1)
        sx1 = min(sx1, sx2)
        ent = []
        if sx1 - fsx1 > 1e-3:
            ent.append((sx1 - 1, np.float32((sx1 - fsx1) / cell)))
        for sx in range(sx1, sx2):
            ent.append((sx, np.float32(1.0 / cell)))
        if fsx2 - sx2 > 1e-3:
            ent.append((sx2, np.float32(min(min(fsx2 - sx2, 1.0), cell) / cell)))
        tab.append(ent)
    return tab


def resize_area_u8(img, out_h, out_w):
    """img (H,W,C) uint8 -> (out_h,out_w,C) uint8, INTER_AREA shrinking."""
    img = np.asarray(img, np.uint8)
    H, W, C = img.shape
    if H < out_h or W < out_w:
        raise ValueError("INTER_AREA restatement covers shrinking only")
    if H % out_h == 0 and W % out_w == 0:
        sy, sx = H // out_h, W // out_w
        box = img.reshape(out_h, sy, out_w, sx, C).astype(np.int64).sum((1, 3))
        if sy == 2 and sx == 2:     # ResizeAreaFastVec: integer mean, round half up
            return ((box + 2) >> 2).astype(np.uint8)
        val = box.astype(np.float32) * np.float32(1.0 / np.float32(sx * sy))
        return np.clip(np.rint(val), 0, 255).astype(np.uint8)
    xt, yt = area_tab(W, out_w), area_tab(H, out_h)
    src = img.astype(np.float32)
    # horizontal reduction of every source row: buf = 0; buf += S * alpha (separate mul and add)
    hbuf = np.zeros((H, out_w, C), np.float32)
    for dx, ent in enumerate(xt):
        acc = np.zeros((H, C), np.float32)
        for sx, a in ent:
            acc = acc + src[:, sx, :] * a
        hbuf[:, dx, :] = acc
    out = np.zeros((out_h, out_w, C), np.uint8)
    for dy, ent in enumerate(yt):
        acc = None
        for sy, b in ent:
            t = b * hbuf[sy]
            acc = t if acc is None else acc + t
        out[dy] = np.clip(np.rint(acc), 0, 255).astype(np.uint8)
    return out


# ---------------------------------------------------------------------------------------------
# cv2.resize(im, (768, 432)) with the DEFAULT interpolation, INTER_LINEAR -- what
# extract_embeddings feeds the TF model (smarter_generate_clips.py:81).  OpenCV's 8-bit bilinear
# path (imgproc/resize.cpp: resizeGeneric_ / HResizeLinear / VResizeLinear<uchar,int,short>):
# two taps per axis, NO antialiasing when shrinking; coefficients are shorts scaled by 2^11
# (INTER_RESIZE_COEF_BITS), fx = float((dx + .5) * scale - .5) - floor; along x the fraction is
# zeroed at the borders, along y the ROWS are clamped instead; the horizontal pass keeps
# S0 * a0 + S1 * a1 as int, the vertical pass is
#   ((b0 * (S0 >> 4)) >> 16) + ((b1 * (S1 >> 4)) >> 16) + 2) >> 2.
# (At exactly 2 x 2 shrinking OpenCV switches INTER_LINEAR to INTER_AREA; the fixed-point form
# above yields the same bytes there -- checked live in the tests.)
# PINNED like INTER_AREA: byte for byte against the live cv2.resize, shrinking and growing.
# ---------------------------------------------------------------------------------------------
def linear_tab(ssize, dsize, clamp_fraction):
    """Per destination index: (first source index, (a0, a1) int coefficients scaled by 2048)."""
    scale = 1.0 / (dsize / ssize)          # OpenCV: inv_scale = dsize/ssize; scale = 1./inv_scale
    idx = np.zeros(dsize, np.int64)
    coef = np.zeros((dsize, 2), np.int64)
    for d in range(dsize):
        f = np.float32((d + 0.5) * scale - 0.5)
        s = int(np.floor(f))
        f = np.float32(f - np.float32(s))
        if clamp_fraction:
            if s < 0:
                f, s = np.float32(0), 0
            if s >= ssize - 1:
                f, s = np.float32(0), ssize - 1
        idx[d] = s
        coef[d, 0] = int(np.rint(np.float32((np.float32(1.0) - f) * np.float32(2048))))
        coef[d, 1] = int(np.rint(np.float32(f * np.float32(2048))))
    return idx, coef


def resize_linear_u8(img, out_h, out_w):
    """img (H,W,C) uint8 -> (out_h,out_w,C) uint8, cv2 INTER_LINEAR (any direction)."""
    img = np.asarray(img, np.uint8)
    H, W, _ = img.shape
    xs, xa = linear_tab(W, out_w, True)
    ys, ya = linear_tab(H, out_h, False)
    S = img.astype(np.int64)
    hb = S[:, xs, :] * xa[:, 0][None, :, None] + S[:, np.minimum(xs + 1, W - 1), :] * xa[:, 1][None, :, None]
    S0, S1 = hb[np.clip(ys, 0, H - 1)], hb[np.clip(ys + 1, 0, H - 1)]
    b0, b1 = ya[:, 0][:, None, None], ya[:, 1][:, None, None]
    out = (((b0 * (S0 >> 4)) >> 16) + ((b1 * (S1 >> 4)) >> 16) + 2) >> 2
    return np.clip(out, 0, 255).astype(np.uint8)


# geometries of the cv2-produced golden file (tests/golden/preprocess_cv2_golden.npz): source
# (h, w) -> the model resolution used with it; frames come from cv2_case_frame (seeded).
CV2_GOLDEN_CASES = {
    "hd1080_to_tf": ((1080, 1920), (432, 768)), "hd720_to_tf": ((720, 1280), (432, 768)),
    "x2_to_tf": ((864, 1536), (432, 768)), "x3_to_tf": ((1296, 2304), (432, 768)),
    "x2x1_to_tf": ((432, 1536), (432, 768)), "odd_to_tf": ((433, 769), (432, 768)),
    "wvga_to_tf": ((480, 854), (432, 768)), "x2_to_224": ((448, 448), (224, 224)),
    "x5_to_224": ((1120, 1120), (224, 224)), "odd_to_224": ((1000, 225), (224, 224)),
}


def cv2_case_frame(name):
    (h, w), _ = CV2_GOLDEN_CASES[name]
    return np.random.default_rng(sum(map(ord, name)) + 17).integers(0, 256, (h, w, 3), dtype=np.uint8)
