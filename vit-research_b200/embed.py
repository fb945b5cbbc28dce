"""Embedding entry points with the reference's call signatures, backed by libvitk (sm_100a).

Mirrors, for the hot path only:
  hf_vit_embed_batch(frames_np) -> (N,768) float32, unit rows
      nba_proj/ratt_write_clips_to_ragdb.py:135-159 (+10 copies)
  extract_embeddings(vit_model, image_paths, batch_size=1024) -> (N,768)
      nba_proj/smarter_generate_clips.py:69-94
  model.predict(x, batch_size=, verbose=)['pre_logits'] -> (B,1,1,768)
      nba_proj/write_embeddings.py:185-188
  vit_model(pixel_values=...).last_hidden_state[:, 0, :]
      nba_proj/ratt_write_clips_to_ragdb.py:146-148
PyTorch is used for device memory and streams only; all arithmetic runs in libvitk's kernels.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass
from types import SimpleNamespace

import numpy as np

from . import _lib
from ._lib import VitkConfig, VitkError, check, ptr, stream_ptr


@dataclass
class ViTGeometry:
    image_h: int = 224
    image_w: int = 224
    patch: int = 16
    channels: int = 3
    hidden: int = 768
    heads: int = 12
    mlp: int = 3072
    layers: int = 12
    ln_eps: float = 1e-12
    gelu_tanh: bool = False
    u8_mean: float = 127.5   # uint8 pixels enter the model as (x - u8_mean) / u8_std: the HF
    u8_std: float = 127.5    # processor's fused rescale+normalise; the TF scripts feed raw pixels (0, 1)

    @property
    def tokens(self) -> int:
        return (self.image_h // self.patch) * (self.image_w // self.patch) + 1

    @classmethod
    def vit_b16(cls) -> "ViTGeometry":
        return cls()

    @classmethod
    def vit_l16(cls) -> "ViTGeometry":
        return cls(hidden=1024, heads=16, mlp=4096, layers=24)

    @classmethod
    def tf_432x768(cls, patch: int = 32) -> "ViTGeometry":
        """The reference's TF-path hyper-parameters (generate_clips_hmm.py:48-56,
        models/vit_backbone.py:158-177): 432x768 raw-pixel frames, patch 32 'valid' (313 tokens),
        LayerNorm eps 1e-6, tanh-GELU.  (patch 256 -> 3 patches + CLS is the bank-writing variant,
        write_embeddings.py:101-109.)"""
        return cls(image_h=432, image_w=768, patch=patch, ln_eps=1e-6, gelu_tanh=True,
                   u8_mean=0.0, u8_std=1.0)

    @classmethod
    def from_hf_config(cls, cfg) -> "ViTGeometry":
        img = cfg.image_size if isinstance(cfg.image_size, (tuple, list)) else (cfg.image_size,) * 2
        act = getattr(cfg, "hidden_act", "gelu")
        if act not in ("gelu", "gelu_new", "gelu_pytorch_tanh"):
            raise VitkError(f"unsupported hidden_act {act!r}")
        return cls(image_h=int(img[0]), image_w=int(img[1]), patch=int(cfg.patch_size),
                   channels=int(cfg.num_channels), hidden=int(cfg.hidden_size),
                   heads=int(cfg.num_attention_heads), mlp=int(cfg.intermediate_size),
                   layers=int(cfg.num_hidden_layers), ln_eps=float(cfg.layer_norm_eps),
                   gelu_tanh=act != "gelu")


def random_init_state_dict(geometry: "ViTGeometry", seed: int = 0, cls_zeros: bool = False) -> dict:
    """Random-init weights with HF ViT's initialiser statistics (modeling_vit.py:384-399:
    trunc-normal(0, 0.02) for Linear/Conv/pos-emb/CLS, zero biases, LayerNorm gamma=1 beta=0),
    keyed like the HF state dict.  Used by bench.py / smoke() — there is no network for
    checkpoints, and the product path must not depend on transformers or the oracle.
    cls_zeros=True: the TF backbone's TokenLayer initialises the CLS token to zeros
    (models/vit_backbone.py:93-104) -- the state a freshly built reference TF model is in when
    write_embeddings.py:101-109,243 saves `vit_random_weights.h5`."""
    import torch

    g = geometry
    gen = torch.Generator().manual_seed(seed)
    D, F, ps, C_ = g.hidden, g.mlp, g.patch, g.channels

    def tn(*shape):
        t = torch.empty(*shape)
        torch.nn.init.trunc_normal_(t, mean=0.0, std=0.02, generator=gen)
        return t

    sd = {
        "embeddings.cls_token": torch.zeros(1, 1, D) if cls_zeros else tn(1, 1, D),
        "embeddings.position_embeddings": tn(1, g.tokens, D),
        "embeddings.patch_embeddings.projection.weight": tn(D, C_, ps, ps),
        "embeddings.patch_embeddings.projection.bias": torch.zeros(D),
        "layernorm.weight": torch.ones(D), "layernorm.bias": torch.zeros(D),
    }
    for i in range(g.layers):
        p = f"encoder.layer.{i}."
        for n in ("query", "key", "value"):
            sd[p + f"attention.attention.{n}.weight"] = tn(D, D)
            sd[p + f"attention.attention.{n}.bias"] = torch.zeros(D)
        sd[p + "attention.output.dense.weight"] = tn(D, D)
        sd[p + "attention.output.dense.bias"] = torch.zeros(D)
        sd[p + "intermediate.dense.weight"] = tn(F, D)
        sd[p + "intermediate.dense.bias"] = torch.zeros(F)
        sd[p + "output.dense.weight"] = tn(D, F)
        sd[p + "output.dense.bias"] = torch.zeros(D)
        for n in ("layernorm_before", "layernorm_after"):
            sd[p + n + ".weight"] = torch.ones(D)
            sd[p + n + ".bias"] = torch.zeros(D)
    return sd


class ViTEmbedder:
    """Owns one vitk handle (weights + workspace on one GPU)."""

    def __init__(self, geometry: ViTGeometry, state_dict, device: int = 0, max_batch: int = 256):
        import torch

        self.lib = _lib.load()
        self.geometry = geometry
        self.device = int(device)
        self.max_batch = int(max_batch)
        self.handle = C.c_void_p()
        if not torch.cuda.is_available():
            raise VitkError("no CUDA device visible: vitk has no CPU fallback")
        torch.cuda.set_device(self.device)
        g = geometry
        cfg = VitkConfig(g.image_h, g.image_w, g.patch, g.channels, g.hidden, g.heads, g.mlp,
                         g.layers, g.ln_eps, int(g.gelu_tanh), self.max_batch, float(g.u8_mean),
                         float(g.u8_std))
        check(self.lib.vitk_create(C.byref(cfg), self.device, C.byref(self.handle)), None,
              "vitk_create")
        self._weights = {}          # HF key -> float32 host array (what save_weights writes)
        self._set_weights(state_dict)

    def _set_weights(self, state_dict):
        for name, t in state_dict.items():
            if name.startswith("vit."):
                name = name[4:]
            if name.startswith("pooler."):
                continue
            a = t.detach().cpu().float().contiguous().numpy() if hasattr(t, "detach") else \
                np.ascontiguousarray(t, dtype=np.float32)
            shape = (C.c_int64 * a.ndim)(*a.shape)
            check(self.lib.vitk_set_weight(self.handle, name.encode(), ptr(a), shape, a.ndim),
                  self.handle, f"vitk_set_weight({name})")
            self._weights[name] = a
        check(self.lib.vitk_finalize_weights(self.handle), self.handle, "vitk_finalize_weights")

    # ---- Keras-style weight files: model.save_weights(path) / model.load_weights(path) ------
    def save_weights(self, path):
        """`model.save_weights('vit_random_weights.h5')` (write_embeddings.py:243): every tensor
        under its HF ViTModel state-dict key, float32, in one .npz (the container store.py also
        uses for the banks).  HDF5 is not written: this image has no h5py, and the reference's
        .h5 holds TF Model Garden variable names."""
        path = str(path)
        if path.endswith((".h5", ".hdf5", ".keras")):
            raise VitkError("save_weights writes .npz (HF state-dict keys); HDF5 needs h5py, which is absent")
        np.savez(path, **self._weights)
        return path if path.endswith(".npz") else path + ".npz"

    def load_weights(self, path):
        """`model.load_weights(path)` (chroma.py:159, generate_clips_hmm.py:59): reads the .npz
        written by save_weights (or any npz / dict keyed by HF state-dict names), re-uploads and
        re-folds the LayerNorms.  Shapes are checked by vitk_set_weight."""
        if isinstance(path, dict):
            sd = path
        else:
            path = str(path)
            if path.endswith((".h5", ".hdf5", ".keras")):
                raise VitkError("load_weights reads .npz files keyed by HF ViTModel names; an HDF5 "
                                "checkpoint of the TF model needs h5py + the Model Garden name map (absent here)")
            z = np.load(path if path.endswith(".npz") else path + ".npz")
            sd = {k: z[k] for k in z.files}
        missing = set(self._weights) - {k[4:] if k.startswith("vit.") else k for k in sd}
        if missing:
            raise VitkError(f"load_weights: {len(missing)} tensors missing, e.g. {sorted(missing)[0]}")
        self._set_weights(sd)
        return self

    @classmethod
    def from_hf(cls, hf_model, device: int = 0, max_batch: int = 256) -> "ViTEmbedder":
        """hf_model: transformers.ViTModel (the reference's `vit_model`)."""
        return cls(ViTGeometry.from_hf_config(hf_model.config), hf_model.state_dict(), device,
                   max_batch)

    def close(self):
        if getattr(self, "handle", None) and self.handle.value:
            self.lib.vitk_destroy(self.handle)
            self.handle = C.c_void_p()

    def __del__(self):  # pragma: no cover
        try:
            self.close()
        except Exception:
            pass

    @property
    def launch_count(self) -> int:
        return int(self.lib.vitk_launch_count(self.handle))

    # ---- device-resident path -------------------------------------------------------------
    def embed(self, pixels, l2_normalise: bool = True, out=None, out_bf16=None, stream=None):
        """pixels: CUDA tensor — (N,3,H,W) float32/bfloat16 (already processed like
        ViTImageProcessor) or (N,H,W,3) uint8 raw frames.  Returns (N,hidden) float32 CUDA."""
        import torch

        g = self.geometry
        if not pixels.is_cuda:
            raise VitkError("embed() wants a CUDA tensor; use embed_frames() for host arrays")
        pixels = pixels.contiguous()
        if pixels.dtype == torch.uint8:
            fmt, want = _lib.PIX_U8_NHWC, (g.image_h, g.image_w, g.channels)
        elif pixels.dtype == torch.float32:
            fmt, want = _lib.PIX_F32_NCHW, (g.channels, g.image_h, g.image_w)
        elif pixels.dtype == torch.bfloat16:
            fmt, want = _lib.PIX_BF16_NCHW, (g.channels, g.image_h, g.image_w)
        else:
            raise VitkError(f"unsupported pixel dtype {pixels.dtype}")
        if pixels.ndim != 4 or tuple(pixels.shape[1:]) != want:
            # same failure mode as HF ViTPatchEmbeddings (modeling_vit.py:155-165)
            raise ValueError(f"Input image size {tuple(pixels.shape)} doesn't match model "
                             f"(N,{want[0]},{want[1]},{want[2]}) for dtype {pixels.dtype}.")
        n = pixels.shape[0]
        if out is None:
            out = torch.empty((n, g.hidden), dtype=torch.float32, device=pixels.device)
        check(self.lib.vitk_embed(self.handle, ptr(pixels), fmt, n, ptr(out), ptr(out_bf16),
                                  int(l2_normalise), stream_ptr(stream)), self.handle, "vitk_embed")
        return out

    def resize_frames(self, frames, out=None, stream=None, interpolation="bilinear_aa"):
        """CUDA uint8 (N,H,W,C) frames of any size -> CUDA uint8 (N,image_h,image_w,C).
        interpolation="bilinear_aa": the HF processor's resize (BILINEAR, antialias, on uint8) as
        one kernel, bit-identical.  interpolation="area": cv2.INTER_AREA shrinking, the TF path's
        front-end (generate_clips_hmm.py:392).  interpolation="linear": cv2.resize's default
        INTER_LINEAR, what extract_embeddings applies (smarter_generate_clips.py:81).  Both cv2
        forms are byte-identical to OpenCV 4.13."""
        import torch

        g = self.geometry
        if not frames.is_cuda or frames.dtype != torch.uint8 or frames.dim() != 4 or \
                frames.shape[3] != g.channels:
            raise VitkError(f"resize_frames() wants a CUDA uint8 (N,H,W,{g.channels}) tensor")
        frames = frames.contiguous()
        n, ih, iw = frames.shape[:3]
        if out is None:
            out = torch.empty((n, g.image_h, g.image_w, g.channels), dtype=torch.uint8, device=frames.device)
        if interpolation == "area":
            check(self.lib.vitk_preprocess_resize_area_u8(self.handle, ptr(frames), n, ih, iw, ptr(out),
                                                          stream_ptr(stream)), self.handle,
                  "vitk_preprocess_resize_area_u8")
            return out
        if interpolation == "linear":
            check(self.lib.vitk_preprocess_resize_linear_u8(self.handle, ptr(frames), n, ih, iw, ptr(out),
                                                            stream_ptr(stream)), self.handle,
                  "vitk_preprocess_resize_linear_u8")
            return out
        if interpolation != "bilinear_aa":
            raise VitkError(f"unknown interpolation {interpolation!r}")
        check(self.lib.vitk_preprocess_resize_u8(self.handle, ptr(frames), n, ih, iw, ptr(out),
                                                 stream_ptr(stream)), self.handle,
              "vitk_preprocess_resize_u8")
        return out

    # ---- host path (what hf_vit_embed_batch does) -----------------------------------------
    def embed_frames(self, frames_np: np.ndarray, l2_normalise: bool = True,
                     interpolation: str = "bilinear_aa") -> np.ndarray:
        """(N,H,W,3) uint8 host frames of ANY size -> (N,hidden) float32 host.  Frames that are not
        at the model's resolution are resized on the GPU: interpolation="bilinear_aa" is the HF
        processor's resize (hf_vit_embed_batch), "area" / "linear" are the cv2 resizes of the TF
        scripts (generate_clips_hmm.py:392 / smarter_generate_clips.py:81)."""
        g = self.geometry
        frames_np = np.ascontiguousarray(frames_np)
        if frames_np.dtype != np.uint8:
            raise VitkError("embed_frames wants uint8")
        if frames_np.ndim != 4 or frames_np.shape[3] != g.channels:
            raise ValueError(f"frames {frames_np.shape} are not (N,H,W,{g.channels})")
        n = frames_np.shape[0]
        raw = frames_np.shape[1:3] != (g.image_h, g.image_w)
        out = np.empty((n, g.hidden), dtype=np.float32)
        if n > self.max_batch or (raw and (interpolation != "bilinear_aa" or not l2_normalise)):
            # the reference's 1024-frame batches (FRAME_BATCH_SIZE, ratt_write_clips_to_ragdb.py:29):
            # stream max_batch-sized chunks so uploads / resizes / downloads overlap the forward
            chunks = (frames_np[a:a + self.max_batch] for a in range(0, n, self.max_batch))
            a = 0
            for emb, *_ in self.embed_and_vote_stream(chunks, interpolation=interpolation,
                                                      l2_normalise=l2_normalise):
                out[a:a + emb.shape[0]] = emb
                a += emb.shape[0]
            return out
        if raw:
            # the processor's antialiased bilinear uint8 resize runs on the GPU (preprocess.cuh),
            # bit-identical to ViTImageProcessor
            ih, iw = frames_np.shape[1:3]
            check(self.lib.vitk_embed_resize_host_u8(self.handle, ptr(frames_np), n, ih, iw, ptr(out),
                                                     int(l2_normalise)), self.handle,
                  "vitk_embed_resize_host_u8")
            return out
        check(self.lib.vitk_embed_host_u8(self.handle, ptr(frames_np), n, ptr(out), int(l2_normalise)),
              self.handle, "vitk_embed_host_u8")
        return out

    def embed_and_vote(self, frames_np: np.ndarray, index, k: int = 25, out=None):
        """Host frames -> (embeddings (N,hidden) f32, labels (N,) i32, idx (N,k) i32,
        score (N,k) f32, counts (N,classes) i32), all host arrays: the per-batch work of the
        reference's labelling loops (generate_clips_hmm.py:387-412) in one call.  `index` is the
        BankIndex set on this embedder; `out` may hold preallocated arrays (e.g. pinned)."""
        g = self.geometry
        frames_np = np.ascontiguousarray(frames_np)
        if frames_np.dtype != np.uint8 or frames_np.ndim != 4 or \
                frames_np.shape[1:] != (g.image_h, g.image_w, g.channels):
            raise ValueError(f"frames {frames_np.shape}/{frames_np.dtype} do not match uint8 "
                             f"(N,{g.image_h},{g.image_w},{g.channels})")
        n, nc = frames_np.shape[0], max(index.n_classes, 1)
        if k > _lib.STREAM_MAX_K:
            raise VitkError(f"embed_and_vote: k={k} > {_lib.STREAM_MAX_K}; use BankIndex.search for larger k")
        index.select()
        if out is None:
            out = (np.empty((n, g.hidden), np.float32), np.empty((n,), np.int32),
                   np.empty((n, k), np.int32), np.empty((n, k), np.float32),
                   np.empty((n, nc), np.int32))
        emb, label, idx, score, counts = out
        check(self.lib.vitk_embed_vote_host_u8(self.handle, ptr(frames_np), n, int(k), index.metric,
                                               ptr(emb), ptr(label), ptr(idx), ptr(score),
                                               ptr(counts)), self.handle, "vitk_embed_vote_host_u8")
        return emb, label, idx, score, counts

    def embed_and_vote_stream(self, batches, index=None, k: int = 25, outs=None,
                              interpolation: str = "bilinear_aa", l2_normalise: bool = True):
        """Streaming form of embed_and_vote for long frame lists (the reference's batched
        drivers: build_frame_store, ratt_write_clips_to_ragdb.py:197-209).  `batches` yields
        uint8 (n<=max_batch,H,W,C) arrays (pinned for true copy/compute overlap); results are
        yielded in order as (emb, label, idx, score, counts) host arrays, with two batches in
        flight: the upload of batch i+1 and the download of batch i-1 overlap the forward of
        batch i.  index=None (or k=0) skips the vote.  `outs`: optional pair of preallocated
        output tuples, reused alternately (valid until the next-but-one result is produced).
        Batches whose frames are not at the model's resolution (what build_frame_store feeds:
        PIL frames at the video's size) are uploaded raw and resized on the GPU inside the same
        overlapped pipeline (vitk_stream_submit_raw), `interpolation` as in embed_frames."""
        g = self.geometry
        interp = {"bilinear_aa": _lib.RESIZE_BILINEAR_AA, "area": _lib.RESIZE_AREA,
                  "linear": _lib.RESIZE_LINEAR}.get(interpolation)
        if interp is None:
            raise VitkError(f"unknown interpolation {interpolation!r}")
        k = int(k) if index is not None else 0
        if k > _lib.STREAM_MAX_K:
            raise VitkError(f"embed_and_vote_stream: k={k} > {_lib.STREAM_MAX_K}; query the index with "
                            f"BankIndex.search (k <= {_lib.KNN_MAX_K}) instead")
        if index is not None:
            index.select()
        nc = max(index.n_classes, 1) if index is not None else 1
        metric = index.metric if index is not None else 0
        pending = []  # (slot, frames_ref, out, n)

        def finish():
            slot, _keep, out, n = pending.pop(0)
            check(self.lib.vitk_stream_wait(self.handle, slot), self.handle, "vitk_stream_wait")
            emb, label, idx, score, counts = out
            return emb[:n], label[:n], idx[:n], score[:n], counts[:n]

        for i, frames_np in enumerate(batches):
            frames_np = np.ascontiguousarray(frames_np)
            if frames_np.dtype != np.uint8 or frames_np.ndim != 4 or frames_np.shape[3] != g.channels:
                raise ValueError(f"frames {frames_np.shape}/{frames_np.dtype} are not uint8 (N,H,W,{g.channels})")
            n = frames_np.shape[0]
            if n > self.max_batch:
                raise ValueError(f"a streamed batch holds at most max_batch={self.max_batch} frames, got {n}")
            raw = frames_np.shape[1:3] != (g.image_h, g.image_w)
            slot = i & 1
            if len(pending) == 2:
                yield finish()
            if outs is not None:
                out = outs[slot]
            else:
                kk = max(k, 1)
                out = (np.empty((n, g.hidden), np.float32), np.empty((n,), np.int32),
                       np.empty((n, kk), np.int32), np.empty((n, kk), np.float32),
                       np.empty((n, nc), np.int32))
            emb, label, idx, score, counts = out
            if index is not None:
                index.select()      # the caller may have queried another bank between batches
            if raw or not l2_normalise:
                ih, iw = frames_np.shape[1:3]
                check(self.lib.vitk_stream_submit_raw(self.handle, slot, ptr(frames_np), n, ih, iw, interp,
                                                      int(l2_normalise), k, metric, ptr(emb), ptr(label),
                                                      ptr(idx), ptr(score), ptr(counts)),
                      self.handle, "vitk_stream_submit_raw")
            else:
                check(self.lib.vitk_stream_submit(self.handle, slot, ptr(frames_np), n, k, metric,
                                                  ptr(emb), ptr(label), ptr(idx), ptr(score), ptr(counts)),
                      self.handle, "vitk_stream_submit")
            pending.append((slot, frames_np, out, n))
        while pending:
            yield finish()

    # ---- HF-style call: vit_model(pixel_values=...).last_hidden_state[:,0,:] ---------------
    def __call__(self, pixel_values=None, **_):
        """Returns an object whose .last_hidden_state is (N,1,hidden): only the CLS row exists
        (the reference never reads the other 196, ratt_write_clips_to_ragdb.py:148)."""
        cls = self.embed(pixel_values, l2_normalise=False)
        return SimpleNamespace(last_hidden_state=cls[:, None, :], pooler_output=None)

    # ---- Keras-style call: model.predict(x)['pre_logits'] ---------------------------------
    def predict(self, x, batch_size: int = 32, verbose: int = 0):
        """x: (B,H,W,3) uint8/float host array.  Returns {'pre_logits': (B,1,1,hidden)}
        (write_embeddings.py:185-188, rag_vit.py:463-464).  Pixels enter the model as
        (x - geometry.u8_mean) / geometry.u8_std: the HF processor's values by default, raw 0..255
        like the TF scripts with ViTGeometry.tf_432x768()."""
        x = np.asarray(x)
        if x.dtype != np.uint8:
            x = np.clip(x, 0, 255).astype(np.uint8)
        emb = self.embed_frames(x, l2_normalise=False)
        return {"pre_logits": emb.reshape(-1, 1, 1, self.geometry.hidden)}


# ---- module-global style used by the reference scripts ----------------------------------------
_default: ViTEmbedder | None = None


def install(embedder: ViTEmbedder) -> None:
    """Set the module-global model, like the reference's `vit_model = ViTModel...to(device)`."""
    global _default
    _default = embedder


def hf_vit_embed_batch(frames_np):
    """
    frames_np: (N, H, W, 3) uint8 or float32
    returns: (N, 768) float32          [ratt_write_clips_to_ragdb.py:135-159]
    """
    if _default is None:
        raise VitkError("call vit_research_b200.install(ViTEmbedder...) first")
    frames_np = np.asarray(frames_np)
    if frames_np.dtype != np.uint8:
        frames_np = np.clip(frames_np, 0, 255).astype(np.uint8)
    return _default.embed_frames(frames_np, l2_normalise=True)


def extract_embeddings(vit_model, image_paths, batch_size=1024, loader=None):
    """smarter_generate_clips.py:69-94: load each path, resize to the model resolution, predict in
    batches, stack the un-normalised pre_logits.

    The reference resizes with `cv2.resize(im, (768, 432))` -- cv2's DEFAULT interpolation,
    INTER_LINEAR (smarter_generate_clips.py:81; the other TF call sites use INTER_AREA).  Here the
    default loader only decodes (imread + BGR2RGB, :79-80); a batch of equally sized frames is
    uploaded raw and resized on the GPU by the INTER_LINEAR kernel (byte-identical to cv2), frames
    of mixed sizes fall back to cv2.resize per frame.  `loader(path) -> (H,W,3) uint8` may return
    frames of any size."""
    g = vit_model.geometry
    if loader is None:
        import cv2

        def loader(p):
            return cv2.cvtColor(cv2.imread(p), cv2.COLOR_BGR2RGB)
    embs, batch = [], []
    for i, p in enumerate(image_paths):
        batch.append(np.asarray(loader(p)))
        if len(batch) == batch_size or i == len(image_paths) - 1:
            if len({b.shape for b in batch}) > 1:
                import cv2
                batch = [b if b.shape[:2] == (g.image_h, g.image_w) else cv2.resize(b, (g.image_w, g.image_h))
                         for b in batch]
            x = np.array(batch)
            if x.dtype != np.uint8:
                x = np.clip(x, 0, 255).astype(np.uint8)
            out = vit_model.embed_frames(x, l2_normalise=False, interpolation="linear")
            embs.append(out.reshape(len(batch), -1))
            batch = []
    return np.vstack(embs) if embs else np.zeros((0, g.hidden), np.float32)
