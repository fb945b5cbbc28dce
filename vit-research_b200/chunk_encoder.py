"""ChunkEncoder on the B200 op layer: the temporal transformer that turns the
T frame embeddings of a chunk into one chunk embedding + proxy logit.

Mirrors nba_proj/models/chunk_encoder.py:9-97 — `ChunkEncoder(hidden_size, num_layers, num_heads,
max_frames)`, `model(frame_embeddings, return_attention=False) -> (chunk_emb (B,768),
class_logit (B,1)[, attention_scores_all])`.  The linear layers run on the tcgen05 GEMM
(vitk_op_gemm: bias / tanh-GELU / residual epilogues), the 8 x 96 attention over <= 25 tokens, the
fp32 LayerNorm of the CLS rows and the class head on the kernels of csrc/chunk.cuh.  torch is used
for device buffers and the [CLS, frames] + positions assembly only.

Weights: a dict of Keras-layout arrays (Dense kernels (in,out); attention kernels (D,H,dh) and
(H,dh,D)) under the names of oracle/chunk_encoder_ref.random_weights; reading the reference's TF
checkpoints needs TensorFlow, which this image does not have.  TF semantics are therefore
unpinned; parity is against the restated block."""
from __future__ import annotations

import numpy as np

from . import _lib
from ._lib import VitkError, ptr, stream_ptr

EPI_BIAS_BF16, EPI_BIAS_GELUT_BF16, EPI_BIAS_RESID_F32 = 0, 2, 3


class ChunkEncoder:
    def __init__(self, weights, hidden_size=768, num_layers=3, num_heads=8, max_frames=24, device=0,
                 norm_epsilon=1e-6):
        import torch

        if not torch.cuda.is_available():
            raise VitkError("ChunkEncoder needs an sm_100 device: there is no CPU fallback")
        if hidden_size not in (768, 1024) or hidden_size % num_heads or (hidden_size // num_heads) % 32:
            raise VitkError("ChunkEncoder: hidden_size must be 768 or 1024 and the head width a multiple of 32")
        self.lib = _lib.load()
        self.D, self.L, self.H, self.dh = hidden_size, num_layers, num_heads, hidden_size // num_heads
        self.max_tokens, self.eps = 1 + max_frames, float(norm_epsilon)
        self.device = torch.device("cuda", device)
        D = self.D

        def f32(a, *shape):
            return torch.as_tensor(np.ascontiguousarray(a, np.float32)).reshape(*shape).to(self.device).contiguous()

        def bf16_t(a_in_out):   # Keras (in, out) kernel -> the GEMM's [N, K] bf16 weight
            return torch.as_tensor(np.ascontiguousarray(a_in_out, np.float32)).t().contiguous().to(
                self.device, torch.bfloat16).contiguous()

        w = weights
        self.cls = f32(w["cls_token"], 1, 1, D)
        self.pos = f32(w["pos_embedding"], 1, self.max_tokens, D)
        self.blocks = []
        for i in range(num_layers):
            p = f"block{i}/"
            qkv = np.concatenate([np.asarray(w[p + "w" + n], np.float32).reshape(D, D) for n in "qkv"], 1)
            self.blocks.append({
                "ln1": (f32(w[p + "ln1_g"], D), f32(w[p + "ln1_b"], D)),
                "wqkv": bf16_t(qkv), "bqkv": f32(np.concatenate([np.asarray(w[p + "b" + n]).reshape(D) for n in "qkv"]), 3 * D),
                "wo": bf16_t(np.asarray(w[p + "wo"], np.float32).reshape(D, D)), "bo": f32(w[p + "bo"], D),
                "ln2": (f32(w[p + "ln2_g"], D), f32(w[p + "ln2_b"], D)),
                "w1": bf16_t(w[p + "w1"]), "b1": f32(w[p + "b1"], 4 * D),
                "w2": bf16_t(w[p + "w2"]), "b2": f32(w[p + "b2"], D)})
        self.norm = (f32(w["norm_g"], D), f32(w["norm_b"], D))
        self.head_hidden = int(np.asarray(w["head_b1"]).size)
        self.head = (f32(np.asarray(w["head_w1"], np.float32).T, self.head_hidden, D), f32(w["head_b1"], self.head_hidden),
                     f32(w["head_w2"], self.head_hidden), f32(w["head_b2"], 1))

    def _check(self, rc, what):
        if rc != 0:
            msg = self.lib.vitk_last_error(None)
            raise VitkError(f"{what} failed: {msg.decode() if msg else 'unknown error'}")

    def __call__(self, frame_embeddings, training=False, return_attention=False):
        """frame_embeddings (B,T,D) numpy or CUDA tensor -> (chunk_emb (B,D), class_logit (B,1)) as
        numpy float32 (CUDA tensors when the input was one), plus the per-block attention scores
        (B,H,T+1,T+1) when return_attention."""
        import torch

        if training:
            raise VitkError("ChunkEncoder: inference only (dropout / stochastic depth are identities)")
        on_dev = hasattr(frame_embeddings, "is_cuda")
        f = frame_embeddings if on_dev else torch.as_tensor(np.ascontiguousarray(frame_embeddings, np.float32))
        if f.dim() != 3:
            raise ValueError(f"Expected frame_embeddings to have shape (B, T, D), got {tuple(f.shape)}")
        B, T, D = f.shape
        if D != self.D or T + 1 > self.max_tokens:
            raise VitkError(f"ChunkEncoder: got (T={T}, D={D}) for max_frames={self.max_tokens - 1}, hidden={self.D}")
        f = f.to(self.device, torch.float32)
        S, M = T + 1, B * (T + 1)
        lib, bf = self.lib, torch.bfloat16
        empty = lambda *s, dt=torch.float32: torch.empty(s, device=self.device, dtype=dt)  # noqa: E731
        if B == 0:
            z = (empty(0, D), empty(0, 1))
            return z if on_dev else tuple(t.cpu().numpy() for t in z)
        with torch.cuda.device(self.device):
            st = stream_ptr(None)
            x = (torch.cat([self.cls.expand(B, 1, D), f], 1) + self.pos[:, :S]).reshape(M, D).contiguous()
            y, qkv, att = empty(M, D, dt=bf), empty(M, 3 * D, dt=bf), empty(M, D, dt=bf)
            h = empty(M, 4 * D, dt=bf)
            scores = []
            for blk in self.blocks:
                self._check(lib.vitk_op_layernorm(ptr(x), ptr(blk["ln1"][0]), ptr(blk["ln1"][1]), ptr(y), M, D,
                                                  self.eps, st), "vitk_op_layernorm")
                self._check(lib.vitk_op_gemm(ptr(y), ptr(blk["wqkv"]), ptr(qkv), M, 3 * D, D, EPI_BIAS_BF16,
                                             ptr(blk["bqkv"]), None, None, 1, st), "vitk_op_gemm")
                pr = empty(B, self.H, S, S) if return_attention else None
                self._check(lib.vitk_op_attention_small(ptr(qkv), ptr(att), ptr(pr), B, S, self.H, self.dh, st),
                            "vitk_op_attention_small")
                if pr is not None:
                    scores.append(pr)
                self._check(lib.vitk_op_gemm(ptr(att), ptr(blk["wo"]), ptr(x), M, D, D, EPI_BIAS_RESID_F32,
                                             ptr(blk["bo"]), ptr(x), None, 1, st), "vitk_op_gemm")
                self._check(lib.vitk_op_layernorm(ptr(x), ptr(blk["ln2"][0]), ptr(blk["ln2"][1]), ptr(y), M, D,
                                                  self.eps, st), "vitk_op_layernorm")
                self._check(lib.vitk_op_gemm(ptr(y), ptr(blk["w1"]), ptr(h), M, 4 * D, D, EPI_BIAS_GELUT_BF16,
                                             ptr(blk["b1"]), None, None, 1, st), "vitk_op_gemm")
                self._check(lib.vitk_op_gemm(ptr(h), ptr(blk["w2"]), ptr(x), M, D, 4 * D, EPI_BIAS_RESID_F32,
                                             ptr(blk["b2"]), ptr(x), None, 1, st), "vitk_op_gemm")
            emb, logit = empty(B, D), empty(B, 1)
            self._check(lib.vitk_op_layernorm_rows_f32(ptr(x), ptr(self.norm[0]), ptr(self.norm[1]), ptr(emb), B,
                                                       S * D, D, self.eps, st), "vitk_op_layernorm_rows_f32")
            self._check(lib.vitk_op_chunk_head(ptr(emb), ptr(self.head[0]), ptr(self.head[1]), ptr(self.head[2]),
                                               ptr(self.head[3]), ptr(logit), B, D, self.head_hidden, st),
                        "vitk_op_chunk_head")
            torch.cuda.current_stream().synchronize()
        out = (emb, logit) + ((scores,) if return_attention else ())
        if on_dev:
            return out
        res = (emb.cpu().numpy(), logit.cpu().numpy())
        return res + (([s.cpu().numpy() for s in scores],) if return_attention else ())
