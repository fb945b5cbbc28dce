"""TEST INFRASTRUCTURE ONLY — CPU fp32 restatement of the reference's ChunkEncoder.

Follows nba_proj/models/chunk_encoder.py:9-97: learned CLS + T frame tokens (:72-77), + positional
embedding (:80), `num_layers` Model Garden `nn_blocks.TransformerEncoderBlock(norm_first=True,
inner_activation=activations.gelu, norm_epsilon=1e-6)` (:17-31), LayerNormalization(1e-6) (:35, :83),
CLS row (:86), class_head Dense(256, relu) -> Dropout -> Dense(1) (:37-41, :89).

PARITY UNPINNED: tensorflow / tf-models-official are absent from this image and the reference pins
no version, so the block is restated from its published semantics: pre-LN; Keras MultiHeadAttention
(per-head kernels [D,H,dh], query scaled by dh^-1/2, softmax over keys, output kernel [H,dh,D]);
x = x + attn; y = LN(x); x + Dense(gelu_tanh(Dense(y))) with `official.modeling.activations.gelu` =
the tanh approximation; dropout and stochastic depth are identities at inference.
Only tests/ may import this module.
"""
import math

import numpy as np
import torch


def random_weights(seed=0, hidden=768, layers=3, heads=8, max_frames=24, randomise_ln=True):
    """Keras-layout weights: Dense kernels (in, out), attention kernels (D,H,dh) / (H,dh,D);
    glorot-uniform kernels, N(0, 0.02) CLS / positions (chunk_encoder.py:43-55)."""
    g = torch.Generator().manual_seed(seed)
    dh = hidden // heads

    def glorot(*shape, fan_in, fan_out):
        lim = math.sqrt(6.0 / (fan_in + fan_out))
        return ((torch.rand(shape, generator=g) * 2 - 1) * lim).numpy()

    def ln():
        if randomise_ln:
            return ((1 + 0.1 * torch.randn(hidden, generator=g)).numpy(), (0.1 * torch.randn(hidden, generator=g)).numpy())
        return np.ones(hidden, np.float32), np.zeros(hidden, np.float32)

    def small(n):
        return (0.02 * torch.randn(n, generator=g)).numpy()

    w = {"cls_token": (0.02 * torch.randn(1, 1, hidden, generator=g)).numpy(),
         "pos_embedding": (0.02 * torch.randn(1, 1 + max_frames, hidden, generator=g)).numpy()}
    for i in range(layers):
        p = f"block{i}/"
        w[p + "ln1_g"], w[p + "ln1_b"] = ln()
        for n in "qkv":
            w[p + "w" + n] = glorot(hidden, heads, dh, fan_in=hidden, fan_out=hidden)
            w[p + "b" + n] = small(hidden).reshape(heads, dh)
        w[p + "wo"] = glorot(heads, dh, hidden, fan_in=hidden, fan_out=hidden)
        w[p + "bo"] = small(hidden)
        w[p + "ln2_g"], w[p + "ln2_b"] = ln()
        w[p + "w1"] = glorot(hidden, 4 * hidden, fan_in=hidden, fan_out=4 * hidden)
        w[p + "b1"] = small(4 * hidden)
        w[p + "w2"] = glorot(4 * hidden, hidden, fan_in=4 * hidden, fan_out=hidden)
        w[p + "b2"] = small(hidden)
    w["norm_g"], w["norm_b"] = ln()
    w["head_w1"] = glorot(hidden, 256, fan_in=hidden, fan_out=256)
    w["head_b1"] = small(256)
    w["head_w2"] = glorot(256, 1, fan_in=256, fan_out=1)
    w["head_b2"] = small(1)
    return {k: np.asarray(v, np.float32) for k, v in w.items()}


def _ln(x, g, b, eps=1e-6):
    m = x.mean(-1, keepdim=True)
    v = ((x - m) ** 2).mean(-1, keepdim=True)
    return (x - m) / torch.sqrt(v + eps) * g + b


def forward(weights, frame_embeddings, layers=3, heads=8):
    """(B,T,D) -> chunk_emb (B,D), class_logit (B,1), attention scores [layers x (B,H,S,S)]."""
    w = {k: torch.from_numpy(np.asarray(v, np.float32)) for k, v in weights.items()}
    f = torch.from_numpy(np.asarray(frame_embeddings, np.float32))
    B, T, D = f.shape
    dh = D // heads
    x = torch.cat([w["cls_token"].expand(B, 1, D), f], 1) + w["pos_embedding"][:, : T + 1]
    scores = []
    for i in range(layers):
        p = f"block{i}/"
        y = _ln(x, w[p + "ln1_g"], w[p + "ln1_b"])
        q = torch.einsum("bsd,dhe->bhse", y, w[p + "wq"]) + w[p + "bq"][None, :, None, :]
        k = torch.einsum("bsd,dhe->bhse", y, w[p + "wk"]) + w[p + "bk"][None, :, None, :]
        v = torch.einsum("bsd,dhe->bhse", y, w[p + "wv"]) + w[p + "bv"][None, :, None, :]
        a = torch.softmax(torch.einsum("bhse,bhte->bhst", q * dh ** -0.5, k), -1)
        scores.append(a)
        o = torch.einsum("bhst,bhte->bshe", a, v)
        x = x + torch.einsum("bshe,hed->bsd", o, w[p + "wo"]) + w[p + "bo"]
        y = _ln(x, w[p + "ln2_g"], w[p + "ln2_b"])
        h = torch.nn.functional.gelu(y @ w[p + "w1"] + w[p + "b1"], approximate="tanh")
        x = x + h @ w[p + "w2"] + w[p + "b2"]
    emb = _ln(x, w["norm_g"], w["norm_b"])[:, 0]
    logit = torch.relu(emb @ w["head_w1"] + w["head_b1"]) @ w["head_w2"] + w["head_b2"]
    return emb.numpy(), logit.numpy(), [s.numpy() for s in scores]
