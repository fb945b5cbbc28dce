"""GPU parity of the ChunkEncoder pieces (models/chunk_encoder.py:9-97) against
the fp32 CPU restatement in oracle/chunk_encoder_ref.py (TF semantics unpinned: TensorFlow is not in
this image).  Tolerances: bf16 GEMM operands -> cosine >= 0.999 on the chunk embedding, 2e-2 abs on
the logit and 1e-2 abs on the attention probabilities (q, k are bf16 GEMM outputs); the three small kernels alone are fp32 and
are held to 1e-5."""
import numpy as np
import pytest
import torch

from oracle import chunk_encoder_ref as ref

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def v():
    import vit_research_b200 as v
    assert torch.cuda.is_available()
    v.load()
    return v


def _p(t):
    from vit_research_b200._lib import ptr
    return ptr(t)


def _s():
    from vit_research_b200._lib import stream_ptr
    return stream_ptr(None)


@pytest.mark.parametrize("n_seq,S,heads,dh", [(3, 25, 8, 96), (1, 1, 2, 32), (5, 64, 4, 128), (2, 33, 3, 64), (7, 9, 8, 96)])
def test_small_attention_matches_fp32(v, n_seq, S, heads, dh):
    lib = v.load()
    g = torch.Generator().manual_seed(S * 7 + dh)
    D = heads * dh
    qkv = (torch.randn(n_seq, S, 3 * D, generator=g) * 0.7).bfloat16()
    q, k, val = [qkv[..., i * D:(i + 1) * D].float().reshape(n_seq, S, heads, dh).permute(0, 2, 1, 3) for i in range(3)]
    pw = torch.softmax((q * dh ** -0.5) @ k.transpose(-1, -2), -1)
    want = (pw @ val).permute(0, 2, 1, 3).reshape(n_seq, S, D)
    out = torch.empty(n_seq, S, D, device="cuda", dtype=torch.bfloat16)
    probs = torch.empty(n_seq, heads, S, S, device="cuda")
    assert lib.vitk_op_attention_small(_p(qkv.cuda()), _p(out), _p(probs), n_seq, S, heads, dh, _s()) == 0
    torch.cuda.synchronize()
    torch.testing.assert_close(probs.cpu(), pw, atol=1e-5, rtol=1e-4)
    torch.testing.assert_close(out.float().cpu(), want, atol=2e-2, rtol=1e-2)   # bf16 output rounding
    out2 = torch.empty_like(out)
    assert lib.vitk_op_attention_small(_p(qkv.cuda()), _p(out2), None, n_seq, S, heads, dh, _s()) == 0
    torch.cuda.synchronize()
    assert torch.equal(out, out2)
    assert lib.vitk_op_attention_small(_p(qkv.cuda()), _p(out), None, n_seq, 65, heads, dh, _s()) != 0
    assert lib.vitk_op_attention_small(_p(qkv.cuda()), _p(out), None, n_seq, S, heads, 48, _s()) != 0


def test_strided_layernorm_and_head_match_fp32(v):
    lib = v.load()
    g = torch.Generator().manual_seed(5)
    B, S, D, hid = 9, 13, 768, 256
    x = torch.randn(B, S, D, generator=g) * 2 + 0.5
    gam, bet = 1 + 0.1 * torch.randn(D, generator=g), 0.1 * torch.randn(D, generator=g)
    want = torch.nn.functional.layer_norm(x[:, 0], (D,), gam, bet, 1e-6)
    y = torch.empty(B, D, device="cuda")
    assert lib.vitk_op_layernorm_rows_f32(_p(x.cuda()), _p(gam.cuda()), _p(bet.cuda()), _p(y), B, S * D, D,
                                          1e-6, _s()) == 0
    torch.cuda.synchronize()
    torch.testing.assert_close(y.cpu(), want, atol=1e-5, rtol=1e-5)
    w1, b1 = torch.randn(hid, D, generator=g) * 0.05, torch.randn(hid, generator=g) * 0.1
    w2, b2 = torch.randn(hid, generator=g) * 0.1, torch.randn(1, generator=g)
    logit = torch.empty(B, device="cuda")
    assert lib.vitk_op_chunk_head(_p(y), _p(w1.cuda()), _p(b1.cuda()), _p(w2.cuda()), _p(b2.cuda()), _p(logit), B,
                                  D, hid, _s()) == 0
    torch.cuda.synchronize()
    wl = torch.relu(want @ w1.t() + b1) @ w2 + b2
    torch.testing.assert_close(logit.cpu(), wl, atol=1e-4, rtol=1e-4)


@pytest.mark.parametrize("B,T,layers", [(6, 8, 3), (33, 12, 4), (1, 24, 3), (300, 8, 3)])
def test_chunk_encoder_matches_oracle(v, B, T, layers):
    w = ref.random_weights(seed=layers, layers=layers)
    rng = np.random.default_rng(B + T)
    frames = rng.standard_normal((B, T, 768)).astype(np.float32)
    frames /= np.linalg.norm(frames, axis=-1, keepdims=True)      # frame embeddings are unit rows
    frames *= 27.7                                                 # TF pre_logits scale (un-normalised)
    enc = v.ChunkEncoder(w, num_layers=layers)
    emb, logit, scores = enc(frames, return_attention=True)
    wemb, wlogit, wscores = ref.forward(w, frames, layers=layers)
    assert emb.shape == (B, 768) and logit.shape == (B, 1) and len(scores) == layers
    cos = (emb * wemb).sum(1) / (np.linalg.norm(emb, axis=1) * np.linalg.norm(wemb, axis=1))
    assert cos.min() >= 0.999, cos.min()
    assert np.abs(logit - wlogit).max() <= 2e-2 * max(1.0, np.abs(wlogit).max())
    for a, b in zip(scores, wscores):
        assert a.shape == (B, 8, T + 1, T + 1)
        assert np.abs(a - b).max() <= 1e-2
        np.testing.assert_allclose(a.sum(-1), 1.0, atol=1e-5)
    # device tensors in -> device tensors out, same numbers; no attention scores unless asked
    e2, l2 = enc(torch.from_numpy(frames).cuda())
    assert e2.is_cuda and np.array_equal(e2.cpu().numpy(), emb) and np.array_equal(l2.cpu().numpy(), logit)
    # each chunk is independent of the batch it sits in
    e1, _ = enc(frames[:1])
    assert np.array_equal(e1, emb[:1])


def test_chunk_encoder_argument_errors(v):
    w = ref.random_weights(seed=0)
    enc = v.ChunkEncoder(w)
    with pytest.raises(ValueError):
        enc(np.zeros((4, 768), np.float32))
    with pytest.raises(v.VitkError):
        enc(np.zeros((2, 25, 768), np.float32))          # more frames than max_frames
    with pytest.raises(v.VitkError):
        enc(np.zeros((2, 8, 512), np.float32))
    e, l = enc(np.zeros((0, 8, 768), np.float32))
    assert e.shape == (0, 768) and l.shape == (0, 1)
