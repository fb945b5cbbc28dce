"""GPU operator tests: each libvitk kernel, through the C ABI, against a plain PyTorch fp32
reference of the same op (tolerances = bf16 input/output rounding, written per test)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def lib():
    import vit_research_b200 as v
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    lib = v.load()
    assert lib.vitk_device_ok(0) == 1, "not an sm_100 device"
    torch.backends.cuda.matmul.allow_tf32 = False
    return lib


def _p(t):
    from vit_research_b200._lib import ptr
    return ptr(t)


def _s():
    from vit_research_b200._lib import stream_ptr
    return stream_ptr()


def _gemm_ref(acc, epi, bias, resid, table):
    M = acc.shape[0]
    if epi == 0: return acc + bias
    if epi == 1: return torch.nn.functional.gelu(acc + bias)
    if epi == 2: return torch.nn.functional.gelu(acc + bias, approximate="tanh")
    if epi == 3: return acc + bias + resid
    if epi == 4: return acc + table[torch.arange(M, device=acc.device) % table.shape[0]]
    return acc


@pytest.mark.parametrize("M,N,K,epi", [
    (128, 256, 64, 5), (1, 256, 64, 5), (300, 768, 768, 0), (300, 768, 768, 1), (300, 768, 768, 2),
    (300, 768, 768, 3), (197 * 3, 768, 768, 4), (129, 1472, 768, 5), (197 * 40, 2304, 768, 0),
    (197 * 40, 3072, 768, 1), (197 * 40, 768, 3072, 3), (777, 1024, 4096, 3), (50, 36, 128, 5),
])
def test_gemm_epilogues(lib, M, N, K, epi):
    g = torch.Generator().manual_seed(M * 7 + N + K + epi)
    dev = "cuda:0"
    A = (torch.randn(M, K, generator=g) * 0.5).to(dev).bfloat16()
    W = (torch.randn(N, K, generator=g) * 0.05).to(dev).bfloat16()
    bias = torch.randn(N, generator=g).to(dev)
    resid = torch.randn(M, N, generator=g).to(dev)
    table = torch.randn(7, N, generator=g).to(dev)
    bf16_out = epi in (0, 1, 2)
    out = torch.full((M, N), float("nan"), device=dev, dtype=torch.bfloat16 if bf16_out else torch.float32)
    if epi == 3:
        out.copy_(resid)  # in-place residual, as the forward uses it
    rc = lib.vitk_op_gemm(_p(A), _p(W), _p(out), M, N, K, epi, _p(bias), _p(out) if epi == 3 else None,
                          _p(table), 7, _s())
    torch.cuda.synchronize()
    assert rc == 0, lib.vitk_last_error(None)
    ref = _gemm_ref(A.float() @ W.float().t(), epi, bias, resid, table)
    if bf16_out:  # one bf16 rounding of the output: 2^-8 relative
        torch.testing.assert_close(out.float(), ref, rtol=8e-3, atol=8e-3)
    else:          # fp32 accumulate, different summation order only
        torch.testing.assert_close(out, ref, rtol=1e-4, atol=1e-4 * max(1.0, K / 768))


def test_gemm_is_linear_in_a(lib):
    """Size-independent property at the full C2 shape: gemm(A1+A2) == gemm(A1)+gemm(A2) when the
    inputs are exactly representable (small integers in bf16, fp32 accumulation is exact)."""
    dev = "cuda:0"
    M, N, K = 197 * 256, 768, 768
    g = torch.Generator().manual_seed(1)
    A1 = torch.randint(-3, 4, (M, K), generator=g).to(dev).bfloat16()
    A2 = torch.randint(-3, 4, (M, K), generator=g).to(dev).bfloat16()
    W = torch.randint(-2, 3, (N, K), generator=g).to(dev).bfloat16()
    outs = []
    for A in (A1, A2, A1 + A2):
        o = torch.empty(M, N, device=dev)
        assert lib.vitk_op_gemm(_p(A), _p(W), _p(o), M, N, K, 5, None, None, None, 1, _s()) == 0
        outs.append(o)
    torch.cuda.synchronize()
    assert torch.equal(outs[0] + outs[1], outs[2])
    # and exact against integer arithmetic on a row sample
    rows = torch.arange(0, M, 997, device=dev)
    assert torch.equal(outs[2][rows], (A1 + A2)[rows].float() @ W.float().t())


@pytest.mark.parametrize("rows,d", [(1000, 768), (77, 1024), (1, 768)])
def test_layernorm(lib, rows, d):
    g = torch.Generator().manual_seed(rows + d)
    x = (torch.randn(rows, d, generator=g) * 3 + 0.5).cuda()
    gm, bt = torch.randn(d, generator=g).cuda(), torch.randn(d, generator=g).cuda()
    y = torch.empty(rows, d, device="cuda", dtype=torch.bfloat16)
    assert lib.vitk_op_layernorm(_p(x), _p(gm), _p(bt), _p(y), rows, d, 1e-12, _s()) == 0
    torch.cuda.synchronize()
    ref = torch.nn.functional.layer_norm(x, (d,), gm, bt, 1e-12)
    torch.testing.assert_close(y.float(), ref, rtol=8e-3, atol=8e-3)  # bf16 output rounding


@pytest.mark.parametrize("frames,S,heads", [(1, 197, 1), (3, 197, 12), (2, 4, 12), (3, 64, 2),
                                            (2, 128, 2), (2, 129, 3), (2, 256, 2), (1, 250, 16),
                                            # wide form (keys in two halves, 512 TMEM columns): the
                                            # reference's TF-path geometry has 313 tokens
                                            (2, 257, 2), (3, 313, 12), (1, 320, 3), (2, 385, 2), (1, 448, 2)])
def test_attention(lib, frames, S, heads):
    D = heads * 64
    g = torch.Generator().manual_seed(frames * 1000 + S + heads)
    qkv = torch.randn(frames, S, 3 * D, generator=g).cuda().bfloat16()
    out = torch.full((frames, S, D), float("nan"), device="cuda", dtype=torch.bfloat16)
    rc = lib.vitk_op_attention(_p(qkv), _p(out), frames, S, heads, _s())
    torch.cuda.synchronize()
    assert rc == 0, lib.vitk_last_error(None)
    q, k, v = qkv.float().split(D, dim=-1)
    sh = lambda t: t.view(frames, S, heads, 64).transpose(1, 2)
    p = torch.softmax(sh(q) @ sh(k).transpose(-1, -2) * 0.125, dim=-1)
    ref = (p @ sh(v)).transpose(1, 2).reshape(frames, S, D)
    # bf16 P and bf16 output: ~2^-8 relative of values <= max|v|
    torch.testing.assert_close(out.float(), ref, rtol=1e-2, atol=1e-2)


@pytest.mark.parametrize("mult", [3.0, 8.0, 20.0])
def test_attention_large_logits_are_stable(lib, mult):
    """Row-max handling: logits of +-10 .. +-400 nats must neither overflow nor lose the row
    maximum.  The kernel's one-pass fast path shifts by the max of the first 32 scores; rows whose
    true maximum is > 100 binades above that are detected and redone with the exact two-pass
    form — mult = 8 and 20 exercise that path, mult = 3 the boundary between the two."""
    frames, S, heads = 2, 197, 2
    D = heads * 64
    g = torch.Generator().manual_seed(9)
    qkv = torch.randn(frames, S, 3 * D, generator=g)
    qkv[..., : 2 * D] *= mult
    qkv = qkv.cuda().bfloat16()
    out = torch.empty(frames, S, D, device="cuda", dtype=torch.bfloat16)
    assert lib.vitk_op_attention(_p(qkv), _p(out), frames, S, heads, _s()) == 0
    torch.cuda.synchronize()
    q, k, v = qkv.float().split(D, dim=-1)
    sh = lambda t: t.view(frames, S, heads, 64).transpose(1, 2)
    ref = (torch.softmax(sh(q) @ sh(k).transpose(-1, -2) * 0.125, -1) @ sh(v)).transpose(1, 2).reshape(frames, S, D)
    assert torch.isfinite(out.float()).all()
    torch.testing.assert_close(out.float(), ref, rtol=2e-2, atol=2e-2)


@pytest.mark.parametrize("fmt", [0, 1, 2])
def test_patchify_is_exact(lib, fmt):
    n, C_, H, W, ps = 3, 3, 224, 224, 16
    g = torch.Generator().manual_seed(fmt)
    if fmt == 2:
        pix = torch.randint(0, 256, (n, H, W, C_), generator=g, dtype=torch.uint8).cuda()
        x = ((pix.float() - 127.5) / 127.5).permute(0, 3, 1, 2)
    else:
        x = (torch.randn(n, C_, H, W, generator=g) * 0.5).clamp(-1, 1).cuda()
        pix = x if fmt == 0 else x.bfloat16()
        x = pix.float()
    P = (H // ps) * (W // ps)
    out = torch.full((n * (P + 1), C_ * ps * ps), float("nan"), device="cuda", dtype=torch.bfloat16)
    assert lib.vitk_op_patchify(_p(pix.contiguous()), fmt, _p(out), n, C_, H, W, ps, _s()) == 0
    torch.cuda.synchronize()
    ref = torch.nn.functional.unfold(x, ps, stride=ps).transpose(1, 2)
    ref = torch.cat([torch.zeros(n, 1, ref.shape[-1], device="cuda"), ref], 1).reshape(n * (P + 1), -1)
    assert torch.equal(out, ref.bfloat16())  # pure data movement + one rounding: bit exact


def test_bad_arguments_fail_loudly(lib):
    a = torch.zeros(128, 100, device="cuda", dtype=torch.bfloat16)
    o = torch.zeros(128, 256, device="cuda")
    assert lib.vitk_op_gemm(_p(a), _p(a), _p(o), 128, 256, 100, 5, None, None, None, 1, _s()) != 0
    assert b"K=100" in lib.vitk_last_error(None)
    assert lib.vitk_op_attention(_p(a), _p(o), 1, 500, 1, _s()) != 0   # S > 448


# ---------------------------------------------------------------------------------------------
# fused forms the forward actually runs
# ---------------------------------------------------------------------------------------------
@pytest.mark.parametrize("M,N,K,epi", [(300, 768, 768, 0), (197 * 16, 3072, 768, 1), (200, 2048, 1024, 0),
                                       (129, 2304, 768, 0), (1, 768, 768, 1)])
def test_layernorm_folded_gemm(lib, M, N, K, epi):
    """LN(x) @ W^T + b (-> gelu) computed as the GEMM of the RAW rows against gamma-scaled weights
    with the per-row (mean, rstd) and per-column (c1, c2) applied in the epilogue
    (HF modeling_vit.py:325-346) vs torch layer_norm -> linear in fp32."""
    g = torch.Generator().manual_seed(M + N + K + epi)
    dev = "cuda:0"
    x = (torch.randn(M, K, generator=g) * 2 + 0.3).to(dev)
    gm = (torch.randn(K, generator=g) * 0.3 + 1).to(dev)
    bt = (torch.randn(K, generator=g) * 0.2).to(dev)
    W = (torch.randn(N, K, generator=g) * 0.05).to(dev)
    b = torch.randn(N, generator=g).to(dev)
    parts = 2 * K // 256
    xb = torch.empty(M, K, device=dev, dtype=torch.bfloat16)
    stats = torch.full((M, parts, 2), float("nan"), device=dev)
    assert lib.vitk_op_rowstats(_p(x), _p(xb), _p(stats), M, K, parts, _s()) == 0
    Wf = (W * gm).bfloat16()
    c1 = Wf.float().sum(1).contiguous()
    c2 = (b + W @ bt).contiguous()
    out = torch.full((M, N), float("nan"), device=dev, dtype=torch.bfloat16)
    rc = lib.vitk_op_gemm_fused(_p(xb), _p(Wf), _p(out), M, N, K, epi, _p(c2), _p(c1), _p(stats), parts,
                                1e-12, None, None, None, _s())
    torch.cuda.synchronize()
    assert rc == 0, lib.vitk_last_error(None)
    ref = torch.nn.functional.layer_norm(x, (K,), gm, bt, 1e-12) @ W.t() + b
    if epi == 1:
        ref = torch.nn.functional.gelu(ref)
    # bf16 rounding of the raw rows, of gamma*W and of the output
    torch.testing.assert_close(out.float(), ref, rtol=2e-2, atol=3e-2)
    assert torch.equal(xb, x.bfloat16())
    sref = torch.stack([x.sum(1), (x * x).sum(1)], 1)
    torch.testing.assert_close(stats[:, 0, :], sref, rtol=1e-5, atol=1e-3)
    assert (stats[:, 1:, :] == 0).all()


@pytest.mark.parametrize("M,N,K,inplace", [(300, 768, 768, True), (300, 768, 768, False),
                                           (197 * 16, 768, 3072, True), (333, 1024, 1024, True),
                                           (1, 768, 768, True)])
def test_residual_gemm_emits_copy_and_row_sums(lib, M, N, K, inplace):
    """out = acc + bias + resid (TMA-loaded residual, in place or not), bf16 copy == bf16(out),
    per-row partial (sum, sumsq) per 128 columns of out (HF modeling_vit.py:337,343-346)."""
    g = torch.Generator().manual_seed(M * 3 + N + K)
    dev = "cuda:0"
    A = (torch.randn(M, K, generator=g) * 0.5).to(dev).bfloat16()
    W = (torch.randn(N, K, generator=g) * 0.05).to(dev).bfloat16()
    bias = torch.randn(N, generator=g).to(dev)
    resid = torch.randn(M, N, generator=g).to(dev)
    ref = A.float() @ W.float().t() + bias + resid
    out = resid.clone() if inplace else torch.full((M, N), float("nan"), device=dev)
    parts = 2 * ((N + 255) // 256)
    xb = torch.full((M, N), float("nan"), device=dev, dtype=torch.bfloat16)
    st = torch.full((M, parts, 2), float("nan"), device=dev)
    rc = lib.vitk_op_gemm_fused(_p(A), _p(W), _p(out), M, N, K, 3, _p(bias), None, None, 0, 0.0,
                                _p(out) if inplace else _p(resid), _p(xb), _p(st), _s())
    torch.cuda.synchronize()
    assert rc == 0, lib.vitk_last_error(None)
    torch.testing.assert_close(out, ref, rtol=1e-4, atol=1e-4 * max(1.0, K / 768))
    assert torch.equal(xb, out.bfloat16())
    cols = torch.arange(N, device=dev) // 128
    sref = torch.zeros(M, parts, 2, device=dev)
    sref[:, :, 0].index_add_(1, cols, out)
    sref[:, :, 1].index_add_(1, cols, out * out)
    torch.testing.assert_close(st, sref, rtol=1e-5, atol=1e-3)


def test_gelu_polynomial_matches_erf_form(lib):
    """The epilogue exponent polynomial (packed-fp32 degree-5 form) vs torch exact erf GELU over [-9, 9]."""
    dev = "cuda:0"
    M, N, K = 256, 4096, 64
    xs = torch.linspace(-9, 9, M * N, device=dev).reshape(M, N)
    A = torch.zeros(M, K, device=dev, dtype=torch.bfloat16)
    W = torch.zeros(N, K, device=dev, dtype=torch.bfloat16)
    # acc == 0, so out = gelu(bias column) is not what we want; drive x through the LN-free bias:
    # use K=64 identity-free trick: out[m,n] = gelu(acc + bias[n]) with acc = A[m,:]·W[n,:]
    A[:, 0] = 1.0
    col = torch.linspace(-9, 9, N, device=dev)
    W[:, 0] = col.bfloat16()
    bias = (col - W[:, 0].float()).contiguous()  # acc + bias == col exactly in fp32
    out = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
    assert lib.vitk_op_gemm(_p(A), _p(W), _p(out), M, N, K, 1, _p(bias), None, None, 1, _s()) == 0
    torch.cuda.synchronize()
    ref = torch.nn.functional.gelu(col.double()).float()
    got = out[0].float()
    assert (got - ref.bfloat16().float()).abs().max() <= ref.abs().max() * 2 ** -8
    assert (out == out[0:1]).all() and xs.numel() == M * N


def test_streaming_api_matches_synchronous_call(lib):
    """vitk_stream_submit/wait (two batches in flight) returns exactly what the synchronous
    host call returns for the same frames, in order, including a ragged last batch."""
    import vit_research_b200 as v
    geom = v.ViTGeometry(image_h=224, image_w=224, patch=16, channels=3, hidden=768, heads=12, mlp=3072,
                         layers=2, ln_eps=1e-12, gelu_tanh=False)
    emb = v.ViTEmbedder(geom, v.random_init_state_dict(geom, seed=3), device=0, max_batch=16)
    g = torch.Generator().manual_seed(11)
    bank = torch.nn.functional.normalize(torch.randn(300, 768, generator=g), dim=1).numpy()
    labels = (np.arange(300) % 3).astype(np.int32)
    index = v.BankIndex(emb, bank, labels, 3, metric="cosine", normalise_rows=True)
    batches = [torch.randint(0, 256, (n, 224, 224, 3), generator=g, dtype=torch.uint8).numpy()
               for n in (16, 16, 16, 5)]
    want = [emb.embed_and_vote(b, index, 7) for b in batches]
    got = [tuple(a.copy() for a in r) for r in emb.embed_and_vote_stream(iter(batches), index, 7)]
    assert len(got) == len(want)
    for w, r in zip(want, got):
        for a, b in zip(w, r):
            assert np.array_equal(a, b)
    only = [r[0].copy() for r in emb.embed_and_vote_stream(iter(batches[:2]))]
    assert np.array_equal(only[0], want[0][0]) and np.array_equal(only[1], want[1][0])
    emb.close()
