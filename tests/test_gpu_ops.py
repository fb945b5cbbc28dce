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
                                            (2, 128, 2), (2, 129, 3), (2, 256, 2), (1, 250, 16)])
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
        x = (((pix.float() * (1 / 255.0)) - 0.5) / 0.5).permute(0, 3, 1, 2)
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
    assert lib.vitk_op_attention(_p(a), _p(o), 1, 300, 1, _s()) != 0
