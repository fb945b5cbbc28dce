"""Developer diagnostic: run every libvitk operator on the GPU against a torch fp32 reference and
print error statistics (never asserts, so one gpurun call reports on all kernels)."""
import ctypes as C
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import vit_research_b200 as v
from vit_research_b200 import _lib
from vit_research_b200._lib import ptr, stream_ptr

lib = v.load()
torch.backends.cuda.matmul.allow_tf32 = False
dev = torch.device("cuda:0")
g = torch.Generator(device="cpu").manual_seed(0)


def report(name, got, ref, tol):
    got, ref = got.float(), ref.float()
    err = (got - ref).abs()
    bad = (err > tol * (1 + ref.abs())).sum().item()
    print(f"[{name}] max_abs_err={err.max().item():.4e} mean_abs_err={err.mean().item():.4e} "
          f"ref_absmax={ref.abs().max().item():.3e} bad={bad}/{err.numel()} nan={torch.isnan(got).sum().item()}",
          flush=True)
    if bad:
        idx = torch.nonzero(err > tol * (1 + ref.abs()))
        print("   first bad idx:", idx[:5].tolist(), " last bad idx:", idx[-3:].tolist())
        rows = torch.unique(idx[:, 0])
        print("   bad rows: n=%d min=%d max=%d" % (rows.numel(), rows.min().item(), rows.max().item()),
              " bad cols: min=%d max=%d" % (idx[:, -1].min().item(), idx[:, -1].max().item()))
        i = idx[0]
        print("   sample got/ref:", got[tuple(i)].item(), ref[tuple(i)].item())
    return bad == 0


def gemm_case(M, N, K, epi):
    A = (torch.randn(M, K, generator=g) * 0.5).to(dev).bfloat16()
    W = (torch.randn(N, K, generator=g) * 0.05).to(dev).bfloat16()
    bias = torch.randn(N, generator=g).to(dev)
    acc = A.float() @ W.float().t()
    resid = torch.randn(M, N, generator=g).to(dev)
    table = torch.randn(7, N, generator=g).to(dev)
    bf16_out = epi in (0, 1, 2)
    out = torch.full((M, N), float("nan"), device=dev, dtype=torch.bfloat16 if bf16_out else torch.float32)
    if epi == 3:
        out.copy_(resid)
    rc = lib.vitk_op_gemm(ptr(A), ptr(W), ptr(out), M, N, K, epi, ptr(bias),
                          ptr(out) if epi == 3 else None, ptr(table), 7, stream_ptr())
    torch.cuda.synchronize()
    if rc:
        print("gemm rc", rc, lib.vitk_last_error(None))
        return False
    if epi == 0: ref = acc + bias
    elif epi == 1: ref = torch.nn.functional.gelu(acc + bias)
    elif epi == 2: ref = torch.nn.functional.gelu(acc + bias, approximate="tanh")
    elif epi == 3: ref = acc + bias + resid
    elif epi == 4: ref = acc + table[torch.arange(M, device=dev) % 7]
    else: ref = acc
    return report(f"gemm M={M} N={N} K={K} epi={epi}", out, ref, 2e-2 if bf16_out else 2e-3)


def ln_case(rows, d):
    x = (torch.randn(rows, d, generator=g) * 3 + 0.5).to(dev)
    gm = torch.randn(d, generator=g).to(dev); bt = torch.randn(d, generator=g).to(dev)
    y = torch.empty(rows, d, device=dev, dtype=torch.bfloat16)
    rc = lib.vitk_op_layernorm(ptr(x), ptr(gm), ptr(bt), ptr(y), rows, d, 1e-12, stream_ptr())
    torch.cuda.synchronize()
    ref = torch.nn.functional.layer_norm(x, (d,), gm, bt, 1e-12)
    return report(f"layernorm rows={rows} d={d} rc={rc}", y, ref, 1e-2)


def attn_case(frames, S, heads):
    D = heads * 64
    qkv = (torch.randn(frames, S, 3 * D, generator=g) * 1.0).to(dev).bfloat16()
    out = torch.full((frames, S, D), float("nan"), device=dev, dtype=torch.bfloat16)
    rc = lib.vitk_op_attention(ptr(qkv), ptr(out), frames, S, heads, stream_ptr())
    torch.cuda.synchronize()
    if rc:
        print("attn rc", rc, lib.vitk_last_error(None)); return False
    q, k, vv = qkv.float().split(D, dim=-1)
    sh = lambda t: t.view(frames, S, heads, 64).transpose(1, 2)
    p = torch.softmax(sh(q) @ sh(k).transpose(-1, -2) * 0.125, dim=-1)
    ref = (p @ sh(vv)).transpose(1, 2).reshape(frames, S, D)
    return report(f"attention frames={frames} S={S} heads={heads}", out, ref, 2e-2)


def patchify_case(n, fmt):
    C_, H, W, ps = 3, 224, 224, 16
    if fmt == 2:
        pix = torch.randint(0, 256, (n, H, W, C_), generator=g, dtype=torch.uint8).to(dev)
        x = ((pix.float() * (1 / 255.0)) - 0.5) / 0.5
        x = x.permute(0, 3, 1, 2)
    else:
        x = (torch.randn(n, C_, H, W, generator=g) * 0.5).clamp(-1, 1).to(dev)
        pix = x if fmt == 0 else x.bfloat16()
        x = pix.float()
    P = (H // ps) * (W // ps)
    out = torch.full((n * (P + 1), C_ * ps * ps), float("nan"), device=dev, dtype=torch.bfloat16)
    rc = lib.vitk_op_patchify(ptr(pix.contiguous()), fmt, ptr(out), n, C_, H, W, ps, stream_ptr())
    torch.cuda.synchronize()
    ref = torch.nn.functional.unfold(x, ps, stride=ps).transpose(1, 2)  # (n, P, C*ps*ps)
    ref = torch.cat([torch.zeros(n, 1, ref.shape[-1], device=dev), ref], 1).reshape(n * (P + 1), -1)
    return report(f"patchify n={n} fmt={fmt} rc={rc}", out, ref.bfloat16(), 1e-6)


def fused_ln_case(M, N, K, epi):
    """LayerNorm folded into the GEMM epilogue vs torch LN -> linear (-> gelu)."""
    x = (torch.randn(M, K, generator=g) * 2 + 0.3).to(dev)
    gm = (torch.randn(K, generator=g) * 0.3 + 1).to(dev); bt = (torch.randn(K, generator=g) * 0.2).to(dev)
    W = (torch.randn(N, K, generator=g) * 0.05).to(dev); b = torch.randn(N, generator=g).to(dev)
    parts = 2 * K // 256
    xb = torch.empty(M, K, device=dev, dtype=torch.bfloat16)
    stats = torch.full((M, parts, 2), float("nan"), device=dev)
    rc = lib.vitk_op_rowstats(ptr(x), ptr(xb), ptr(stats), M, K, parts, stream_ptr())
    Wf = (W * gm).bfloat16()
    c1 = Wf.float().sum(1).contiguous()
    c2 = (b + W @ bt).contiguous()
    out = torch.full((M, N), float("nan"), device=dev, dtype=torch.bfloat16)
    rc |= lib.vitk_op_gemm_fused(ptr(xb), ptr(Wf), ptr(out), M, N, K, epi, ptr(c2), ptr(c1), ptr(stats), parts,
                                 1e-12, None, None, None, stream_ptr())
    torch.cuda.synchronize()
    if rc:
        print("fused ln rc", rc, lib.vitk_last_error(None)); return False
    ref = torch.nn.functional.layer_norm(x, (K,), gm, bt, 1e-12) @ W.t() + b
    if epi == 1: ref = torch.nn.functional.gelu(ref)
    ok = report(f"fused LN gemm M={M} N={N} K={K} epi={epi}", out, ref, 3e-2)
    sref = torch.stack([x.sum(1), (x * x).sum(1)], 1)
    ok &= report("   rowstats sums", stats[:, 0, :], sref, 1e-4)
    ok &= report("   rowstats xb", xb, x.bfloat16(), 1e-6)
    return ok


def fused_resid_case(M, N, K, inplace=True):
    A = (torch.randn(M, K, generator=g) * 0.5).to(dev).bfloat16()
    W = (torch.randn(N, K, generator=g) * 0.05).to(dev).bfloat16()
    bias = torch.randn(N, generator=g).to(dev)
    resid = torch.randn(M, N, generator=g).to(dev)
    ref = A.float() @ W.float().t() + bias + resid
    out = resid.clone() if inplace else torch.full((M, N), float("nan"), device=dev)
    parts = 2 * ((N + 255) // 256)
    xb = torch.full((M, N), float("nan"), device=dev, dtype=torch.bfloat16)
    st = torch.full((M, parts, 2), float("nan"), device=dev)
    rc = lib.vitk_op_gemm_fused(ptr(A), ptr(W), ptr(out), M, N, K, 3, ptr(bias), None, None, 0, 0.0,
                                ptr(out) if inplace else ptr(resid), ptr(xb), ptr(st), stream_ptr())
    torch.cuda.synchronize()
    if rc:
        print("fused resid rc", rc, lib.vitk_last_error(None)); return False
    ok = report(f"fused resid gemm M={M} N={N} K={K} inplace={inplace}", out, ref, 2e-3)
    ok &= report("   xb copy", xb, out.bfloat16(), 1e-6)
    cols = torch.arange(N, device=dev) // 128
    sref = torch.zeros(M, parts, 2, device=dev)
    sref[:, :, 0].index_add_(1, cols, out)
    sref[:, :, 1].index_add_(1, cols, out * out)
    ok &= report("   partial stats", st, sref, 1e-4)
    return ok


if __name__ == "__main__":
    print(torch.cuda.get_device_name(0), lib.vitk_version(), "device_ok", lib.vitk_device_ok(0), flush=True)
    which = sys.argv[1:] or ["gemm", "fused", "ln", "attn", "patch"]
    ok = True
    if "gemm" in which:
        ok &= gemm_case(128, 256, 64, 5)
        ok &= gemm_case(128, 256, 128, 5)
        ok &= gemm_case(256, 512, 768, 5)
        for epi in (0, 1, 2, 3, 4, 5):
            ok &= gemm_case(300, 768, 768, epi)
        ok &= gemm_case(197 * 64, 2304, 768, 0)
        ok &= gemm_case(197 * 64, 768, 3072, 3)
        ok &= gemm_case(256, 1472, 768, 5)
    if "fused" in which:
        ok &= fused_ln_case(300, 768, 768, 0)
        ok &= fused_ln_case(197 * 16, 3072, 768, 1)
        ok &= fused_ln_case(200, 2048, 1024, 0)
        ok &= fused_resid_case(300, 768, 768, True)
        ok &= fused_resid_case(300, 768, 768, False)
        ok &= fused_resid_case(197 * 16, 768, 3072, True)
        ok &= fused_resid_case(333, 1024, 1024, True)
    if "ln" in which:
        ok &= ln_case(1000, 768); ok &= ln_case(77, 1024)
    if "patch" in which:
        for fmt in (0, 1, 2): ok &= patchify_case(3, fmt)
    if "attn" in which:
        ok &= attn_case(1, 197, 1)
        ok &= attn_case(2, 197, 12)
        ok &= attn_case(3, 64, 2)
        ok &= attn_case(2, 128, 2)
        ok &= attn_case(2, 256, 2)
    print("ALL OK" if ok else "SOME FAILED", flush=True)
