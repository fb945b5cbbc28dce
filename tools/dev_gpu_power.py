"""Developer diagnostic: sustained (1 s) loops of each hot kernel with SM clock + board power
sampled during the loop -> TFLOP/s, W and TFLOP/J per kernel (the forward is power-capped)."""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import vit_research_b200 as v
import bench
from vit_research_b200._lib import ptr, stream_ptr

lib = v.load()
n = 256
M, D, F, S, H = n * 197, 768, 3072, 197, 12
dev, bf = "cuda", torch.bfloat16
x = torch.randn(M, D, device=dev).to(bf); h = (torch.randn(M, F, device=dev) * 0.1).to(bf)
qkv = torch.randn(M, 3 * D, device=dev).to(bf)
out_qkv = torch.empty(M, 3 * D, device=dev, dtype=bf); out_h = torch.empty(M, F, device=dev, dtype=bf)
resid = torch.randn(M, D, device=dev); att_out = torch.empty(M, D, device=dev, dtype=bf)
w_qkv = (torch.randn(3 * D, D, device=dev) * 0.02).to(bf); w_o = (torch.randn(D, D, device=dev) * 0.02).to(bf)
w_1 = (torch.randn(F, D, device=dev) * 0.02).to(bf); w_2 = (torch.randn(D, F, device=dev) * 0.02).to(bf)
b3, b1, bF = torch.zeros(3 * D, device=dev), torch.zeros(D, device=dev), torch.zeros(F, device=dev)
parts = 2 * D // 256
xb = torch.empty(M, D, device=dev, dtype=bf); stats = torch.empty(M, parts, 2, device=dev); stats2 = torch.empty_like(stats)
lib.vitk_op_rowstats(ptr(resid), ptr(xb), ptr(stats), M, D, parts, stream_ptr())
st = stream_ptr(); fused = lib.vitk_op_gemm_fused
ops = {
    "gemm_qkv": (lambda: fused(ptr(xb), ptr(w_qkv), ptr(out_qkv), M, 3 * D, D, 0, ptr(b3), ptr(b3), ptr(stats), parts, 1e-12, None, None, None, st), 2.0 * M * 3 * D * D),
    "gemm_proj_resid": (lambda: fused(ptr(x), ptr(w_o), ptr(resid), M, D, D, 3, ptr(b1), None, None, 0, 0.0, ptr(resid), ptr(xb), ptr(stats2), st), 2.0 * M * D * D),
    "gemm_fc1_gelu": (lambda: fused(ptr(xb), ptr(w_1), ptr(out_h), M, F, D, 1, ptr(bF), ptr(bF), ptr(stats), parts, 1e-12, None, None, None, st), 2.0 * M * F * D),
    "gemm_fc2_resid": (lambda: fused(ptr(h), ptr(w_2), ptr(resid), M, D, F, 3, ptr(b1), None, None, 0, 0.0, ptr(resid), ptr(xb), ptr(stats2), st), 2.0 * M * D * F),
    "attention": (lambda: lib.vitk_op_attention(ptr(qkv), ptr(att_out), n, S, H, st), 4.0 * n * H * S * S * 64),
    "cublas_fc1": (lambda: torch.matmul(xb, w_1.t(), out=out_h), 2.0 * M * F * D),
    "cublas_fc2": (lambda: torch.matmul(h, w_2.t()), 2.0 * M * D * F),
}
sel = sys.argv[1:] or list(ops)
for name in sel:
    fn, flops = ops[name]
    for _ in range(20): fn()
    torch.cuda.synchronize()
    # ~3 s of back-to-back launches; clocks / power taken from the last two thirds of the samples
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t_end = time.time() + 3.0
    reps = 0
    with bench.ClockSampler(0) as clk:
        while time.time() < t_end:
            if reps == 2000: e0.record()
            for _ in range(200): fn()
            reps += 200
            if reps % 2000 == 0: torch.cuda.synchronize()
        e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / max(reps - 2000, 1)
    k = len(clk.samples) // 3
    mhz = sorted(clk.samples[k:])[len(clk.samples[k:]) // 2]
    pw = sorted(clk.power[k:])[len(clk.power[k:]) // 2]
    tf = flops / ms / 1e9
    print(f"{name:16s} {ms*1e3:7.1f} us  {tf:7.1f} TFLOP/s  {mhz} MHz  {pw:6.1f} W  {tf/max(pw,1):.3f} TFLOP/J  "
          f"{ms*pw:.1f} mJ/launch  reasons={sorted(clk.reasons)}", flush=True)
    time.sleep(3.0)
