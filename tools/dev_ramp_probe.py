"""Developer probe: per-launch duration of the fc1 GEMM for 60 consecutive launches after a 2 s
idle (does the clock ramp / power cap show inside a short timing loop?)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import vit_research_b200 as v
from vit_research_b200._lib import ptr, stream_ptr
lib = v.load()
M, D, F = 256 * 197, 768, 3072
bf = torch.bfloat16
st = stream_ptr()
xb = torch.randn(M, D, device="cuda").to(bf)
w = (torch.randn(F, D, device="cuda") * 0.02).to(bf)
out = torch.empty(M, F, device="cuda", dtype=bf)
bF, cF = torch.zeros(F, device="cuda"), torch.zeros(F, device="cuda")
stats = torch.zeros(M, 6, 2, device="cuda"); stats[:, 0, 1] = 768.0
run = lambda: lib.vitk_op_gemm_fused(ptr(xb), ptr(w), ptr(out), M, F, D, 1, ptr(bF), ptr(cF), ptr(stats), 6,
                                     1e-12, None, None, None, st)
for idle in (2.0, 0.0, 2.0):
    torch.cuda.synchronize(); time.sleep(idle)
    n = 60
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(n + 1)]
    ev[0].record()
    for i in range(n):
        run(); ev[i + 1].record()
    torch.cuda.synchronize()
    t = [round(ev[i].elapsed_time(ev[i + 1]) * 1000) for i in range(n)]
    print(f"idle {idle}s: us/launch {t}")
