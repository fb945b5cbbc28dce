"""Developer probe: fc1-shaped GEMM with and without the GELU / LayerNorm-fold epilogue work."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import vit_research_b200 as v
from vit_research_b200._lib import ptr, stream_ptr
lib = v.load()
M, D, F = 256 * 197, 768, 3072
bf = torch.bfloat16
xb = torch.randn(M, D, device="cuda").to(bf)
w = (torch.randn(F, D, device="cuda") * 0.02).to(bf)
out = torch.empty(M, F, device="cuda", dtype=bf)
bF, cF = torch.zeros(F, device="cuda"), torch.zeros(F, device="cuda")
parts = 6
stats = torch.zeros(M, parts, 2, device="cuda"); stats[:, 0, 1] = 768.0
st = stream_ptr()
def run(epi, fold):
    return lib.vitk_op_gemm_fused(ptr(xb), ptr(w), ptr(out), M, F, D, epi, ptr(bF), ptr(cF) if fold else None,
                                  ptr(stats) if fold else None, parts if fold else 0, 1e-12, None, None, None, st)
for name, epi, fold in [("bias only", 0, False), ("ln fold", 0, True), ("gelu", 1, False), ("ln fold + gelu", 1, True)]:
    for _ in range(5):
        assert run(epi, fold) == 0
    res = []
    for rep in range(3):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize(); e0.record()
        for _ in range(20):
            run(epi, fold)
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 20
        res.append(round(2.0 * M * F * D / ms / 1e9))
    print(f"{name:16s} TFLOP/s {res}")
