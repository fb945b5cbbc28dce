"""Developer probe: the four encoder GEMMs at the bench shape (M = 256 x 197), L2 flushed between
launches.  Run once per setting of the VITK_GEMM_* switches to A/B a kernel variant."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import vit_research_b200 as v
from vit_research_b200._lib import ptr, stream_ptr
lib = v.load()
M, D, F = 256 * 197, 768, 3072
bf = torch.bfloat16
st = stream_ptr()
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
parts = 6
stats = torch.zeros(M, parts, 2, device="cuda"); stats[:, 0, 1] = 768.0
x = torch.randn(M, D, device="cuda")
xb = torch.empty(M, D, device="cuda", dtype=bf)
so = torch.empty(M, parts, 2, device="cuda")
def mk(N, K, epi):
    A = torch.randn(M, K, device="cuda").to(bf)
    W = (torch.randn(N, K, device="cuda") * 0.02).to(bf)
    b, c1 = torch.zeros(N, device="cuda"), torch.zeros(N, device="cuda")
    if epi == 3:
        return lambda: lib.vitk_op_gemm_fused(ptr(A), ptr(W), ptr(x), M, N, K, 3, ptr(b), None, None, 0, 0.0,
                                              ptr(x), ptr(xb), ptr(so), st), (A, W, b)
    out = torch.empty(M, N, device="cuda", dtype=bf)
    return lambda: lib.vitk_op_gemm_fused(ptr(A), ptr(W), ptr(out), M, N, K, epi, ptr(b), ptr(c1), ptr(stats),
                                          parts, 1e-12, None, None, None, st), (A, W, b, c1, out)
tag = " ".join(f"{k}={os.environ[k]}" for k in sorted(os.environ) if k.startswith("VITK_GEMM"))
for name, N, K, epi in [("qkv", 3 * D, D, 0), ("proj", D, D, 3), ("fc1", F, D, 1), ("fc2", D, F, 3)]:
    run, keep = mk(N, K, epi)
    for _ in range(5):
        assert run() == 0, lib.vitk_last_error(None)
    res = []
    for rep in range(3):
        tot = 0.0
        for _ in range(10):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); run(); e1.record(); torch.cuda.synchronize()
            tot += e0.elapsed_time(e1)
        res.append(round(tot / 10 * 1000, 1))
    print(f"[{tag}] {name:5s}: us {res}  {2.0 * M * N * K / min(res) / 1e6:.0f} TFLOP/s")
    del run, keep
