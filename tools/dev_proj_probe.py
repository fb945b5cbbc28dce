"""Developer probe: the residual GEMMs (attention out-projection K=768, fc2 K=3072) at the bench
shape, in place on the fp32 stream, decomposed: full (bf16 copy + row sums), without the copy,
without copy and sums, fp32 out with no residual, bf16 out.  L2 is flushed between launches."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import vit_research_b200 as v
from vit_research_b200._lib import ptr, stream_ptr
lib = v.load()
M, D = 256 * 197, 768
bf = torch.bfloat16
st = stream_ptr()
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for K in (768, 3072):
    A = torch.randn(M, K, device="cuda").to(bf)
    W = (torch.randn(D, K, device="cuda") * 0.02).to(bf)
    bias = torch.zeros(D, device="cuda")
    x = torch.randn(M, D, device="cuda")
    xb = torch.empty(M, D, device="cuda", dtype=bf)
    stats = torch.empty(M, 6, 2, device="cuda")
    def g(epi, resid, xbo, so, out):
        return lambda: lib.vitk_op_gemm_fused(ptr(A), ptr(W), ptr(out), M, D, K, epi, ptr(bias), None, None, 0,
                                              0.0, resid, xbo, so, st)
    variants = [("full", g(3, ptr(x), ptr(xb), ptr(stats), x)),
                ("no copy", g(3, ptr(x), None, ptr(stats), x)),
                ("no copy/sums", g(3, ptr(x), None, None, x)),
                ("f32 out, no resid", g(5, None, None, None, x)),
                ("bf16 out", g(0, None, None, None, xb))]
    for name, run in variants:
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
        print(f"XBT={os.environ.get('VITK_GEMM_XBT', 'default')} K={K} {name:18s}: us {res}")
        if os.environ.get("PROBE_FULL_ONLY"):
            break
