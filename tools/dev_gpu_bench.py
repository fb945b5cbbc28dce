"""Developer diagnostic: per-kernel device times at the bench shapes (bench.time_kernels)."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import vit_research_b200 as v
import bench

lib = v.load()
n = int(sys.argv[1]) if len(sys.argv) > 1 else 256
warm = int(os.environ.get("KB_WARM", "3")); reps = int(os.environ.get("KB_REPS", "10"))
print(json.dumps(bench.time_kernels(v, lib, torch, n, warm, reps), indent=1))

# library bar: cuBLAS (torch.matmul) on the same GEMM shapes, same timing loop, no epilogue
if os.environ.get("KB_CUBLAS", "1") == "1":
    M, D, F = n * 197, 768, 3072
    res = {}
    for name, (m_, n_, k_) in {"qkv": (M, 3 * D, D), "proj": (M, D, D), "fc1": (M, F, D), "fc2": (M, D, F)}.items():
        a = torch.randn(m_, k_, device="cuda").bfloat16()
        w = torch.randn(n_, k_, device="cuda").bfloat16()
        for _ in range(3):
            torch.matmul(a, w.t())
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize(); e0.record()
        for _ in range(10):
            torch.matmul(a, w.t())
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 10
        res[name] = {"ms": round(ms, 4), "tflops": round(2.0 * m_ * n_ * k_ / ms / 1e9, 1)}
    print("cublas_plain_gemm", json.dumps(res))
