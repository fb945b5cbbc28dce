"""Developer diagnostic: per-work-unit clock64 timeline of CTA 0 of the persistent GEMM (MMA thread,
epilogue warps 4 and 8, TMA producer) at the four layer shapes.
Needs a developer build of the library: VITK_NVCC_EXTRA=-DVITK_GEMM_TIMELINE=1 (the stamps are compiled out of production builds)."""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import vit_research_b200 as v
from vit_research_b200._lib import ptr, stream_ptr
lib = v.load()
raw = C.CDLL(str(v._lib.LIB_PATH))
M, D, F = 256 * 197, 768, 3072
bf = torch.bfloat16
parts = 6
st = stream_ptr()
shapes = {"qkv": (3 * D, D, 0, True), "fc1": (F, D, 1, True), "fc1_nogelu": (F, D, 0, True),
          "fc1_plain": (F, D, 0, False), "fc2": (D, F, 7, False), "proj": (D, D, 7, False), "patch": (D, D, 8, False)}
which = sys.argv[1:] or list(shapes)
for name in which:
    N, K, epi, fold = shapes[name]
    a = torch.randn(M, K, device="cuda").to(bf)
    w = (torch.randn(N, K, device="cuda") * 0.02).to(bf)
    out = torch.zeros(M, N, device="cuda", dtype=bf)
    bias, c1 = torch.zeros(N, device="cuda"), torch.zeros(N, device="cuda")
    stats = torch.zeros(M, parts, 2, device="cuda"); stats[:, 0, 1] = float(K)
    so = torch.zeros(M, 2 * ((N + 255) // 256), 2, device="cuda")
    table = torch.randn(197, N, device="cuda")
    def run():
        if epi == 8:   # patch embedding: acc + position / CLS / bias table -> bf16 stream
            return lib.vitk_op_gemm(ptr(a), ptr(w), ptr(out), M, N, K, epi, None, None, ptr(table), 197, st)
        if epi == 7:
            return lib.vitk_op_gemm_fused(ptr(a), ptr(w), ptr(out), M, N, K, epi, ptr(bias), None, None, 0, 1e-12,
                                          ptr(out), None, ptr(so), st)
        return lib.vitk_op_gemm_fused(ptr(a), ptr(w), ptr(out), M, N, K, epi, ptr(bias), ptr(c1) if fold else None,
                                      ptr(stats) if fold else None, parts if fold else 0, 1e-12, None, None, None, st)
    for _ in range(3):
        assert run() == 0, v.last_error() if hasattr(v, "last_error") else "gemm failed"
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    for _ in range(10):
        run()
    e1.record(); torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 100
    import time
    buf = torch.zeros(41 * 8, device="cuda", dtype=torch.int64)
    # the same launch in three power states: right after the ten above, after 2 s of idle, inside a burst of 40
    for state in ("hot", "idle 2 s", "launch 40 of a burst"):
        if state.startswith("idle"):
            time.sleep(2.0)
        if state.startswith("launch"):
            for _ in range(39):
                run()
        raw.vitk_dev_gemm_timeline(C.c_void_p(buf.data_ptr()))
        run(); torch.cuda.synchronize()
        raw.vitk_dev_gemm_timeline(None)
        tt = buf.cpu().numpy()
        nn = int((tt[:320].reshape(40, 8)[:, 2] > 0).sum())
        iss = tt[:320].reshape(40, 8)
        cyc, ns = tt[322] - tt[320], tt[323] - tt[321]
        print(f"   [{state:22s}] mainloop {cyc} cycles in {ns} ns = {cyc / max(ns, 1) * 1000:.0f} MHz; "
              f"median cycles per unit {sorted(iss[1:nn, 1] - iss[0:nn-1, 1])[max(nn // 2 - 1, 0)]}")
    t = buf.cpu()[:320].view(40, 8).numpy()
    n = int((t[:, 2] > 0).sum())
    t0 = t[0, 0]
    print(f"== {name}: N={N} K={K} epi={epi} fold={fold}  {us:.1f} us/launch  {2.0*M*N*K/us/1e6:.0f} TFLOP/s  units on CTA 0: {n}")
    print("unit  mma: wait_acc  issue   | epi: wait_full  busy(w4) busy(w8) | period  | producer ahead of mma start")
    for i in range(n):
        r = t[i] - t0
        period = (t[i + 1, 1] - t[i, 1]) if i + 1 < n else 0
        print(f"{i:3d}   {r[1]-r[0]:8d} {r[2]-r[1]:7d}  | {r[4]-r[3]:9d} {r[5]-r[4]:8d} {r[6]-r[4]:8d} | {period:6d} | {r[1]-r[7]:6d}")
    if n > 4:
        import numpy as np
        mid = slice(2, n - 1)
        print(f"median: mma wait_acc {np.median(t[mid,1]-t[mid,0]):.0f}  issue {np.median(t[mid,2]-t[mid,1]):.0f}  "
              f"epi wait_full {np.median(t[mid,4]-t[mid,3]):.0f}  busy w4 {np.median(t[mid,5]-t[mid,4]):.0f} "
              f"w8 {np.median(t[mid,6]-t[mid,4]):.0f}  period {np.median(np.diff(t[1:n,1])):.0f}")
