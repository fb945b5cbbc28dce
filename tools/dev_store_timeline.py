"""Developer diagnostic: per-unit timeline (CTA 0) of the LAST threshold-filter GEMM stage of a
1 M-row store search (the stage that walks 83 % of the rows).
Needs a developer build of the library: VITK_NVCC_EXTRA=-DVITK_GEMM_TIMELINE=1 (the stamps are compiled out of production builds)."""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import vit_research_b200 as v
from oracle import vit_ref
m, nq, k = 1_000_000, 256, 10
model = vit_ref.build_hf_model("vit_b16_l2", 0, randomise_ln=True)
emb = v.ViTEmbedder.from_hf(model, device=0, max_batch=8)
raw = C.CDLL(str(v._lib.LIB_PATH))
g = torch.Generator(device="cuda").manual_seed(99)
store = torch.randn((m, 768), generator=g, device="cuda")
index = v.BankIndex(emb, store, None, 0, metric="cosine", normalise_rows=True)
del store
q = torch.randn((nq, 768), generator=g, device="cuda")
for _ in range(3):
    index.search(q, k, normalise_queries=True)
torch.cuda.synchronize()
buf = torch.zeros(41 * 8, device="cuda", dtype=torch.int64)
raw.vitk_dev_gemm_timeline(C.c_void_p(buf.data_ptr()))
index.search(q, k, normalise_queries=True)
torch.cuda.synchronize()
raw.vitk_dev_gemm_timeline(None)
tt = buf.cpu().numpy()
t = tt[:320].reshape(40, 8)
n = int((t[:, 2] > 0).sum())
t0 = t[0, 0]
cyc, ns = tt[322] - tt[320], tt[323] - tt[321]
print(f"mainloop of CTA 0: {cyc} cycles in {ns} ns = {cyc / max(ns, 1) * 1000:.0f} MHz")
print("unit  mma: wait_acc  issue   | epi: wait_full  busy(w0) busy(w4) | period  | producer ahead of mma start")
for i in range(min(n, 40)):
    r = t[i] - t0
    period = (t[i + 1, 1] - t[i, 1]) if i + 1 < n else 0
    print(f"{i:3d}   {r[1]-r[0]:8d} {r[2]-r[1]:7d}  | {r[4]-r[3]:9d} {r[5]-r[4]:8d} {r[6]-r[4]:8d} | {period:6d} | {r[1]-r[7]:6d}")
