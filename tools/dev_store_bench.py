"""Developer timing of the large-store top-k (BASELINE configs[4] shape): 1 M x 768 store, 256
queries, cosine top-10.  CUDA events around repeated vitk_knn_vote calls."""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import vit_research_b200 as v
from oracle import vit_ref

m = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
nq = int(sys.argv[2]) if len(sys.argv) > 2 else 256
k = int(sys.argv[3]) if len(sys.argv) > 3 else 10
model = vit_ref.build_hf_model("vit_b16_l2", 0, randomise_ln=True)
emb = v.ViTEmbedder.from_hf(model, device=0, max_batch=8)
g = torch.Generator(device="cuda").manual_seed(99)
store = torch.randn((m, 768), generator=g, device="cuda")
index = v.BankIndex(emb, store, None, 0, metric="cosine", normalise_rows=True)
del store
qs = [torch.randn((nq, 768), generator=g, device="cuda") for _ in range(4)]
for i in range(3):
    index.search(qs[i % 4], k, normalise_queries=True)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
reps = 20
e0.record()
for i in range(reps):
    index.search(qs[i % 4], k, normalise_queries=True)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / reps
flops = 2.0 * nq * m * 768
print(json.dumps({"m": m, "queries": nq, "k": k, "ms_per_call": ms, "queries_per_s": nq / ms * 1e3,
                  "tflops": flops / ms / 1e9, "store_gbs": m * 768 * 2 / ms / 1e6,
                  "redo": v.load().vitk_knn_redo_count(emb.handle)}))
