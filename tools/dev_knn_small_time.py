"""Developer timing of the small-bank k-NN vote (1 472-row bank, 256 queries, k = 25), CUDA events around repeated calls."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import vit_research_b200 as v
from oracle import vit_ref
model = vit_ref.build_hf_model("vit_b16_l2", 0, randomise_ln=True)
emb = v.ViTEmbedder.from_hf(model, device=0, max_batch=8)
g = torch.Generator(device="cuda").manual_seed(5)
bank = torch.randn((1472, 768), generator=g, device="cuda")
lab = torch.randint(0, 3, (1472,), generator=g, device="cuda").int().cpu().numpy()
for metric in ("cosine", "l2"):
    index = v.BankIndex(emb, bank, lab, 3, metric=metric, normalise_rows=True)
    q = torch.randn((256, 768), generator=g, device="cuda")
    for _ in range(3):
        out = index.search(q, 25, normalise_queries=True)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(50):
        out = index.search(q, 25, normalise_queries=True)
    e1.record(); torch.cuda.synchronize()
    print(metric, "us per 256-query vote:", round(e0.elapsed_time(e1) * 20, 1), "checksum", int(out[0].sum()), float(out[1].sum()))
