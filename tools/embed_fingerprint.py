"""Prints the embeddings of a fixed tiny workload as JSON (used by the variant tests, which run it
under VITK_GEMM_CG / VITK_FULL_LAST_LAYER / VITK_ATTN to compare the code paths)."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import vit_research_b200 as v

geom = v.ViTGeometry(layers=3)
emb = v.ViTEmbedder(geom, v.random_init_state_dict(geom, seed=5), device=0, max_batch=8)
g = torch.Generator().manual_seed(21)
frames = torch.randint(0, 256, (11, 224, 224, 3), generator=g, dtype=torch.uint8).numpy()
out = emb.embed_frames(frames, l2_normalise=True)
emb.close()
print(json.dumps({"emb": np.asarray(out, np.float64).round(9).tolist()}))
