"""Generate tests/golden/* from the reference, in the build container only (/root/reference and
the reference's own dependency transformers.ViTModel are imported here; neither is needed at
test time).  Re-run: python tools/make_golden.py
"""
import hashlib, json, os, shutil, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from oracle import knn_ref, vit_ref

REF = "/root/reference/nba_proj"
OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")
os.makedirs(os.path.join(OUT, "banks"), exist_ok=True)
manifest = {}

# 1. the reference's only data fixtures: the three label banks, copied verbatim
for name in knn_ref.LABELS:
    src = f"{REF}/{name}_embeddings.npz"
    dst = f"{OUT}/banks/{name}_embeddings.npz"
    shutil.copyfile(src, dst)
    manifest[f"banks/{name}_embeddings.npz"] = hashlib.sha256(open(dst, "rb").read()).hexdigest()

# 2. leave-one-out exact top-51 tables on the banks (fp64 brute force), cosine and L2
rows, labels = knn_ref.load_banks(f"{OUT}/banks")
K = 51
ci, cs, _ = knn_ref.cosine_topk(rows, rows, K, exclude_self=True)
li, ls, _ = knn_ref.l2_topk(rows, rows, K, exclude_self=True)
rates = {}
for k in (5, 25, 50):
    lab = knn_ref.vote_pass2(knn_ref.class_counts(li[:, :k], labels))
    rates[str(k)] = float((lab == labels).mean())
np.savez_compressed(f"{OUT}/banks_loo_topk.npz", cos_idx=ci.astype(np.int16), cos_sim=cs,
                    l2_idx=li.astype(np.int16), l2_dist=ls)
manifest["loo_own_label_rate_l2"] = rates
print("leave-one-out own-label rates (L2):", rates)

# 3. ViT half: HF ViTModel (the reference's dependency) on seeded weights and inputs
torch.set_num_threads(os.cpu_count())
for name, seed, n in (("vit_b16", 0, 8), ("vit_l16_l2", 1, 4)):
    model = vit_ref.build_hf_model(name, seed, randomise_ln=True)
    px = vit_ref.synthetic_pixels(n, seed=1234)
    raw = vit_ref.hf_embed(model, px, normalise=False)
    u8 = vit_ref.synthetic_frames_u8(n, seed=4321)
    raw_u8 = vit_ref.hf_embed(model, vit_ref.process_u8(u8), normalise=False)
    np.savez_compressed(f"{OUT}/{name}_seed{seed}.npz", cls=raw, cls_u8=raw_u8,
                        fingerprint=np.float64(vit_ref.weights_fingerprint(model)))
    manifest[f"{name}_seed{seed}"] = {"n": n, "pixel_seed": 1234, "u8_seed": 4321,
                                      "fingerprint": vit_ref.weights_fingerprint(model),
                                      "torch": torch.__version__}
    print(name, "cls norm", np.linalg.norm(raw, axis=1)[:3])
# 4. full-depth ViT-L/16 (BASELINE configs[3]: 24 layers) and the TF bank-writing geometry
#    (patch 256 -> 4 tokens, raw pixels, CLS zeros) -- parity at the depths / shapes that are run
model = vit_ref.build_hf_model("vit_l16", 1, randomise_ln=True)
raw = vit_ref.hf_embed(model, vit_ref.synthetic_pixels(4, seed=1234), normalise=False)
np.savez_compressed(f"{OUT}/vit_l16_seed1.npz", cls=raw, fingerprint=np.float64(vit_ref.weights_fingerprint(model)))
manifest["vit_l16_seed1"] = {"n": 4, "pixel_seed": 1234, "layers": 24,
                             "fingerprint": vit_ref.weights_fingerprint(model), "torch": torch.__version__}
print("vit_l16 (24 layers) cls norm", np.linalg.norm(raw, axis=1))
model = vit_ref.build_hf_model("vit_tf432_p256", 2, randomise_ln=True, cls_zeros=True)
u8 = vit_ref.synthetic_frames_u8(4, seed=4321, size=(432, 768))
raw = vit_ref.hf_embed(model, vit_ref.raw_pixels(u8), normalise=False)
np.savez_compressed(f"{OUT}/vit_tf432_p256_seed2.npz", cls_u8=raw,
                    fingerprint=np.float64(vit_ref.weights_fingerprint(model)))
manifest["vit_tf432_p256_seed2"] = {"n": 4, "u8_seed": 4321, "tokens": 4, "cls_zeros": True,
                                    "fingerprint": vit_ref.weights_fingerprint(model), "torch": torch.__version__}
print("vit_tf432_p256 cls norm", np.linalg.norm(raw, axis=1))
import transformers
manifest["transformers"] = transformers.__version__
json.dump(manifest, open(f"{OUT}/MANIFEST.json", "w"), indent=1, sort_keys=True)
