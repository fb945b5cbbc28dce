"""The "library kernels" bar of BASELINE.md §4: stock HF ViTModel (the reference's embed path) in bf16
on the same B200 — cuBLAS / cuDNN / SDPA kernels, eager — for 256-frame batches, CUDA events.
Not part of the product or of bench.py; records what our hand-written kernels are compared with."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from oracle import vit_ref

model = vit_ref.build_hf_model("vit_b16", 0).cuda().bfloat16().eval()
px = [vit_ref.synthetic_pixels(256, seed=s).cuda().bfloat16() for s in (1, 2)]
res = {}
with torch.no_grad():
    for name, m in (("eager_sdpa_bf16", model),):
        for i in range(3):
            m(pixel_values=px[i % 2]).last_hidden_state[:, 0]
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize(); e0.record()
        for i in range(10):
            cls = m(pixel_values=px[i % 2]).last_hidden_state[:, 0]
            cls = cls / (cls.norm(dim=1, keepdim=True) + 1e-8)
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 10
        res[name] = {"ms_per_256": round(ms, 3), "frames_per_s": round(256 / ms * 1e3), "tflops": round(256 * 35.126e9 / ms / 1e9, 1)}
print(json.dumps(res))
