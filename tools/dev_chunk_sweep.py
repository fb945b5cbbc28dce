"""Dev probe: time one bench step (256 frames, embed + top-25 vote) with the forward run in
sub-batches of `max_batch` frames, to see whether activations that fit the 126 MB L2 (and the
wave count of the 256-row GEMM tiles) beat one 256-frame pass.  Run under gpurun:
  python tools/dev_chunk_sweep.py 256 128 127 96 64
"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import vit_research_b200 as v

BATCH, K = 256, 25
sizes = [int(a) for a in sys.argv[1:]] or [256, 128, 127, 96, 64]
steps = int(os.environ.get("SWEEP_STEPS", "30"))
v.load()
geom = v.ViTGeometry.vit_b16()
sd = v.random_init_state_dict(geom, seed=0)
rows, labels, _ = v.load_label_banks(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", "banks"))
g = torch.Generator().manual_seed(1234)
px = [(torch.randn(BATCH, 3, 224, 224, generator=g) * 0.5).clamp_(-1, 1).cuda().bfloat16() for _ in range(4)]
out = torch.empty((BATCH, 768), dtype=torch.float32, device="cuda")
ref = None
for mb in sizes:
    emb = v.ViTEmbedder(geom, sd, device=0, max_batch=mb)
    index = v.BankIndex(emb, rows, labels, 3, metric="cosine", normalise_rows=True)
    def step(i):
        emb.embed(px[i % 4], l2_normalise=True, out=out)
        return index.search(out, K)
    for i in range(5):
        step(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(steps):
        step(i)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    step(0)
    cur = out.clone()
    same = None if ref is None else bool(torch.equal(cur, ref))
    if ref is None:
        ref = cur
    print(f"max_batch {mb:4d}: {ms:7.3f} ms/step  {BATCH / ms * 1e3:9.1f} frames/s  identical_to_first={same}", flush=True)
    emb.close()
