"""Developer check: the attention kernel selected by VITK_ATTN on seeded inputs -> .npy, plus timing.
  VITK_ATTN=3 python tools/dev_attn_compare.py /tmp/a3.npy; VITK_ATTN=5 python tools/dev_attn_compare.py /tmp/a5.npy
  python tools/dev_attn_compare.py --diff /tmp/a3.npy /tmp/a5.npy"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

if sys.argv[1] == "--diff":
    a, b = np.load(sys.argv[2]), np.load(sys.argv[3])
    print("identical" if np.array_equal(a, b) else f"DIFFER: max abs {np.abs(a.astype(np.float64) - b).max()}, "
          f"{(a != b).mean() * 100:.3f}% of elements")
    sys.exit(0)
import torch
import vit_research_b200 as v
from vit_research_b200._lib import ptr, stream_ptr
lib = v.load()
outs = []
for frames, S, H, std in [(7, 197, 12, 1.0), (3, 197, 16, 3.0), (2, 16, 12, 1.0), (5, 100, 12, 1.0), (4, 208, 12, 0.5), (2, 197, 12, 40.0)]:
    g = torch.Generator().manual_seed(S * 7 + frames)
    qkv = (torch.randn(frames, S, 3 * H * 64, generator=g) * std).cuda().bfloat16()
    out = torch.zeros(frames, S, H * 64, device="cuda", dtype=torch.bfloat16)
    assert lib.vitk_op_attention(ptr(qkv), ptr(out), frames, S, H, stream_ptr()) == 0
    torch.cuda.synchronize()
    q, k, vv = qkv.float().view(frames, S, 3, H, 64).permute(2, 0, 3, 1, 4)
    ref = torch.softmax(q @ k.transpose(-1, -2) * 0.125, -1) @ vv
    ref = ref.permute(0, 2, 1, 3).reshape(frames, S, H * 64)
    err = (out.float() - ref).abs().max().item()
    print(f"frames {frames} S {S} H {H} std {std}: max abs err vs torch fp32 {err:.4f}")
    outs.append(out.float().cpu().numpy().ravel())
np.save(sys.argv[1], np.concatenate(outs))
frames, S, H = 256, 197, 12
qkv = [torch.randn(frames, S, 3 * H * 64, device="cuda").bfloat16() for _ in range(2)]
out = torch.empty(frames, S, H * 64, device="cuda", dtype=torch.bfloat16)
for _ in range(5):
    lib.vitk_op_attention(ptr(qkv[0]), ptr(out), frames, S, H, stream_ptr())
res = []
for rep in range(3):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for i in range(50):
        lib.vitk_op_attention(ptr(qkv[i & 1]), ptr(out), frames, S, H, stream_ptr())
    e1.record()
    torch.cuda.synchronize()
    res.append(round(e0.elapsed_time(e1) / 50 * 1e3, 1))
print("attention us:", res, "VITK_ATTN", os.environ.get("VITK_ATTN", "3"))
