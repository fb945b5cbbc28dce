"""Developer timing of the attention kernel alone (256 frames, S=197, 12 heads), CUDA events."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import vit_research_b200 as v
from vit_research_b200._lib import ptr, stream_ptr
lib = v.load()
frames, S, H = 256, 197, 12
qkv = [(torch.randn(frames, S, 3 * H * 64, device="cuda") * float(os.environ.get("QKV_STD", "1"))).bfloat16() for _ in range(2)]
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
print("attention us:", res, "VITK_ATTN", os.environ.get("VITK_ATTN", "3"), "handover", os.environ.get("VITK_ATTN_HANDOVER", "end"))
