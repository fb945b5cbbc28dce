"""Developer timing of patchify_kernel at the bench shape (256 frames, bf16 NCHW), each launch after an L2 flush."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import vit_research_b200 as v
from vit_research_b200._lib import ptr, stream_ptr
lib = v.load()
n = 256
pix = torch.randn(n, 3, 224, 224, device="cuda").bfloat16()
out = torch.empty(n * 197, 768, device="cuda", dtype=torch.bfloat16)
big = torch.empty(256 << 20, device="cuda", dtype=torch.uint8)
ts = []
for i in range(8):
    big.zero_()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    assert lib.vitk_op_patchify(ptr(pix), 1, ptr(out), n, 3, 224, 224, 16, stream_ptr()) == 0
    e1.record(); torch.cuda.synchronize()
    ts.append(round(e0.elapsed_time(e1) * 1e3, 1))
print("patchify us (after L2 flush):", ts)
