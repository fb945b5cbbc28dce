"""Developer diagnostic: clock64 timeline of CTA 0 of the v6 attention kernel (VITK_ATTN=6).
Needs a developer build of the library: VITK_NVCC_EXTRA=-DVITK_ATT_TIMELINE=1 (the stamps are compiled out of production builds)."""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import vit_research_b200 as v
from vit_research_b200._lib import ptr, stream_ptr
lib = v.load()
raw = C.CDLL(str(v._lib.LIB_PATH))
frames, S, H = 256, 197, 12
qkv = torch.randn(frames, S, 3 * H * 64, device="cuda").bfloat16()
out = torch.empty(frames, S, H * 64, device="cuda", dtype=torch.bfloat16)
buf = torch.zeros(2 * 16 * 8, device="cuda", dtype=torch.int64)
for _ in range(2):
    lib.vitk_op_attention(ptr(qkv), ptr(out), frames, S, H, stream_ptr())
raw.vitk_dev_attn_timeline(C.c_void_p(buf.data_ptr()))
lib.vitk_op_attention(ptr(qkv), ptr(out), frames, S, H, stream_ptr())
torch.cuda.synchronize()
raw.vitk_dev_attn_timeline(None)
t = buf.cpu().view(2, 16, 8).numpy()
t0 = t[:, :, 0][t[:, :, 0] > 0].min()
for g in range(2):
    for j in range(2, 10):
        r = t[g, j] - t0
        print(f"g{g} j{j:2d} start {r[0]:7d} wait_S {r[6]-r[0]:5d} turn+other_O {r[1]-r[6]:5d} pass {r[2]-r[1]:5d} "
              f"redo/arrive {r[3]-r[2]:5d} wait_O {r[4]-r[3]:5d} epi {r[5]-r[4]:5d} total {r[5]-r[0]:6d}")
