"""Developer diagnostic: clock64 timeline of CTA 0 of the v5 attention kernel (VITK_ATTN=5):
loader warp 4 and math warp 8, first 16 (unit, slot) items.
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
for i in range(12):
    L, M = t[0, i] - t0, t[1, i] - t0
    print(f"item {i:2d} | L start {L[0]:7d} wait_S {L[1]-L[0]:5d} stream {L[2]-L[1]:5d} epi_prev {L[3]-L[2]:5d} "
          f"| M start {M[0]:7d} wait_S {M[1]-M[0]:5d} chunk0 {M[2]-M[1]:5d} chunks {M[3]-M[2]:5d} redo+arrive {M[4]-M[3]:5d}")
