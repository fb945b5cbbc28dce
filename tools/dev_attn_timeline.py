"""Developer diagnostic: per-phase clock64 timeline of CTA 0 of the attention kernel.
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
buf = torch.zeros(3 * 16 * 8, device="cuda", dtype=torch.int64)
for _ in range(2):
    lib.vitk_op_attention(ptr(qkv), ptr(out), frames, S, H, stream_ptr())
raw.vitk_dev_attn_timeline(C.c_void_p(buf.data_ptr()))
lib.vitk_op_attention(ptr(qkv), ptr(out), frames, S, H, stream_ptr())
torch.cuda.synchronize()
raw.vitk_dev_attn_timeline(None)
full = buf.cpu().view(3, 16, 8).numpy()
t = full[:2]
t0 = t[:, :, 0][t[:, :, 0] > 0].min()
print("cols: unit_start  s_full  pass1_done  p_arrived  o_full  unit_end   (cycles since first sample)")
for g in range(2):
    for j in range(12):
        r = t[g, j, :6] - t0
        d = [r[1] - r[0], r[2] - r[1], r[3] - r[2], r[4] - r[3], r[5] - r[4]]
        print(f"g{g} j{j:2d} blk{g ^ (j & 1)}  start {r[0]:7d}  wait_S {d[0]:5d}  pass1 {d[1]:5d}  pass2 {d[2]:5d}  wait_O {d[3]:5d}  epi {d[4]:5d}  total {r[5]-r[0]:6d}  | PV issue after P {t[g, j, 6] - t0 - r[3]:5d}  issue took {t[g, j, 7] - t[g, j, 6]:5d}  O after commit {r[4] - (t[g, j, 7] - t0):5d}")

ld = full[2]
print("loader: unit | Q/K load issued | kv_full seen by the MMA warp of slot 0 / 1 | slot_free seen 0 / 1 | S complete seen by group 0 / 1  (cycles since first sample)")
for j in range(12):
    r = ld[j] - t0
    print(f"  j{j:2d}  QK issued {r[0]:7d} | kv_full {r[1]:7d} / {r[2]:7d} | slot_free {r[3]:7d} / {r[4]:7d} | s_full {r[5]:7d} / {r[6]:7d} | turn {t[0, j, 1]-t0:7d} / {t[1, j, 1]-t0:7d}")
