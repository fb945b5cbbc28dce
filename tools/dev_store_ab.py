"""Developer A/B of the large-store search between two builds of libvitk (raw ctypes, old-ABI calls only):
  python tools/dev_store_ab.py vit-research_b200/libvitk_r01.so vit-research_b200/libvitk.so"""
import ctypes as C, json, sys
import torch

class Cfg(C.Structure):
    _fields_ = [(n, C.c_int32) for n in ("image_h", "image_w", "patch", "channels", "hidden", "heads", "mlp", "layers")] + \
               [("ln_eps", C.c_float), ("gelu_tanh", C.c_int32), ("max_batch", C.c_int32), ("u8_mean", C.c_float), ("u8_std", C.c_float)]

m, nq, k = 1_000_000, 256, 10
g = torch.Generator(device="cuda").manual_seed(99)
store = torch.randn((m, 768), generator=g, device="cuda")
qs = [torch.randn((nq, 768), generator=g, device="cuda") for _ in range(4)]
idx = torch.empty((nq, k), dtype=torch.int32, device="cuda"); sc = torch.empty((nq, k), device="cuda")
st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
for rnd in range(3):
    for path in sys.argv[1:]:
        lib = C.CDLL(path)
        lib.vitk_last_error.restype = C.c_char_p
        h = C.c_void_p()
        cfg = Cfg(224, 224, 16, 3, 768, 12, 3072, 1, 1e-12, 0, 8, 0.0, 0.0)
        assert lib.vitk_create(C.byref(cfg), 0, C.byref(h)) == 0
        assert lib.vitk_bank_set_device(h, C.c_void_p(store.data_ptr()), m, 768, None, 0, 1, st) == 0
        def search(q):
            rc = lib.vitk_knn_vote(h, C.c_void_p(q.data_ptr()), nq, k, 0, 1, C.c_void_p(idx.data_ptr()), C.c_void_p(sc.data_ptr()), None, None, st)
            assert rc == 0, lib.vitk_last_error(h)
        for i in range(3):
            search(qs[i % 4])
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize(); e0.record()
        for i in range(20):
            search(qs[i % 4])
        e1.record(); torch.cuda.synchronize()
        print(rnd, path.split("/")[-1], round(e0.elapsed_time(e1) / 20, 4), "ms/search", int(idx[0, 0]))
        lib.vitk_destroy(h)
