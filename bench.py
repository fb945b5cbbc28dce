#!/usr/bin/env python
"""bench.py — frames/s of the hot path (ViT-B/16@224 embed + k-NN vote) on N B200s.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

One "step" = one batch of 256 synthetic frames through the whole path (BASELINE.json configs[1]):
im2col -> 12-layer ViT-B/16 -> CLS -> L2-normalise -> cosine top-25 against the reference's three
label banks (1472 rows) -> left/right/none vote.  `value` = frames/s with the bf16 frames already
in HBM; `e2e` = the same through the host-buffer API (uint8 frames in pinned host memory, H2D and
D2H inside the timed region).  N > 1: one process per GPU (torchrun), frames sharded, weights and
banks replicated, one NCCL all-gather of labels + embeddings per step (weak scaling).
`--impl reference` times the reference's own CPU path (HF ViTModel fp32 + numpy k-NN, via oracle/).
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

BATCH = 256                 # frames per step per GPU (configs[1])
K_NEIGHBOURS = 25           # chroma.py:27 top_n_closest
FLOP_PER_FRAME = 35.126e9   # BASELINE.md §3 (GEMMs + attention matmuls, 197 tokens, 12 layers)
BANK_DIR = os.path.join(ROOT, "tests", "golden", "banks")
METRIC = "frames/sec ViT-B/16@224 embed+kNN"


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return {"bf16_burst": d["bf16_tflops"], "bf16_sustained": d["bf16_tflops_sustained"],
                "hbm": d["hbm_gbs"], "source": "measured"}
    return {"bf16_burst": 1590.0, "bf16_sustained": 1400.0, "hbm": 6650.0, "source": "fallback"}


# ---------------------------------------------------------------------------------------------
# clocks: sampled DURING the timed region
# ---------------------------------------------------------------------------------------------
class ClockSampler:
    REASONS = {0x4: "sw_power_cap", 0x8: "hw_slowdown", 0x20: "sw_thermal_slowdown",
               0x40: "hw_thermal_slowdown", 0x80: "hw_power_brake_slowdown", 0x2: "applications_clocks_setting",
               0x10: "sync_boost"}

    def __init__(self, device_index: int):
        self.samples, self.reasons, self.max_mhz, self.power = [], set(), None, []
        self._stop = threading.Event()
        self._thread = None
        try:
            import pynvml
            import torch
            pynvml.nvmlInit()
            uuid = str(torch.cuda.get_device_properties(device_index).uuid)
            if not uuid.startswith("GPU-"):
                uuid = "GPU-" + uuid
            self.h = pynvml.nvmlDeviceGetHandleByUUID(uuid.encode() if hasattr(uuid, "encode") else uuid)
            self.nv = pynvml
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception as ex:  # pragma: no cover
            self.nv, self.err = None, repr(ex)

    def _run(self):
        while not self._stop.is_set():
            try:
                self.samples.append(self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_SM))
                self.power.append(self.nv.nvmlDeviceGetPowerUsage(self.h) / 1e3)
                r = self.nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                for bit, name in self.REASONS.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            self._stop.wait(0.02)

    def __enter__(self):
        if self.nv:
            self._thread = threading.Thread(target=self._run, daemon=True)
            self._thread.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        if self._thread:
            self._thread.join()

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": ["unsampled"]}
        s = sorted(self.samples)
        pw = sorted(self.power) or [0.0]
        return {"sm_mhz": s[len(s) // 2], "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(s), "power_w": round(pw[len(pw) // 2], 1)}


# ---------------------------------------------------------------------------------------------
# CPU baseline / reference arm: the reference's own path on the host cores (oracle/)
# ---------------------------------------------------------------------------------------------
def cpu_reference_time(sample: int, repeats: int, warmup: int):
    """Seconds per `sample` frames of HF ViTModel fp32 eager (the reference's embed path,
    ratt_write_clips_to_ragdb.py:145-153) + numpy brute-force k-NN + vote on the real banks."""
    import numpy as np
    import torch
    from oracle import knn_ref, vit_ref  # the one place bench.py executes oracle/

    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    model = vit_ref.build_hf_model("vit_b16", 0)
    rows, labels = knn_ref.load_banks(BANK_DIR)
    px = vit_ref.synthetic_pixels(sample, seed=1234)

    def once():
        t0 = time.perf_counter()
        emb = vit_ref.hf_embed(model, px, normalise=True, batch=sample)
        idx, _, _ = knn_ref.cosine_topk(emb, rows, K_NEIGHBOURS, dtype=np.float32)
        knn_ref.vote_pass2(knn_ref.class_counts(idx, labels))
        return time.perf_counter() - t0

    for _ in range(warmup):
        once()
    times = [once() for _ in range(repeats)]
    return times, cores, torch.get_num_threads()


_JSON_FD = None


def emit_json(line):
    data = (json.dumps(line) + "\n").encode()
    if _JSON_FD is None:
        sys.stdout.write(data.decode()); sys.stdout.flush()
    else:
        os.write(_JSON_FD, data)


def run_reference(args, rank):
    if rank != 0:
        return
    sample = 32
    times, cores, threads = cpu_reference_time(sample, args.steps, args.warmup)
    total = sum(times)
    fps = sample * len(times) / total
    line = {
        "impl": "reference", "metric": METRIC, "value": fps, "unit": "frames/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * total / len(times),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic",
        "config": {"workload": "configs[1]: ViT-B/16 224x224, embed + cosine top-25 vote vs 3 class banks "
                               "(1472 rows); reference CPU arm runs a bounded sample of 32 frames/step",
                   "frames_per_step": sample, "k": K_NEIGHBOURS},
        "cpu_baseline": {"value": fps, "unit": "frames/s", "cores": threads, "kind": "reference",
                         "sample": f"{sample} frames/step x {len(times)} steps; transformers.ViTModel fp32 "
                                   f"eager on {cores} host cores ({threads} torch threads) + numpy k-NN/vote"},
        "e2e": {"value": fps, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit_json(line)


# ---------------------------------------------------------------------------------------------
# our arm
# ---------------------------------------------------------------------------------------------
def time_kernels(v, lib, torch, n_frames, warm=3, reps=10):
    """Per-kernel device times at the workload's shapes, each kernel timed ALONE: every launch has
    its own CUDA-event pair on the launching stream and is preceded by an (untimed) 256 MB L2
    flush, so it starts from a cold L2 and -- like the best-of-10 cuBLAS launch behind the burst
    peak in MEASURED_PEAKS.json -- before the power cap has pulled the clocks down (back to back
    the same launch goes 191 -> 201 us within five launches).  `ms` is the mean over `reps`
    launches, `ms_best` the fastest."""
    from vit_research_b200._lib import ptr, stream_ptr
    M, D, F, S, H = n_frames * 197, 768, 3072, 197, 12
    dev = "cuda"
    bf = torch.bfloat16
    x = torch.randn(M, D, device=dev).to(bf)
    h = torch.randn(M, F, device=dev).to(bf) * 0.1
    qkv = torch.randn(M, 3 * D, device=dev).to(bf)
    out_qkv = torch.empty(M, 3 * D, device=dev, dtype=bf)
    out_h = torch.empty(M, F, device=dev, dtype=bf)
    resid = torch.randn(M, D, device=dev)
    ln_out = torch.empty(M, D, device=dev, dtype=bf)
    w_qkv = (torch.randn(3 * D, D, device=dev) * 0.02).to(bf)
    w_o = (torch.randn(D, D, device=dev) * 0.02).to(bf)
    w_1 = (torch.randn(F, D, device=dev) * 0.02).to(bf)
    w_2 = (torch.randn(D, F, device=dev) * 0.02).to(bf)
    b3, b1, bF = torch.zeros(3 * D, device=dev), torch.zeros(D, device=dev), torch.zeros(F, device=dev)
    g1 = torch.ones(D, device=dev)
    st = stream_ptr()
    parts = 2 * D // 256
    xb = torch.empty(M, D, device=dev, dtype=bf)
    stats = torch.empty(M, parts, 2, device=dev)
    stats2 = torch.empty(M, parts, 2, device=dev)
    assert lib.vitk_op_rowstats(ptr(resid), ptr(xb), ptr(stats), M, D, parts, st) == 0
    c3, cF = torch.zeros(3 * D, device=dev), torch.zeros(F, device=dev)
    fused = lib.vitk_op_gemm_fused
    ops = {
        # the four GEMMs exactly as the forward launches them (LayerNorm folded / residual fused)
        "gemm_qkv": (lambda: fused(ptr(xb), ptr(w_qkv), ptr(out_qkv), M, 3 * D, D, 0, ptr(b3), ptr(c3), ptr(stats), parts, 1e-12, None, None, None, st), 2.0 * M * 3 * D * D),
        "gemm_proj_resid": (lambda: fused(ptr(x), ptr(w_o), ptr(resid), M, D, D, 3, ptr(b1), None, None, 0, 0.0, ptr(resid), ptr(xb), ptr(stats2), st), 2.0 * M * D * D),
        "gemm_fc1_gelu": (lambda: fused(ptr(xb), ptr(w_1), ptr(out_h), M, F, D, 1, ptr(bF), ptr(cF), ptr(stats), parts, 1e-12, None, None, None, st), 2.0 * M * F * D),
        "gemm_fc2_resid": (lambda: fused(ptr(h), ptr(w_2), ptr(resid), M, D, F, 3, ptr(b1), None, None, 0, 0.0, ptr(resid), ptr(xb), ptr(stats2), st), 2.0 * M * D * F),
        "attention": (lambda: lib.vitk_op_attention(ptr(qkv), ptr(ln_out), n_frames, S, H, st), 4.0 * n_frames * H * S * S * 64),
        "rowstats_cast": (lambda: lib.vitk_op_rowstats(ptr(resid), ptr(ln_out), ptr(stats2), M, D, parts, st), None),
    }
    res = {}
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    for name, (fn, flops) in ops.items():
        for _ in range(warm):
            assert fn() == 0, lib.vitk_last_error(None)
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(reps)]
        torch.cuda.synchronize()
        for e0, e1 in ev:
            flush.zero_()
            e0.record()
            fn()
            e1.record()
        torch.cuda.synchronize()
        times = [e0.elapsed_time(e1) for e0, e1 in ev]
        ms = sum(times) / reps
        r = {"ms": round(ms, 4), "ms_best": round(min(times), 4)}
        if flops:
            r["tflops"] = round(flops / ms / 1e9, 1)
        else:
            r["gbs"] = round(M * D * 6 / ms / 1e6, 1)  # f32 read + bf16 write
        if name == "gemm_proj_resid":
            # HBM-bound: A bf16 + residual f32 in/out + bf16 copy (+ weights, negligible)
            r["gbs"] = round(M * D * (2 + 4 + 4 + 2) / ms / 1e6, 1)
        res[name] = r
    return res


def run_ours(args, rank, local_rank, world):
    import numpy as np
    import torch
    import vit_research_b200 as v

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the hot path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    lib = v.load()
    peaks = measured_peaks()
    geom = v.ViTGeometry.vit_b16()
    emb = v.ViTEmbedder(geom, v.random_init_state_dict(geom, seed=0), device=local_rank, max_batch=BATCH)
    rows, labels, _ = v.load_label_banks(BANK_DIR)
    # cosine vote on unit rows (north_star); on these equal-norm banks it selects the same
    # neighbours as the reference's Chroma 'l2' space
    index = v.BankIndex(emb, rows, labels, 3, metric="cosine", normalise_rows=True)

    # synthetic inputs: 4 distinct device batches (rotated: 4 x 77 MB bf16 > L2) + 2 pinned host batches
    g = torch.Generator().manual_seed(1234 + rank)
    n_dev_bufs = 4
    dev_px = [(torch.randn(BATCH, 3, 224, 224, generator=g) * 0.5).clamp_(-1, 1).to("cuda").bfloat16()
              for _ in range(n_dev_bufs)]
    host_frames = [torch.randint(0, 256, (BATCH, 224, 224, 3), generator=g, dtype=torch.uint8).pin_memory()
                   for _ in range(2)]
    host_out = (torch.empty((BATCH, 768), dtype=torch.float32).pin_memory().numpy(),
                torch.empty((BATCH,), dtype=torch.int32).pin_memory().numpy(),
                torch.empty((BATCH, K_NEIGHBOURS), dtype=torch.int32).pin_memory().numpy(),
                torch.empty((BATCH, K_NEIGHBOURS), dtype=torch.float32).pin_memory().numpy(),
                torch.empty((BATCH, 3), dtype=torch.int32).pin_memory().numpy())
    emb_out = torch.empty((BATCH, 768), dtype=torch.float32, device="cuda")
    gather_emb = torch.empty((world * BATCH, 768), dtype=torch.float32, device="cuda") if world > 1 else None
    gather_lab = torch.empty((world * BATCH,), dtype=torch.int32, device="cuda") if world > 1 else None

    def step(i):
        emb.embed(dev_px[i % n_dev_bufs], l2_normalise=True, out=emb_out)
        idx, score, label, counts = index.search(emb_out, K_NEIGHBOURS)
        if world > 1:
            dist.all_gather_into_tensor(gather_emb, emb_out)
            dist.all_gather_into_tensor(gather_lab, label)
        return label

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for i in range(args.warmup):
        step(i)
    barrier()
    launches0 = emb.launch_count
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    if rank == 0:
        lib.vitk_profile_fc1(emb.handle, 1)   # event pair around every fc1 launch of the timed steps
    with ClockSampler(local_rank) as clk:
        barrier()
        e0.record()
        for i in range(args.steps):
            step(i)
        e1.record()
        barrier()
    ms_total = e0.elapsed_time(e1)
    launches = emb.launch_count - launches0
    fc1_in_step = None
    if rank == 0:
        import ctypes
        tot, cnt = ctypes.c_double(0.0), ctypes.c_int(0)
        lib.vitk_profile_fc1(emb.handle, 0)
        if lib.vitk_profile_fc1_read(emb.handle, ctypes.byref(tot), ctypes.byref(cnt)) == 0 and cnt.value:
            avg_ms = tot.value / cnt.value
            tf = 2.0 * BATCH * 197 * 3072 * 768 / avg_ms / 1e9
            fc1_in_step = {"launches": cnt.value, "avg_ms": round(avg_ms, 4), "tflops": round(tf, 1),
                           "share_of_step": round(tot.value / ms_total, 4),
                           "frac_of_sustained_peak": round(tf / peaks["bf16_sustained"], 4)}
    if world > 1:
        t = torch.tensor([ms_total], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_total = float(t.item())

    # e2e: the public host-buffer STREAMING API (embed_and_vote_stream: what a user walking a
    # frame list calls).  Every step uploads its own 256 uint8 frames from pinned host memory
    # and downloads its embeddings + labels + neighbours inside the timed region; two steps are
    # in flight so the copies overlap the previous / next step's forward.
    host_out2 = tuple(torch.empty(a.shape, dtype=torch.from_numpy(a).dtype).pin_memory().numpy()
                      for a in host_out)
    outs = (host_out, host_out2)

    def e2e_run(n_steps):
        acc = 0
        frames_iter = (host_frames[i % 2].numpy() for i in range(n_steps))
        for res in emb.embed_and_vote_stream(frames_iter, index, K_NEIGHBOURS, outs=outs):
            acc += int(res[1][0])  # touch the step's result on the host
        return acc

    e2e_run(max(2, min(args.warmup, 3)))
    barrier()
    t0 = time.perf_counter()
    e2e_run(args.steps)
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    if world > 1:
        t = torch.tensor([e2e_s], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_s = float(t.item())
    h2d = BATCH * 224 * 224 * 3
    d2h = sum(a.nbytes for a in host_out)

    frames = world * BATCH * args.steps
    value = frames / (ms_total / 1e3)
    line = {
        "metric": METRIC, "value": value, "unit": "frames/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_total / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": {"workload": "configs[1]: ViT-B/16 224x224 bf16, 256 synthetic frames/step/GPU, embed + "
                               "cosine top-25 vote vs the reference's 3 class banks (1472 rows)",
                   "frames_per_step_per_gpu": BATCH, "k": K_NEIGHBOURS, "weights": "random-init (HF init statistics)",
                   "l2_flush": "4 rotating 77 MB bf16 input batches + ~0.9 GB of activations per step exceed the 126 MB L2",
                   "parallelism": f"dp{world} (frames sharded, weights+banks replicated, all-gather of labels+embeddings)"},
        "e2e": {"value": frames / e2e_s, "unit": "frames/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h},
        "gpu_launches": int(launches),
        "clocks": clk.summary(),
        "tensor_frac_of_sustained_peak": value / world * FLOP_PER_FRAME / 1e12 / peaks["bf16_sustained"],
    }
    if rank == 0 and world == 1:
        # per-kernel numbers are "kernel timed alone": let the board's power average recover from
        # the timed loop first (the step itself runs under sw_power_cap, see `clocks`), so they are
        # comparable with the burst peak they are divided by
        time.sleep(2.0)
        kern = time_kernels(v, lib, torch, BATCH)
        dom = kern["gemm_fc1_gelu"]
        line["kernels"] = kern
        traffic = None
        tp = os.path.join(ROOT, "profiles", "roofline_traffic.json")
        if os.path.exists(tp):
            traffic = json.load(open(tp)).get("gemm_fc1_gelu_dram_bytes_per_launch")
        line["roofline"] = {"kernel": "gemm_bf16_tn_kernel<256,BIAS_GELU_BF16> M=50432 N=3072 K=768",
                            "bound": "tensor", "achieved": dom["tflops"], "peak": peaks["bf16_burst"],
                            "unit": "TFLOP/s", "frac": dom["tflops"] / peaks["bf16_burst"],
                            "in_step": fc1_in_step, "traffic": traffic, "peak_source": peaks["source"] + " (burst; achieved = mean of 10 launches, each timed alone after an L2 flush)"}
        times, cores, threads = cpu_reference_time(32, 2, 1)
        best = min(times)
        line["cpu_baseline"] = {"value": 32 / best, "unit": "frames/s", "cores": threads, "kind": "reference",
                                "sample": f"32 frames, best of 2 after 1 warm-up; transformers.ViTModel fp32 eager "
                                          f"(the reference's embed path) + numpy k-NN/vote on {cores} host cores"}
    if rank == 0:
        emit_json(line)
    emb.close()
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    # stdout carries the ONE JSON line and nothing else: libraries that write to fd 1 (NCCL prints
    # "NCCL version ..." there when the box exports NCCL_DEBUG) are sent to stderr instead.
    global _JSON_FD
    sys.stdout.flush()
    _JSON_FD = os.dup(1)
    os.dup2(2, 1)
    if args.impl == "reference":
        run_reference(args, rank)
    else:
        if args.warmup < 3:
            args.warmup = 3
        run_ours(args, rank, local_rank, world)


if __name__ == "__main__":
    main()
