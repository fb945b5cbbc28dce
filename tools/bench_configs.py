"""The other BASELINE.json configs at full size (they are parity-test cases, not bench lines; this
tool records what they run at).  One JSON line per config, rank 0 prints.
  python tools/bench_configs.py c3|c4|c5|c5s        (1 GPU)
  torchrun --nproc-per-node N tools/bench_configs.py c3|c4|c5|c5s
c3: ViT-B/16, 108 000-frame stream sharded by contiguous chunk, host uint8 frames through the
    streaming API (H2D/D2H inside), top-25 vote, then the per-video Viterbi on the gathered counts.
c4: ViT-L/16, batch 512, embedding only (device-resident bf16 pixels).
c5: ViT-B/16 embed of 256 frames + cosine top-10 against a replicated 1 M-row store.
"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import vit_research_b200 as v  # noqa: E402


def main():
    which = sys.argv[1]
    rank, local, world = (int(os.environ.get(k, d)) for k, d in (("RANK", 0), ("LOCAL_RANK", 0), ("WORLD_SIZE", 1)))
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def sync_max(x):
        if world == 1:
            return x
        t = torch.tensor([x], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    g = torch.Generator().manual_seed(7 + rank)
    out = {"config": which, "n_gpus": world}
    if which == "c3":
        total, B, k = 108_000, 256, 25
        geom = v.ViTGeometry.vit_b16()
        emb = v.ViTEmbedder(geom, v.random_init_state_dict(geom, 0), device=local, max_batch=B)
        rows, labels, _ = v.load_label_banks(os.path.join(ROOT, "tests", "golden", "banks"))
        index = v.BankIndex(emb, rows, labels, 3, metric="cosine", normalise_rows=True)
        lo, hi = v.frame_ranges(total, world)[rank]
        frames = [torch.randint(0, 256, (B, 224, 224, 3), generator=g, dtype=torch.uint8).pin_memory().numpy()
                  for _ in range(4)]
        n_mine = hi - lo
        counts_all = torch.empty((n_mine, 3), dtype=torch.int32).pin_memory().numpy()
        labels_all = torch.empty((n_mine,), dtype=torch.int32).pin_memory().numpy()

        def batches():
            for i, a in enumerate(range(0, n_mine, B)):
                yield frames[i % 4][: min(B, n_mine - a)]

        def run():
            a = 0
            for e, lab, idx, sc, cnt in emb.embed_and_vote_stream(batches(), index, k):
                counts_all[a:a + len(lab)] = cnt
                labels_all[a:a + len(lab)] = lab
                a += len(lab)
            c = torch.from_numpy(counts_all).cuda()
            if world > 1:
                parts = [torch.empty((r[1] - r[0], 3), dtype=torch.int32, device="cuda") for r in v.frame_ranges(total, world)]
                dist.all_gather(parts, c)
                c = torch.cat(parts)
            states = v.viterbi_from_counts(emb, c, k) if rank == 0 else None
            torch.cuda.synchronize()
            return states

        # warm-up on a short prefix
        for _ in emb.embed_and_vote_stream((frames[i % 4] for i in range(4)), index, k):
            pass
        barrier()
        t0 = time.perf_counter()
        states = run()
        barrier()
        dt = sync_max(time.perf_counter() - t0)
        out.update({"workload": "ViT-B/16 108k-frame stream, host uint8 frames, embed + top-25 vote + Viterbi",
                    "frames": total, "seconds": dt, "frames_per_s": total / dt,
                    "decoded_frames": int((states >= 0).sum()) if states is not None else None})
    elif which == "c4":
        B, steps = 512, 10
        geom = v.ViTGeometry.vit_l16()
        emb = v.ViTEmbedder(geom, v.random_init_state_dict(geom, 1), device=local, max_batch=B)
        px = [(torch.randn(B, 3, 224, 224, generator=g) * 0.5).clamp_(-1, 1).cuda().bfloat16() for _ in range(2)]
        o = torch.empty((B, geom.hidden), device="cuda")
        for i in range(3):
            emb.embed(px[i % 2], out=o)
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(steps):
            emb.embed(px[i % 2], out=o)
        e1.record()
        barrier()
        ms = sync_max(e0.elapsed_time(e1)) / steps
        fps = world * B / ms * 1e3
        out.update({"workload": "ViT-L/16 224x224 bf16 batch 512/GPU embedding only", "ms_per_step": ms,
                    "frames_per_s": fps, "tflops_per_gpu": fps / world * 123.107e9 / 1e12,
                    "finite": bool(torch.isfinite(o).all())})
    elif which == "c5":
        B, steps, k, m = 256, 10, 10, 1_000_000
        geom = v.ViTGeometry.vit_b16()
        emb = v.ViTEmbedder(geom, v.random_init_state_dict(geom, 0), device=local, max_batch=B)
        gd = torch.Generator(device="cuda").manual_seed(99)
        store = torch.randn((m, 768), generator=gd, device="cuda")
        index = v.BankIndex(emb, store, None, 0, metric="cosine", normalise_rows=True)
        del store
        px = [(torch.randn(B, 3, 224, 224, generator=g) * 0.5).clamp_(-1, 1).cuda().bfloat16() for _ in range(2)]
        o = torch.empty((B, 768), device="cuda")

        def step(i):
            emb.embed(px[i % 2], out=o)
            return index.search(o, k)

        for i in range(3):
            step(i)
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(steps):
            r = step(i)
        e1.record()
        barrier()
        ms = sync_max(e0.elapsed_time(e1)) / steps
        e0.record()
        for i in range(steps):
            r = index.search(o, k)
        e1.record()
        barrier()
        ms_s = sync_max(e0.elapsed_time(e1)) / steps
        out.update({"workload": "ViT-B/16 embed 256 frames/GPU + cosine top-10 vs replicated 1M-row store",
                    "ms_per_step": ms, "frames_per_s": world * B / ms * 1e3, "search_only_ms": ms_s,
                    "search_queries_per_s": world * B / ms_s * 1e3,
                    "redo": int(v.load().vitk_knn_redo_count(emb.handle)), "top1_score_mean": float(r[1][:, 0].mean())})
    elif which == "c5s":
        # the same search with the store ROW-SHARDED over the ranks and one exchange step
        # (all-gather queries -> local top-k -> all-gather (row, score) lists -> merge), checked
        # against the replicated store on every rank
        B, steps, k, m = 256, 10, 10, 1_000_000
        geom = v.ViTGeometry.vit_b16()
        emb = v.ViTEmbedder(geom, v.random_init_state_dict(geom, 0), device=local, max_batch=B)
        import dataclasses
        small = dataclasses.replace(v.ViTGeometry.vit_b16(), layers=1)
        emb2 = v.ViTEmbedder(small, v.random_init_state_dict(small, 0), device=local, max_batch=8)
        gd = torch.Generator(device="cuda").manual_seed(99)
        store = torch.randn((m, 768), generator=gd, device="cuda")
        full = v.BankIndex(emb, store, None, 0, metric="cosine", normalise_rows=True)
        lo, hi = v.frame_ranges(m, world)[rank]
        shard = v.BankIndex(emb2, store[lo:hi].contiguous(), None, 0, metric="cosine", normalise_rows=True)
        del store
        gq = torch.Generator().manual_seed(4321 + rank)
        px = [(torch.randn(B, 3, 224, 224, generator=gq) * 0.5).clamp_(-1, 1).cuda().bfloat16() for _ in range(2)]
        o = torch.empty((B, 768), device="cuda")

        def step(i):
            emb.embed(px[i % 2], out=o)
            return v.shard.search_sharded_store(lambda q, kk: shard.search(q, kk, normalise_queries=True)[:2],
                                                o, k, lo, world * B)

        for i in range(3):
            r = step(i)
        want = full.search(o, k, normalise_queries=True)
        same = float((r[0] == want[0]).float().mean())
        score_err = float((r[1] - want[1]).abs().max())
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(steps):
            r = step(i)
        e1.record()
        barrier()
        ms = sync_max(e0.elapsed_time(e1)) / steps
        e0.record()
        for i in range(steps):
            r = v.shard.search_sharded_store(lambda q, kk: shard.search(q, kk, normalise_queries=True)[:2],
                                             o, k, lo, world * B)
        e1.record()
        barrier()
        ms_s = sync_max(e0.elapsed_time(e1)) / steps
        out.update({"workload": "ViT-B/16 embed 256 frames/GPU + cosine top-10 vs a 1M-row store row-sharded over the "
                                "ranks (NCCL: all-gather queries, local top-10, all-gather lists, merge)",
                    "ms_per_step": ms, "frames_per_s": world * B / ms * 1e3, "search_only_ms": ms_s,
                    "search_queries_per_s": world * B / ms_s * 1e3, "rows_per_rank": hi - lo,
                    "idx_mismatch_fraction_vs_replicated": sync_max(1.0 - same), "score_max_abs_diff": sync_max(score_err)})
    if rank == 0:
        print(json.dumps(out), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
