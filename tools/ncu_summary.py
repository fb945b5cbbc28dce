"""Summarise ncu outputs into small text files for profiles/ (run in the build container).
  python tools/ncu_summary.py launches gpurun_out/launches.csv > profiles/rNN_launches.txt
  python tools/ncu_summary.py full gpurun_out/prof.ncu-rep   > profiles/rNN_ncu_x.txt
  python tools/ncu_summary.py inseq gpurun_out/inseq.csv     > profiles/rNN_ncu_in_sequence.txt
"""
import collections, csv, re, subprocess, sys

KEYS = ["gpu__time_duration.sum", "sm__cycles_elapsed.avg.per_second",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct",
        "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
        "launch__shared_mem_per_block_dynamic", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum"]


def launches(path):
    lines = [l for l in open(path) if not l.startswith("==")]
    agg = collections.OrderedDict()
    n = 0
    for row in csv.DictReader(lines):
        if row.get("Metric Name") != "gpu__time_duration.sum":
            continue
        name = re.sub(r"\(.*", "", row["Kernel Name"])  # drop the argument list, keep <...>
        v = float(row["Metric Value"].replace(",", ""))
        v = v / 1e3 if row["Metric Unit"] == "ns" else v * 1e3 if row["Metric Unit"] == "ms" else v
        a = agg.setdefault(name, [0, 0.0])
        a[0] += 1
        a[1] += v
        n += 1
    tot = sum(v[1] for v in agg.values())
    print(f"# ncu --metrics gpu__time_duration.sum --clock-control none  ({n} launches, {tot/1e3:.2f} ms total; "
          "cold-cache, serialised: compare SHARES)")
    print(f"{'share':>7} {'count':>6} {'avg_us':>9}  kernel")
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"{v[1]/tot*100:6.2f}% {v[0]:6d} {v[1]/v[0]:9.1f}  {k[:120]}")


def full(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    r = list(csv.reader(out.splitlines()))
    hdr, units = r[0], r[1]
    ki = hdr.index("Kernel Name")
    for row in r[2:]:
        print("kernel:", row[ki][:150])
        for i, h in enumerate(hdr):
            if h in KEYS:
                print(f"  {h:72s} {row[i]:>18s} {units[i]}")
    src = subprocess.run(["ncu", "-i", path, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(src.splitlines()))
    cur = None
    kern = []
    for rr in rows:
        if rr and rr[0] == "Kernel Name":
            cur = {"name": rr[1], "hdr": None, "rows": []}
            kern.append(cur)
        elif cur is not None and cur["hdr"] is None:
            cur["hdr"] = rr
        elif cur is not None:
            cur["rows"].append(rr)
    seen = set()
    for k in kern:
        if k["name"] in seen:
            continue
        seen.add(k["name"])
        h = k["hdr"]
        si, so = h.index("# Samples"), h.index("Source")
        body = [x for x in k["rows"] if len(x) > si and x[si].isdigit()]
        tot = sum(int(x[si]) for x in body) or 1
        print(f"top stall-sample instructions: {k['name'][:100]}  ({tot} samples)")
        for x in sorted(body, key=lambda x: -int(x[si]))[:12]:
            print(f"  {int(x[si])/tot*100:5.1f}%  {x[so].strip()[:100]}")


def inseq(path):
    """--cache-control none pass (kernels run on the caches their producers left): per kernel the
    mean duration, DRAM bytes read / written and tensor-pipe activity, and the totals."""
    lines = [l for l in open(path) if not l.startswith("==")]
    per = collections.OrderedDict()
    for row in csv.DictReader(lines):
        d = per.setdefault(row["ID"], {"name": re.sub(r"\(.*", "", row["Kernel Name"]).replace("void ", "").replace("vitk::", "")})
        d[row["Metric Name"]] = float(row["Metric Value"].replace(",", ""))
    agg = collections.OrderedDict()
    for d in per.values():
        a = agg.setdefault(d["name"], [0, 0.0, 0.0, 0.0, 0.0])
        a[0] += 1
        a[1] += d.get("gpu__time_duration.sum", 0.0) / 1e3
        a[2] += d.get("dram__bytes_read.sum", 0.0) / 1e6
        a[3] += d.get("dram__bytes_write.sum", 0.0) / 1e6
        a[4] += d.get("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", 0.0)
    print(f"{'count':>5} {'avg_us':>8} {'rd_MB':>8} {'wr_MB':>8} {'tensor%':>8}  kernel")
    for k, a in agg.items():
        n = a[0]
        print(f"{n:5d} {a[1]/n:8.1f} {a[2]/n:8.1f} {a[3]/n:8.1f} {a[4]/n:8.1f}  {k[:110]}")
    rd = sum(a[2] for a in agg.values()) / 1e3
    wr = sum(a[3] for a in agg.values()) / 1e3
    print(f"# total DRAM read {rd:.2f} GB, write {wr:.2f} GB over these {len(per)} launches "
          f"({sum(a[1] for a in agg.values())/1e3:.2f} ms of kernel time)")


if __name__ == "__main__":
    {"launches": launches, "full": full, "inseq": inseq}[sys.argv[1]](sys.argv[2])
