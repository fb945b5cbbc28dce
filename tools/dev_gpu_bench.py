"""Developer diagnostic: per-kernel device times at the bench shapes (bench.time_kernels)."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import vit_research_b200 as v
import bench

lib = v.load()
n = int(sys.argv[1]) if len(sys.argv) > 1 else 256
warm = int(os.environ.get("KB_WARM", "3")); reps = int(os.environ.get("KB_REPS", "10"))
print(json.dumps(bench.time_kernels(v, lib, torch, n, warm, reps), indent=1))
