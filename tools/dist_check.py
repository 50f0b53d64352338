#!/usr/bin/env python3
"""torchrun --nproc-per-node N tools/dist_check.py : NCCL run of prune_grid_sharded, checked on rank 0 against a
single-GPU run of the whole grid (development aid for the multi-GPU path)."""
import os, sys
import numpy as np, torch, torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fairify_b200 import engine
from fairify_b200.dist import prune_grid_sharded
from fairify_b200.domains import DOMAINS

local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
z = np.load(os.path.join(os.path.dirname(__file__), "..", "tests", "golden", "weights.npz"))
w = [z[f"AC-1/w{l}"] for l in range(3)]; b = [z[f"AC-1/b{l}"] for l in range(3)]
m = engine.Model(w, b); D = DOMAINS["AC"]
grid = engine.Grid(D.columns, D.ranges, 8)          # 67 392 boxes
v, d, c = prune_grid_sharded(m, grid, S=200, seed=3)
if dist.get_rank() == 0:
    one = m.prune(grid.boxes(), S=200, seed=3, want=("cand", "dead", "verdict"))
    ok = (torch.equal(v, one.t["verdict"]) and torch.equal(d, one.t["dead"]) and torch.equal(c, one.t["cand"]))
    print("dist_check", "OK" if ok else "MISMATCH", "world", dist.get_world_size(), "n", len(grid))
dist.destroy_process_group()
