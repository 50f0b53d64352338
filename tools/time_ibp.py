#!/usr/bin/env python3
"""Time fb_ibp (bounds + masks, no counts) for one model (development aid).  usage: time_ibp.py MODEL [n]"""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fairify_b200 import engine
from fairify_b200.domains import DOMAINS, STRESS_THRESHOLD
name = sys.argv[1]; n = int(sys.argv[2]) if len(sys.argv) > 2 else 4_000_000
z = np.load(os.path.join(os.path.dirname(__file__), "..", "tests", "golden", "weights.npz"))
L = int(z[f"{name}/L"]); w = [z[f"{name}/w{l}"] for l in range(L)]; b = [z[f"{name}/b{l}"] for l in range(L)]
m = engine.Model(w, b); D = DOMAINS[name.split("-")[0]]
g = engine.Grid(D.columns, D.ranges, 4 if D.name == "AC" else STRESS_THRESHOLD[D.name])
n = min(n, len(g)); boxes = g.boxes(0, n)
for it in range(3):
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record(); r = m.ibp(boxes, want=("b_dead", "verdict")); e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
print(name, "n", n, "ms", round(ms, 3), "evals/s", f"{n / ms * 1e3:.3e}", "FP64 ops/s", f"{n * (8 * m.P + 4 * m.N) / ms * 1e3:.3e}")
