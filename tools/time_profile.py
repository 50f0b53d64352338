#!/usr/bin/env python3
"""Time fb_profile for one model on the AC/GC/BM grid (development aid).  usage: time_profile.py MODEL [n]"""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fairify_b200 import engine
from fairify_b200.domains import DOMAINS, STRESS_THRESHOLD
name = sys.argv[1]; n = int(sys.argv[2]) if len(sys.argv) > 2 else 2_000_000
z = np.load(os.path.join(os.path.dirname(__file__), "..", "tests", "golden", "weights.npz"))
L = int(z[f"{name}/L"]); w = [z[f"{name}/w{l}"] for l in range(L)]; b = [z[f"{name}/b{l}"] for l in range(L)]
m = engine.Model(w, b); D = DOMAINS[name.split("-")[0]]
g = engine.Grid(D.columns, D.ranges, 4 if D.name == "AC" else STRESS_THRESHOLD[D.name])
n = min(n, len(g)); boxes = g.boxes(0, n)
for it in range(3):
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record(); c = m.profile(boxes, S=1000, seed=7); e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
pk = max(engine.fp64_peak(4096)[0], engine.fp64_peak(4096)[0])
tf = n * 1000 * (2 * m.P + m.N) / (ms * 1e-3) / 1e12
print(name, "spec" if m.specialised else "generic", os.environ.get("FAIRIFY_B200_TILED_PROFILE", "0"), "n", n, "ms", round(ms, 2),
      "evals/s", f"{n / ms * 1e3:.3e}", "TF", round(tf, 2), "frac", round(tf / pk, 3), "chk", int(c.sum().item()))
