#!/usr/bin/env python3
"""Time fb_partition_screen next to fb_profile for one model on its stress grid (development aid).
usage: time_screen.py MODEL [n]"""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fairify_b200 import engine
from fairify_b200.domains import DOMAINS, STRESS_THRESHOLD
name = sys.argv[1]; n = int(sys.argv[2]) if len(sys.argv) > 2 else 100_000
z = np.load(os.path.join(os.path.dirname(__file__), "..", "tests", "golden", "weights.npz"))
L = int(z[f"{name}/L"]); w = [z[f"{name}/w{l}"] for l in range(L)]; b = [z[f"{name}/b{l}"] for l in range(L)]
m = engine.Model(w, b); D = DOMAINS[name.split("-")[0]]
g = engine.Grid(D.columns, D.ranges, 4 if D.name == "AC" else STRESS_THRESHOLD[D.name])
n = min(n, len(g)); boxes = g.boxes(0, n)
pa = D.columns.index(D.pa[0])
hb = boxes[:, pa].cpu().numpy(); width = float((hb[:, 1] - hb[:, 0] + 1).mean())
ev = lambda: torch.cuda.Event(enable_timing=True)
for it in range(2):
    a, c = ev(), ev(); a.record(); m.profile(boxes, S=1000, seed=7); c.record(); torch.cuda.synchronize(); pms = a.elapsed_time(c)
    a, c = ev(), ev(); a.record(); r = m.partition_screen(boxes, pa, S=1000, seed=7); c.record(); torch.cuda.synchronize(); sms = a.elapsed_time(c)
ce = int((r[0] >= 0).sum().item())
fl = n * 1000 * width * (2 * m.P + m.N)
print(name, "spec" if m.specialised else "generic", "n", n, "pa", D.pa[0], "mean pa width", round(width, 2), "profile ms", round(pms, 2),
      "screen ms", round(sms, 2), "screen FP32 TFLOP/s (no early exit)", round(fl / (sms * 1e-3) / 1e12, 2), "partitions with CE", ce)
