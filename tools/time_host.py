#!/usr/bin/env python3
"""Time fb_prune_host against fb_prune on the bench workload (development aid).  usage: time_host.py [n]"""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fairify_b200 import engine
from fairify_b200.domains import DOMAINS
n = int(sys.argv[1]) if len(sys.argv) > 1 else 10_000_000
z = np.load(os.path.join(os.path.dirname(__file__), "..", "tests", "golden", "weights.npz"))
w = [z[f"AC-1/w{l}"] for l in range(3)]; b = [z[f"AC-1/b{l}"] for l in range(3)]
m = engine.Model(w, b); D = DOMAINS["AC"]; g = engine.Grid(D.columns, D.ranges, 4)
boxes = g.boxes(0, n)
lo = [D.ranges[c][0] for c in D.columns]; hi = [D.ranges[c][1] for c in D.columns]
tol = torch.as_tensor(m.tolerance(lo, hi))
hb = boxes.cpu().pin_memory(); hout = m.alloc_host_outputs(n, engine.DEFAULT_OUTPUTS)
ev = lambda: torch.cuda.Event(enable_timing=True)
for it in range(3):
    a, c = ev(), ev(); a.record(); r = m.prune(boxes, S=1000, seed=1, tol=tol.cuda()); c.record(); torch.cuda.synchronize()
    dev_ms = a.elapsed_time(c)
for it in range(3):
    a, c = ev(), ev(); a.record(); m.prune_host(hb, S=1000, seed=1, tol=tol, out=hout); c.record(); torch.cuda.synchronize()
    host_ms = a.elapsed_time(c)
print("slice", os.environ.get("FAIRIFY_B200_HOST_SLICE", "default"), "n", n, "device ms", round(dev_ms, 2), "host ms", round(host_ms, 2))
