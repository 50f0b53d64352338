#!/usr/bin/env python3
"""Per-model table of the pre-solver kernels on the GPU at hand (development aid / evidence for DESIGN.md section 5).

    python tools/profile_table.py > profiles/r2_model_table.md

For every reference model with weights in tests/golden/weights.npz: the activation-profile kernel (evals/s, % of the
measured FP64 peak, which kernel family ran) and the bounds kernel (evals/s, algorithmic FP64 ops/s)."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fairify_b200 import engine  # noqa: E402
from fairify_b200.domains import DOMAINS, STRESS_THRESHOLD  # noqa: E402

MMA_SPEC = {(13, 100, 100, 1), (13, 64, 64, 1), (16, 150, 100, 50, 1), (16, 64, 64, 1), (16, 64, 16, 1),
            (13, 64, 32, 16, 8, 4, 1), (16, 64, 32, 16, 8, 4, 1), (20, 64, 32, 16, 8, 4, 1)}      # fb_spec_lists.h


def timed(fn, reps=5):
    """median of `reps` timed calls after one warm-up call (the first call of a kernel also pays its module load)"""
    ts = []
    for k in range(reps + 1):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        if k:
            ts.append(a.elapsed_time(b))
    return sorted(ts)[len(ts) // 2]


def main():
    z = np.load(os.path.join(os.path.dirname(__file__), "..", "tests", "golden", "weights.npz"))
    peak = max(engine.fp64_peak(4096)[0] for _ in range(3))
    names = [f"AC-{k}" for k in range(1, 13)] + [f"GC-{k}" for k in range(1, 6)] + [f"BM-{k}" for k in range(1, 9)]
    if len(sys.argv) > 1:
        names = sys.argv[1:]
    print(f"B200, measured DFMA peak {peak:.2f} TFLOP/s; profile = 1000 Philox samples per partition\n")
    print("| model | layers | weights | profile kernel | partitions | profile evals/s | % of FP64 peak | bounds evals/s | bounds FP64 ops/s |")
    print("|---|---|---|---|---|---|---|---|---|")
    for name in names:
        L = int(z[f"{name}/L"])
        w = [z[f"{name}/w{l}"] for l in range(L)]; b = [z[f"{name}/b{l}"] for l in range(L)]
        m = engine.Model(w, b)
        D = DOMAINS[name.split("-")[0]]
        g = engine.Grid(D.columns, D.ranges, 4 if D.name == "AC" else STRESS_THRESHOLD[D.name])
        n = min(len(g), max(20000, int(6e7 / m.P)))
        boxes = g.boxes(0, n)
        ms = timed(lambda: m.profile(boxes, S=1000, seed=7))
        tf = n * 1000 * (2 * m.P + m.N) / (ms * 1e-3) / 1e12
        kind = "DMMA" if tuple(int(x) for x in m.dims) in MMA_SPEC else ("FMA" if m.specialised else "runtime-shape")
        nb = min(len(g), 4_000_000)
        bx = g.boxes(0, nb)
        ims = timed(lambda: m.ibp(bx, want=("b_dead", "verdict")))
        print(f"| {name} | {'-'.join(str(int(x)) for x in m.dims)} | {m.P} | {kind} | {n} | {n / ms * 1e3:.3e} | {100 * tf / peak:.1f} | "
              f"{nb / ims * 1e3:.3e} | {nb * (8 * m.P + 4 * m.N) / ims * 1e3:.3e} |")
        del m, boxes, bx


if __name__ == "__main__":
    main()
