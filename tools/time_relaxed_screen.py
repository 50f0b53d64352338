import sys, os, numpy as np, torch
sys.path.insert(0, os.getcwd())
from fairify_b200 import engine
from fairify_b200.domains import DOMAINS
z = np.load("tests/golden/weights.npz")
D = DOMAINS["AC"]
g = engine.Grid(D.columns, D.ranges, 6)
for name in sys.argv[1:]:
    L = int(z[f"{name}/L"]); w = [z[f"{name}/w{l}"] for l in range(L)]; b = [z[f"{name}/b{l}"] for l in range(L)]
    m = engine.Model(w, b)
    n = 20000
    bx = g.boxes(len(g)//3, n)
    for mode in ("relaxed", "strict"):
        kw = dict(ra_col=D.columns.index("age"), ra_threshold=5) if mode == "relaxed" else {}
        for _ in range(2):
            a, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(); r = m.partition_screen(bx, D.columns.index("race"), S=1000, seed=1, **kw); e.record(); torch.cuda.synchronize()
        ce = r[0].cpu().numpy()
        for _ in range(2):
            a2, e2 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a2.record(); cert = m.certified(bx, **kw)
            r2 = m.partition_screen(bx, D.columns.index("race"), S=1000, seed=1, skip=cert, **kw); e2.record(); torch.cuda.synchronize()
        assert all(torch.equal(x, y) for x, y in zip(r, r2))
        print(name, mode, f"{n/(a.elapsed_time(e)*1e-3):.4g} partitions/s", "ce frac", (ce>=0).mean(),
              f"| skipping certified ({float(cert.float().mean()):.3f}): {n/(a2.elapsed_time(e2)*1e-3):.4g} partitions/s", flush=True)
