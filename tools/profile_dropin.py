#!/usr/bin/env python3
"""cProfile of the drivers' per-partition calls on the drop-in utils package (development aid; needs a GPU)."""
import contextlib, cProfile, importlib, io, os, pstats, sys, time
import numpy as np, pandas as pd
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fairify_b200.domains import DOMAINS
from utils.input_partition import partition, partitioned_ranges
import utils.prune as UP
from utils.verif_utils import get_y_pred
name = sys.argv[1] if len(sys.argv) > 1 else "AC-1"
z = np.load(os.path.join(os.path.dirname(__file__), "..", "tests", "golden", "weights.npz"))
L = int(z[f"{name}/L"]); w = [z[f"{name}/w{l}"] for l in range(L)]; b = [z[f"{name}/b{l}"] for l in range(L)]
D = DOMAINS[name.split("-")[0]]
fn = {"AC": UP.sound_prune, "GC": UP.sound_prune_german, "BM": UP.sound_prune_bank}[D.name]
mod = importlib.import_module(f"utils.{name}-Model-Functions")
df = pd.DataFrame(np.zeros((2, D.d + 1)), columns=D.columns + [D.label])
rd = D.copy_ranges()
p_list = partitioned_ranges(rd.keys(), D.pa, partition(rd, D.threshold), rd)
def loop(count):
    for pp in p_list[:count]:
        out = fn(df, w, b, 1000, mod.layer_net, pp)
        pr_w, pr_b = UP.prune_neurons(w, b, out[5])
        sim = out[7].to_numpy()
        get_y_pred(mod.net, w, b, sim); get_y_pred(mod.net, pr_w, pr_b, sim); get_y_pred(mod.net, w, b, sim[:200])
with contextlib.redirect_stdout(io.StringIO()):
    n = min(len(p_list), UP._window_size())          # whole batching windows, so the batched launch is amortised as in a run
    loop(n)
    t0 = time.perf_counter(); loop(n); dt = time.perf_counter() - t0
    pr = cProfile.Profile(); pr.enable(); loop(n); pr.disable()
print(f"{name}: {1e3 * dt / n:.3f} ms per partition over {n} partitions")
pstats.Stats(pr).sort_stats("cumulative").print_stats(22)
