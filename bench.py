#!/usr/bin/env python3
"""Headline benchmark: partition IBP+prune evaluations per second (BASELINE.json `metric`).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

One *evaluation* = the whole pre-solver stage for one (partition, model) pair: S = 1000 samples drawn in
the box, activation profile, interval bounds (round-to-nearest + outward-rounded), dead-neuron masks and
the bound verdict -- what `sound_prune` (reference utils/prune.py:671-707) does per partition, minus
its Z3 neuron queries.  A *step* is one pass over this rank's batch of partitions.

Workload (config.workload): the scaled stress/AC experiment of BASELINE.json -- AC-1 weights (the AC-14
.h5 is not in the reference; same 13-16-8-1 stack), 10^7 synthetic partitions per GPU taken from the Adult
grid with PARTITION_THRESHOLD 4 (161.9 M boxes; rank r owns canonical indices [r*10^7, (r+1)*10^7)).
Weak scaling (default): per-GPU work is fixed.  `--scaling strong` reads BASELINE config 4 literally: --partitions in
total, split into contiguous ranges over the GPUs.  With N > 1 every step's packed records -- verdict | dead mask |
candidate mask | first counterexample sample, written by fb_ibp itself -- are all-gathered over NCCL on a second stream,
so that the gather of step k overlaps the kernels of step k+1; the last gather is inside the timed region, and rank 0
checks the gathered bytes of a foreign shard against its own recomputation of that shard.

`value`   device-resident boxes in, device-resident masks out (kernel-only), CUDA events, max over ranks.
`e2e`     the same batch through fb_prune_host with pinned HOST buffers: boxes copied in, masks + verdicts +
          counts copied out, every step.
`--impl reference`  the CPU arm: the oracle port (the reference is Python and does not travel to the GPU
          box) on all host threads, bounded sample per step.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "partition IBP+prune evals/sec"
UNIT = "evals/s"
S_SAMPLES = 1000
GRID_T = 4
SEED = 20260101


def load_ac1():
    z = np.load(os.path.join(ROOT, "tests", "golden", "weights.npz"))
    L = int(z["AC-1/L"])
    return [z[f"AC-1/w{l}"] for l in range(L)], [z[f"AC-1/b{l}"] for l in range(L)]


def workload_name(n_per_gpu):
    return (f"stress/AC scaled: AC-1 weights (13-16-8-1; AC-14 .h5 absent from the reference), {n_per_gpu} synthetic "
            f"partitions per GPU = Adult grid T={GRID_T} canonical range [rank*n, (rank+1)*n), S={S_SAMPLES} Philox samples")


def config_for(n, world, scaling):
    """`config` of the JSON line -- the same dictionary for the GPU arm and the CPU (reference) arm."""
    per = n if scaling == "weak" else f"{n} in total, split evenly"
    return {"workload": workload_name(n) if scaling == "weak" else workload_name(n).replace("per GPU", "in total (strong scaling: split over the GPUs)"),
            "model": "AC-1", "partitions_per_gpu": per, "samples": S_SAMPLES,
            "l2": "inputs (boxes, 104 B x n) and outputs exceed the 126 MB L2; no flush needed",
            "parallelism": f"contiguous partition range per GPU x{world}" +
                           (", all-gather of the packed verdict / mask / counterexample records per step, overlapped with the next step" if world > 1 else "")}


# ---------------------------------------------------------------------------------------------------
# clocks sampler (B200_PROFILING.md recipe)
# ---------------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "200", "-i", str(self.index)], stdout=subprocess.PIPE, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))
            except ValueError:
                continue
            for nm, val in zip(names, f[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


# ---------------------------------------------------------------------------------------------------
# CPU arm (oracle port) -- test infrastructure used as the reported baseline, never as the product
# ---------------------------------------------------------------------------------------------------
def cpu_boxes(first, count):
    from oracle import oracle as O
    from fairify_b200.domains import DOMAINS
    D = DOMAINS["AC"]
    lo = [D.ranges[c][0] for c in D.columns]; hi = [D.ranges[c][1] for c in D.columns]
    ko = [D.columns.index(k) for k in D.ranges.keys()]
    return O.grid_boxes(lo, hi, GRID_T, ko, first, count)


def cpu_rate(target_seconds, fixed_sample=None):
    """evals/s of the oracle port on all host threads, on a bounded slice of the same workload."""
    from oracle import oracle as O
    from fairify_b200.domains import DOMAINS
    O.set_threads(len(os.sched_getaffinity(0)))        # all host threads, whatever OMP_NUM_THREADS torchrun exported
    w, b = load_ac1()
    net = O.Net(w, b)
    D = DOMAINS["AC"]
    lo = [D.ranges[c][0] for c in D.columns]; hi = [D.ranges[c][1] for c in D.columns]
    tol = O.sliver_tolerance(net, lo, hi)
    if fixed_sample is None:
        probe = 2000
        boxes = cpu_boxes(0, probe)
        t0 = time.perf_counter(); O.prune_batch(net, boxes, SEED, 0, S_SAMPLES, tol); dt = time.perf_counter() - t0
        fixed_sample = int(max(probe, min(5_000_000, probe * target_seconds / max(dt, 1e-6))))
    boxes = cpu_boxes(0, fixed_sample)
    t0 = time.perf_counter(); O.prune_batch(net, boxes, SEED, 0, S_SAMPLES, tol); dt = time.perf_counter() - t0
    return fixed_sample / dt, fixed_sample, dt, O.num_threads()


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    from oracle import oracle as O
    O.build()
    rate0, sample, _, cores = cpu_rate(8.0)            # calibrates the per-step sample (about 8 s of CPU work)
    for _ in range(args.warmup):
        cpu_rate(0, fixed_sample=max(1000, sample // 8))
    t_tot = 0.0
    for _ in range(args.steps):
        _, _, dt, _ = cpu_rate(0, fixed_sample=sample)
        t_tot += dt
    value = sample * args.steps / t_tot
    sample_txt = f"first {sample} partitions of the workload per step (full pre-solver, S={S_SAMPLES})"
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * t_tot / args.steps, "higher_is_better": True, "scaling": args.scaling,
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": config_for(args.partitions, args.gpus, args.scaling),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample_txt},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "note": "CPU arm = oracle/fairify_oracle.c (C restatement of the reference's Python/numpy path, OpenMP over "
                "partitions); the Python reference itself measured 0.3-7 partitions/s/core (BASELINE.md)",
    }
    print(json.dumps(line), flush=True)
    return 0


# ---------------------------------------------------------------------------------------------------
# GPU arm
# ---------------------------------------------------------------------------------------------------
def run_gpu(args):
    import torch
    import torch.distributed as dist
    from fairify_b200 import engine
    from fairify_b200.dist import shard_range
    from fairify_b200.domains import DOMAINS

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the product path has no CPU fallback; use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    saved_stdout = None
    if world > 1:
        # NCCL prints its version banner on stdout when the first communicator comes up; stdout must carry the one JSON
        # line and nothing else, so file descriptor 1 points at stderr until that line is printed
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        sys.stdout.flush()
        saved_stdout = os.dup(1)
        os.dup2(2, 1)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    w, b = load_ac1()
    model = engine.Model(w, b)
    D = DOMAINS["AC"]
    grid = engine.Grid(D.columns, D.ranges, GRID_T)
    if args.scaling == "strong":
        n_total = args.partitions                          # BASELINE config 4 read literally: 10^7 partitions in all
        first, n = shard_range(n_total, rank, world)
    else:
        n = args.partitions
        first = rank * n
        n_total = world * n
    if n_total > len(grid):
        raise SystemExit(f"grid of {len(grid)} boxes cannot give {n_total} partitions")
    nmax = shard_range(n_total, 0, world)[1] if args.scaling == "strong" else n
    lo = [D.ranges[c][0] for c in D.columns]; hi = [D.ranges[c][1] for c in D.columns]
    tol = torch.as_tensor(model.tolerance(lo, hi)).cuda()

    # inputs resident in HBM before the timed region
    ev = lambda: torch.cuda.Event(enable_timing=True)
    g0, g1 = ev(), ev()
    g0.record(); boxes = grid.boxes(first, n); g1.record(); torch.cuda.synchronize()
    boxgen_ms = g0.elapsed_time(g1)
    counts32 = torch.empty((n, model.N), dtype=torch.int32, device="cuda")
    # per-partition record [verdict | dead mask | candidate mask | first counterexample sample], written by fb_ibp itself;
    # two copies so that the all-gather of step k (on its own stream) overlaps the kernels of step k+1
    NW = model.NW
    rec_w = 2 * NW + 2
    outs, po = model._alloc_outputs(n, engine.DEFAULT_OUTPUTS + ("record",))
    packed = [outs["record"], torch.empty_like(outs["record"])]
    pos = [po, type(po)()]
    for k_ in engine.ALL_OUTPUTS:
        setattr(pos[1], k_, getattr(po, k_))
    pos[1].record = packed[1].data_ptr()
    # (shards may differ by one partition under strong scaling: the collective works on nmax rows per rank)
    send = [torch.zeros((nmax, rec_w), dtype=torch.int32, device="cuda") for _ in range(2)] if (world > 1 and nmax != n) else None
    gathered = [torch.empty((world * nmax, rec_w), dtype=torch.int32, device="cuda") for _ in range(2)] if world > 1 else None
    comm = torch.cuda.Stream() if world > 1 else None
    gather_done = [None, None]

    import ctypes as C
    from fairify_b200 import _lib
    L = _lib.lib()
    stream = lambda: C.c_void_p(torch.cuda.current_stream().cuda_stream)
    ptr = lambda t: C.c_void_p(t.data_ptr())

    step_no = [0]

    def step(timers=None):
        k = step_no[0] & 1
        step_no[0] += 1
        if timers:
            timers[0].record()
        _lib.check(L.fb_profile(model._h, n, ptr(boxes), None, S_SAMPLES, C.c_uint64(SEED), first, ptr(counts32), stream()))
        if timers:
            timers[1].record()
        if world > 1 and gather_done[k] is not None:
            torch.cuda.current_stream().wait_event(gather_done[k])      # the gather of two steps ago still reads packed[k]
        _lib.check(L.fb_ibp(model._h, n, ptr(boxes), ptr(counts32), ptr(tol), C.byref(pos[k]), stream()))
        if timers:
            timers[2].record()
        if world > 1:
            ready = torch.cuda.Event()
            ready.record()
            comm.wait_event(ready)
            with torch.cuda.stream(comm):
                src = packed[k]
                if send is not None:
                    send[k][:n].copy_(packed[k]); src = send[k]
                dist.all_gather_into_tensor(gathered[k], src)
                gather_done[k] = torch.cuda.Event()
                gather_done[k].record()

    def drain():
        if world > 1:
            torch.cuda.current_stream().wait_stream(comm)

    for _ in range(args.warmup):
        step()
    drain()
    peak_tf, _ = engine.fp64_peak(4096)
    peak_tf = max(peak_tf, engine.fp64_peak(4096)[0])

    sampler = ClockSampler(local)
    timers = [[ev(), ev(), ev()] for _ in range(args.steps)]
    barrier()
    if rank == 0:
        sampler.start()
    t0, t1 = ev(), ev()
    t0.record()
    for k in range(args.steps):
        step(timers[k])
    drain()
    t1.record()
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    ms = torch.tensor([t0.elapsed_time(t1)], device="cuda", dtype=torch.float64)
    prof_ms = torch.tensor([sum(t[0].elapsed_time(t[1]) for t in timers) / args.steps], device="cuda", dtype=torch.float64)
    ibp_ms = sum(t[1].elapsed_time(t[2]) for t in timers) / args.steps
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        dist.all_reduce(prof_ms, op=dist.ReduceOp.MAX)
    total_ms = float(ms.item())
    value = n_total * args.steps / (total_ms * 1e-3)

    # end to end: host boxes in, host masks / verdicts / counts out, every step
    hboxes = boxes.cpu().pin_memory()
    hout = model.alloc_host_outputs(n, engine.DEFAULT_OUTPUTS)
    htol = tol.cpu()
    model.prune_host(hboxes, S=S_SAMPLES, seed=SEED, first_index=first, tol=htol, out=hout)     # warm-up
    barrier()
    e0, e1 = ev(), ev()
    e0.record()
    for _ in range(args.steps):
        model.prune_host(hboxes, S=S_SAMPLES, seed=SEED, first_index=first, tol=htol, out=hout)
    e1.record()
    barrier()
    ems = torch.tensor([e0.elapsed_time(e1)], device="cuda", dtype=torch.float64)
    if world > 1:
        dist.all_reduce(ems, op=dist.ReduceOp.MAX)
    e2e_value = n_total * args.steps / (float(ems.item()) * 1e-3)
    h2d = hboxes.numel() * 4 + htol.numel() * 8
    d2h = sum(t.numel() * t.element_size() for t in hout.values())
    # the device path and the host path must agree byte for byte
    same = all(bool((hout[k] == outs[k].cpu()).all()) for k in ("dead", "verdict", "counts"))

    # IBP-only and pair-screen side numbers (not the headline)
    i0, i1 = ev(), ev()
    i0.record()
    _lib.check(L.fb_ibp(model._h, n, ptr(boxes), None, ptr(tol), C.byref(po), stream()))
    i1.record(); torch.cuda.synchronize()
    ibp_only_ms = i0.elapsed_time(i1)
    pa_col = D.columns.index("sex")
    ns = min(n, 2_000_000)
    s0, s1 = ev(), ev()
    # (warm-up on the full size: the first call of a size also pays for growing the memory pools it draws from)
    model.partition_screen(boxes[:ns], pa_col, S=S_SAMPLES, seed=SEED, first_index=first)
    torch.cuda.synchronize()
    s0.record()
    model.partition_screen(boxes[:ns], pa_col, S=S_SAMPLES, seed=SEED, first_index=first)
    s1.record(); torch.cuda.synchronize()
    screen_ms = s0.elapsed_time(s1)
    # the same screen told to leave alone the boxes the bounds kernel certifies one-signed (no pair of points of such a
    # box can differ in class); the bounds pass that finds them is inside the timed region
    s2, s3 = ev(), ev()
    cert = model.certified(boxes[:ns]); del cert
    torch.cuda.synchronize()
    s2.record()
    cert = model.certified(boxes[:ns])
    model.partition_screen(boxes[:ns], pa_col, S=S_SAMPLES, seed=SEED, first_index=first, skip=cert)
    s3.record(); torch.cuda.synchronize()
    screen_skip_ms = s2.elapsed_time(s3)
    certified_frac = float(cert.float().mean().item())
    del cert

    # N > 1: what the all-gather delivered for a foreign shard must be what this rank computes for that shard itself
    gather_checked = None
    if world > 1 and rank == 0:
        last = (step_no[0] - 1) & 1
        f1, c1 = (shard_range(n_total, 1, world) if args.scaling == "strong" else (n, n))
        chk = min(c1, 200_000)
        bx1 = grid.boxes(f1, chk)
        r1 = model.prune(bx1, S=S_SAMPLES, seed=SEED, first_index=f1, tol=tol, want=("record",))
        gather_checked = bool(torch.equal(gathered[last][nmax: nmax + chk], r1.t["record"])) and \
            bool(torch.equal(gathered[last][:min(n, chk)], packed[last][:min(n, chk)]))
        if not gather_checked:
            raise SystemExit("bench.py: the gathered records differ from a recomputation of the foreign shard")

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return 0

    def timed(fn, reps=2):
        for _ in range(reps):
            a_, b_ = ev(), ev()
            a_.record(); fn(); b_.record(); torch.cuda.synchronize()
        return a_.elapsed_time(b_)

    z = np.load(os.path.join(ROOT, "tests", "golden", "weights.npz"))

    def load(name):
        Lk = int(z[f"{name}/L"])
        return engine.Model([z[f"{name}/w{l}"] for l in range(Lk)], [z[f"{name}/b{l}"] for l in range(Lk)])

    # BASELINE configs 1 and 2: the src/GC grid x GC-1 (201 partitions) and the src/AC grid x AC-1 (16 000): the full
    # pre-solver + sampled counterexample screen in one batch, and the same through the drop-in `utils` API the
    # unchanged drivers call (sound_prune* + prune_neurons + 3 x get_y_pred per partition, src/AC/Verify-AC.py:117-274),
    # whose window batching answers each call from one launch per 4096 partitions
    src_grids = {}
    try:
        import contextlib, io
        import pandas as pd
        import importlib
        from utils.input_partition import partition, partitioned_ranges
        import utils.prune as UP
        from utils.verif_utils import get_y_pred
        for tag, name, fn in (("src/GC x GC-1", "GC-1", UP.sound_prune_german), ("src/AC x AC-1", "AC-1", UP.sound_prune)):
            Dm = DOMAINS[name.split("-")[0]]
            mk = load(name)
            gk = engine.Grid(Dm.columns, Dm.ranges, Dm.threshold)
            bk = gk.boxes()
            tk = torch.as_tensor(mk.tolerance([Dm.ranges[c][0] for c in Dm.columns], [Dm.ranges[c][1] for c in Dm.columns])).cuda()
            pak = Dm.columns.index(Dm.pa[0])
            ms_batch = timed(lambda: (mk.prune(bk, S=S_SAMPLES, seed=SEED, tol=tk, want=engine.DEFAULT_OUTPUTS + ("record",)),
                                      mk.partition_screen(bk, pak, S=S_SAMPLES, seed=SEED)))
            wk = [z[f"{name}/w{l}"] for l in range(mk.L)]; bkk = [z[f"{name}/b{l}"] for l in range(mk.L)]
            mod = importlib.import_module(f"utils.{name}-Model-Functions")
            df = pd.DataFrame(np.zeros((2, Dm.d + 1)), columns=Dm.columns + [Dm.label])
            rd = Dm.copy_ranges()
            t0 = time.perf_counter()
            p_list = partitioned_ranges(rd.keys(), Dm.pa, partition(rd, Dm.threshold), rd)
            t_grid = time.perf_counter() - t0
            count = min(len(p_list), 4096)              # one full batching window
            np.random.seed(1)
            with contextlib.redirect_stdout(io.StringIO()):
                t0 = time.perf_counter()
                for pp in p_list[:count]:
                    out = fn(df, wk, bkk, S_SAMPLES, mod.layer_net, pp)
                    pr_w, pr_b = UP.prune_neurons(wk, bkk, out[5])
                    sim = out[7].to_numpy()
                    get_y_pred(mod.net, wk, bkk, sim); get_y_pred(mod.net, pr_w, pr_b, sim); get_y_pred(mod.net, wk, bkk, sim[:200])
                t_drop = time.perf_counter() - t0
            src_grids[tag] = {"partitions": len(gk), "batch_prune_plus_screen_ms": ms_batch,
                              "batch_evals_per_s": len(gk) / (ms_batch * 1e-3),
                              "dropin_partitions_timed": count, "dropin_ms_per_partition": 1e3 * t_drop / count,
                              "dropin_partitioned_ranges_s": t_grid,
                              "reference_s_per_partition": "0.17-2.9 (BASELINE.md section 2, same calls in the Python reference)"}
    except Exception as exc:
        src_grids = {"error": repr(exc)}

    # BASELINE config 5: the relaxed / targeted / targeted2 Adult grids (3 290 112 / 205 632 / 1 096 704 partitions,
    # PARTITION_THRESHOLD 6) on every Adult model whose .h5 exists (AC-1..12): full pre-solver with the packed record,
    # and on a window of the relaxed grid the relaxed counterexample screen (PA race, RA age +-5)
    ac_sweep = {}
    try:
        DA = DOMAINS["AC"]
        grids5 = {"relaxed": {}, "targeted": {"age": [30, 35]}, "targeted2": {"education": [9, 10]}}
        gs = {}
        for gname, override in grids5.items():
            rd = DA.copy_ranges()
            rd.update(override)
            gs[gname] = engine.Grid(DA.columns, rd, 6)
        tolA = None
        for k in range(1, 13):
            name = f"AC-{k}"
            mk = load(name)
            tk = torch.as_tensor(mk.tolerance([DA.ranges[c][0] for c in DA.columns], [DA.ranges[c][1] for c in DA.columns])).cuda()
            row = {"weights": mk.P}
            for gname, gk in gs.items():
                nk = len(gk) if mk.P < 4000 else min(len(gk), 400_000)       # (the four wide models: a window, extrapolate)
                bk = gk.boxes(0, nk)
                ms_k = timed(lambda: mk.prune(bk, S=S_SAMPLES, seed=SEED, tol=tk, want=("record",)), reps=1 if nk > 1_000_000 else 2)
                rate = nk / (ms_k * 1e-3)
                row[gname] = {"partitions": len(gk), "timed": nk, "evals_per_s": rate,
                              "fp64_frac": rate * S_SAMPLES * (2.0 * mk.P + mk.N) / 1e12 / peak_tf}
                del bk
            wk_ = min(len(gs["relaxed"]), 200_000)
            bk = gs["relaxed"].boxes(0, wk_)
            ms_r = timed(lambda: mk.partition_screen(bk, DA.columns.index("race"), S=S_SAMPLES, seed=SEED,
                                                     ra_col=DA.columns.index("age"), ra_threshold=5))
            row["relaxed_screen_partitions_per_s"] = wk_ / (ms_r * 1e-3)
            ra_, pa_ = DA.columns.index("age"), DA.columns.index("race")
            ms_rs = timed(lambda: mk.partition_screen(bk, pa_, S=S_SAMPLES, seed=SEED, ra_col=ra_, ra_threshold=5,
                                                      skip=mk.certified(bk, ra_col=ra_, ra_threshold=5)))
            row["relaxed_screen_skipping_certified_partitions_per_s"] = wk_ / (ms_rs * 1e-3)
            ac_sweep[name] = row
            del bk, mk
    except Exception as exc:
        ac_sweep = {"error": repr(exc)}

    # side numbers for BASELINE.json's BM-1..8 sweep (full pre-solver + sampled counterexample screen per model on a
    # window of the stress/BM grid); not part of the headline
    bm_sweep = {}
    try:
        from fairify_b200.domains import STRESS_THRESHOLD
        DB = DOMAINS["BM"]
        gb = engine.Grid(DB.columns, DB.ranges, STRESS_THRESHOLD["BM"])
        nb = min(20000, len(gb))
        bboxes = gb.boxes(0, nb)
        blo = [DB.ranges[c][0] for c in DB.columns]; bhi = [DB.ranges[c][1] for c in DB.columns]
        bpa = DB.columns.index(DB.pa[0])
        for k in range(1, 9):
            name = f"BM-{k}"
            mk = load(name)
            tk = torch.as_tensor(mk.tolerance(blo, bhi)).cuda()
            for it in range(2):                                   # first pass warms the kernels up
                b0, b1, b2 = ev(), ev(), ev()
                b0.record(); mk.prune(bboxes, S=S_SAMPLES, seed=SEED, tol=tk); b1.record()
                mk.partition_screen(bboxes, bpa, S=S_SAMPLES, seed=SEED); b2.record(); torch.cuda.synchronize()
            rate_k = nb / (b0.elapsed_time(b1) * 1e-3)
            bm_sweep[name] = {"weights": mk.P, "kernels": "specialised" if mk.specialised else "runtime-shape",
                              "prune_evals_per_s": rate_k,
                              "fp64_frac": rate_k * S_SAMPLES * (2.0 * mk.P + mk.N) / 1e12 / peak_tf,
                              "screen_partitions_per_s": nb / (b1.elapsed_time(b2) * 1e-3)}
    except Exception as exc:                                      # a side measurement must never cost the headline line
        bm_sweep = {"error": repr(exc)}

    P, N = model.P, model.N
    flops_profile = n * S_SAMPLES * (2.0 * P + N)                 # algorithmic FP64 flops of one profile launch
    achieved = flops_profile / (float(prof_ms.item()) * 1e-3) / 1e12
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "roofline_traffic.json")
    if os.path.exists(tpath):
        try:
            # measured DRAM bytes per partition (ncu --set full, profiles/r1_profile_kernel_ncu.csv) x this launch's size
            per_part = json.load(open(tpath)).get("profile_kernel_dram_bytes_per_partition")
            traffic = None if per_part is None else per_part * n
        except Exception:
            traffic = None
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm_peak = peaks.get("hbm_gbs", 6650.0)
    alg_bytes = n * (8.0 * model.d + 4.0 * N)                     # boxes in, uint32 counts out
    cpu_v, cpu_sample, cpu_dt, cores = cpu_rate(12.0)

    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": total_ms / args.steps, "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": config_for(args.partitions, world, args.scaling),
        "clocks": clocks,
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "host_equals_device": same},
        "gpu_launches": 4 * args.steps,         # per step: the profile kernel once per draw class (3) + bounds/masks (1)
        "roofline": {"kernel": "fb_profile_spec_kernel<Spec<13,16,8,1>, philox, 4 features per word> (kernel_ms also holds the two "
                               "launches for the other draw classes, which find no partition of theirs in this grid)",
                     "bound": "fp64", "achieved": achieved,
                     "peak": peak_tf, "unit": "TFLOP/s", "frac": achieved / peak_tf, "traffic": traffic,
                     "peak_source": "measured in this run: fb_fp64_peak (DFMA issue-rate microbenchmark, FMA = 2 flops); "
                                    "MEASURED_PEAKS.json has no FP64 figure; nominal 148 SM x 64 lanes x 1.965 GHz x 2 = 37.2",
                     "algorithmic_flops_per_launch": flops_profile, "kernel_ms": float(prof_ms.item()),
                     "share_of_step": float(prof_ms.item()) / (total_ms / args.steps),
                     "hbm": {"achieved_gbs": alg_bytes / (float(prof_ms.item()) * 1e-3) / 1e9, "peak_gbs": hbm_peak,
                             "frac": alg_bytes / (float(prof_ms.item()) * 1e-3) / 1e9 / hbm_peak,
                             "peak_source": "MEASURED_PEAKS.json hbm_gbs" if "hbm_gbs" in peaks else "fallback"}},
        "cpu_baseline": {"value": cpu_v, "unit": UNIT, "cores": cores, "kind": "port",
                         "sample": f"first {cpu_sample} partitions of the same workload, full pre-solver, {cpu_dt:.1f} s"},
        "extra": {"ibp_masks_kernel_ms": ibp_ms, "ibp_only_evals_per_s": n / (ibp_only_ms * 1e-3),
                  "box_generation_boxes_per_s": n / (boxgen_ms * 1e-3),
                  "partition_screen_pairs_per_s": ns * S_SAMPLES / (screen_ms * 1e-3),
                  "partition_screen_note": "pairs_per_s = boxes x samples / time: boxes whose float32 enclosure is one-signed "
                                           "are decided without evaluating their samples (same answers)",
                  "partition_screen_partitions_per_s": ns / (screen_ms * 1e-3),
                  "partition_screen_skipping_certified_partitions_per_s": ns / (screen_skip_ms * 1e-3),
                  "certified_fraction": certified_frac,
                  "fp64_peak_tflops_measured": peak_tf,
                  "gathered_records_equal_recomputation": gather_checked,
                  "src_grids_configs_1_2": src_grids,
                  "relaxed_targeted_ac_models_config_5": ac_sweep,
                  "bm_sweep_20000_partitions_S1000": bm_sweep},
    }
    if saved_stdout is not None:
        sys.stdout.flush()
        os.dup2(saved_stdout, 1)
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--partitions", type=int, default=10_000_000, help="partitions per GPU (weak) or in total (strong)")
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"],
                    help="weak: --partitions per GPU (default, the headline); strong: --partitions in total, split over the GPUs")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 0)
    if args.impl == "reference":
        return run_reference(args)
    if args.warmup < 3:
        args.warmup = 3                                   # timing rule: at least 3 warm-up steps
    return run_gpu(args)


if __name__ == "__main__":
    sys.exit(main())
