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
Weak scaling: per-GPU work is fixed; with N > 1 each step ends with the NCCL all-gather of the packed
verdict / mask records, inside the timed region.

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
        "warmup": args.warmup, "ms_per_step": 1e3 * t_tot / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": workload_name(args.partitions), "cpu_sample": sample_txt},
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
    from fairify_b200.dist import pack_records
    from fairify_b200.domains import DOMAINS

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the product path has no CPU fallback; use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")     # NCCL's version banner must not share stdout with the JSON line
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    w, b = load_ac1()
    model = engine.Model(w, b)
    D = DOMAINS["AC"]
    grid = engine.Grid(D.columns, D.ranges, GRID_T)
    n = args.partitions
    first = rank * n
    if first + n > len(grid):
        raise SystemExit(f"grid of {len(grid)} boxes cannot give {n} per rank to {world} ranks")
    lo = [D.ranges[c][0] for c in D.columns]; hi = [D.ranges[c][1] for c in D.columns]
    tol = torch.as_tensor(model.tolerance(lo, hi)).cuda()

    # inputs resident in HBM before the timed region
    ev = lambda: torch.cuda.Event(enable_timing=True)
    g0, g1 = ev(), ev()
    g0.record(); boxes = grid.boxes(first, n); g1.record(); torch.cuda.synchronize()
    boxgen_ms = g0.elapsed_time(g1)
    counts32 = torch.empty((n, model.N), dtype=torch.int32, device="cuda")
    outs, po = model._alloc_outputs(n, engine.DEFAULT_OUTPUTS)
    NW = model.NW
    rec_w = 1 + 2 * NW
    packed = torch.empty((n, rec_w), dtype=torch.int32, device="cuda")
    gathered = torch.empty((world * n, rec_w), dtype=torch.int32, device="cuda") if world > 1 else None

    import ctypes as C
    from fairify_b200 import _lib
    L = _lib.lib()
    stream = lambda: C.c_void_p(torch.cuda.current_stream().cuda_stream)
    ptr = lambda t: C.c_void_p(t.data_ptr())

    def step(timers=None):
        if timers:
            timers[0].record()
        _lib.check(L.fb_profile(model._h, n, ptr(boxes), None, S_SAMPLES, C.c_uint64(SEED), first, ptr(counts32), stream()))
        if timers:
            timers[1].record()
        _lib.check(L.fb_ibp(model._h, n, ptr(boxes), ptr(counts32), ptr(tol), C.byref(po), stream()))
        if timers:
            timers[2].record()
        if world > 1:
            pack_records(outs["verdict"], outs["dead"], outs["cand"], out=packed)
            dist.all_gather_into_tensor(gathered, packed)

    for _ in range(args.warmup):
        step()
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
    value = world * n * args.steps / (total_ms * 1e-3)

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
    e2e_value = world * n * args.steps / (float(ems.item()) * 1e-3)
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
    model.partition_screen(boxes[:1000], pa_col, S=S_SAMPLES, seed=SEED, first_index=first)
    s0.record()
    model.partition_screen(boxes[:ns], pa_col, S=S_SAMPLES, seed=SEED, first_index=first)
    s1.record(); torch.cuda.synchronize()
    screen_ms = s0.elapsed_time(s1)

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return 0

    # side numbers for BASELINE.json's BM-1..8 sweep (full pre-solver + sampled counterexample screen per model on a
    # window of the stress/BM grid); not part of the headline
    bm_sweep = {}
    try:
        from fairify_b200.domains import STRESS_THRESHOLD
        z = np.load(os.path.join(ROOT, "tests", "golden", "weights.npz"))
        DB = DOMAINS["BM"]
        gb = engine.Grid(DB.columns, DB.ranges, STRESS_THRESHOLD["BM"])
        nb = min(20000, len(gb))
        bboxes = gb.boxes(0, nb)
        blo = [DB.ranges[c][0] for c in DB.columns]; bhi = [DB.ranges[c][1] for c in DB.columns]
        bpa = DB.columns.index(DB.pa[0])
        for k in range(1, 9):
            name = f"BM-{k}"
            Lk = int(z[f"{name}/L"])
            mk = engine.Model([z[f"{name}/w{l}"] for l in range(Lk)], [z[f"{name}/b{l}"] for l in range(Lk)])
            tk = torch.as_tensor(mk.tolerance(blo, bhi)).cuda()
            for it in range(2):                                   # first pass warms the kernels up
                b0, b1, b2 = ev(), ev(), ev()
                b0.record(); mk.prune(bboxes, S=S_SAMPLES, seed=SEED, tol=tk); b1.record()
                mk.partition_screen(bboxes, bpa, S=S_SAMPLES, seed=SEED); b2.record(); torch.cuda.synchronize()
            bm_sweep[name] = {"weights": mk.P, "kernels": "specialised" if mk.specialised else "runtime-shape",
                              "prune_evals_per_s": nb / (b0.elapsed_time(b1) * 1e-3),
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
        "ms_per_step": total_ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": {"workload": workload_name(n), "model": "AC-1", "partitions_per_gpu": n, "samples": S_SAMPLES,
                   "l2": "inputs (boxes, 104 B x n) and outputs exceed the 126 MB L2; no flush needed",
                   "parallelism": f"contiguous partition range per GPU x{world}" + (", all-gather of verdict+mask records per step" if world > 1 else "")},
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
                  "fp64_peak_tflops_measured": peak_tf,
                  "bm_sweep_20000_partitions_S1000": bm_sweep},
    }
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
    ap.add_argument("--partitions", type=int, default=10_000_000, help="partitions per GPU")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 0)
    if args.impl == "reference":
        return run_reference(args)
    if args.warmup < 3:
        args.warmup = 3                                   # timing rule: at least 3 warm-up steps
    return run_gpu(args)


if __name__ == "__main__":
    sys.exit(main())
