"""Build libfairify_b200.so in-tree with nvcc for sm_100a (no GPU needed: nvcc cross-compiles).

    python -m fairify_b200.build [--force]

The objects are compiled in parallel (the fully unrolled specialisations take about a minute each) and
linked into fairify_b200/libfairify_b200.so, which is git-ignored but travels with the source tree.
"""
from __future__ import annotations

import concurrent.futures as cf
import glob
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(HERE, "_obj")
SO = os.path.join(HERE, "libfairify_b200.so")

NVCC_FLAGS = [
    "-std=c++17", "-O3", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo",
    # products and sums of the interval recurrence must round separately (reference semantics);
    # every fused multiply-add in the kernels is an explicit intrinsic
    "-fmad=false",
    "-Xcompiler", "-fPIC", "-ccbin", "/usr/bin/g++",
]


def _nvcc():
    for cand in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if cand and (os.path.sep not in cand or os.path.exists(cand)):
            return cand
    raise RuntimeError("nvcc not found")


def _stale(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def _includes(path, seen=None):
    """Transitive closure of the quoted #include lines of `path` (repo headers only)."""
    seen = set() if seen is None else seen
    here = os.path.dirname(path)
    with open(path) as f:
        for line in f:
            line = line.strip()
            if line.startswith('#include "'):
                inc = os.path.normpath(os.path.join(here, line.split('"')[1]))
                if inc not in seen and os.path.exists(inc):
                    seen.add(inc)
                    _includes(inc, seen)
    return seen


def build(force=False, verbose=False):
    os.makedirs(OBJ, exist_ok=True)
    sources = sorted(glob.glob(os.path.join(CSRC, "*.cu")))
    nvcc = _nvcc()
    jobs = []
    wanted = {os.path.basename(s)[:-3] + ".o" for s in sources}
    for stale in glob.glob(os.path.join(OBJ, "*.o")):              # objects of units that no longer exist
        if os.path.basename(stale) not in wanted:
            os.remove(stale)
    for src in sources:
        obj = os.path.join(OBJ, os.path.basename(src)[:-3] + ".o")
        if force or _stale(obj, [src, os.path.abspath(__file__)] + sorted(_includes(src))):
            jobs.append((src, obj))
    # the profile units take longest: start them first
    jobs.sort(key=lambda j: (0 if "fb_specP_" in j[0] else 1, j[0]))

    def compile_one(job):
        src, obj = job
        cmd = [nvcc] + NVCC_FLAGS + os.environ.get("FAIRIFY_B200_NVCC_EXTRA", "").split() + ["-c", src, "-o", obj]
        if verbose:
            print(" ".join(cmd), flush=True)
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"nvcc failed on {src}:\n{r.stdout}\n{r.stderr}")
        return obj

    if jobs:
        with cf.ThreadPoolExecutor(max_workers=min(len(jobs), os.cpu_count() or 4)) as ex:
            list(ex.map(compile_one, jobs))
    objs = [os.path.join(OBJ, os.path.basename(s)[:-3] + ".o") for s in sources]
    if force or jobs or _stale(SO, objs):
        cmd = [nvcc, "-shared", "-o", SO] + objs + ["-gencode", "arch=compute_100a,code=sm_100a", "-ccbin", "/usr/bin/g++"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    return SO


if __name__ == "__main__":
    so = build(force="--force" in sys.argv, verbose=True)
    print(so)
