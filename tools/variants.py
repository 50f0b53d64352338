#!/usr/bin/env python3
"""Build tuning variants of the specialised AC kernels (only fb_spec_a.cu is recompiled) and, on a GPU box,
time fb_profile for each.  Development aid, not part of the product.

    python tools/variants.py build  tag:-DFB_PROFILE_SPT=3 tag2:"-DFB_SPEC_JB=2 -DFB_PROFILE_MINB=2" ...
    python tools/variants.py run    tag tag2 ...          (on the GPU box)
"""
import concurrent.futures as cf, glob, os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
VDIR = os.path.join(ROOT, "fairify_b200", "_variants")


def build(specs):
    from fairify_b200 import build as B
    B.build()
    os.makedirs(VDIR, exist_ok=True)
    objs = [o for o in glob.glob(os.path.join(B.OBJ, "*.o")) if not o.endswith("fb_spec_a.o")]

    def one(spec):
        tag, flags = spec.split(":", 1)
        obj = os.path.join(VDIR, f"fb_spec_a_{tag}.o"); so = os.path.join(VDIR, f"lib_{tag}.so")
        log = os.path.join(VDIR, f"{tag}.ptxas.log")
        cmd = [B._nvcc()] + B.NVCC_FLAGS + flags.split() + ["-Xptxas", "-v", "-c", os.path.join(B.CSRC, "fb_spec_a.cu"), "-o", obj]
        r = subprocess.run(cmd, capture_output=True, text=True)
        open(log, "w").write(r.stdout + r.stderr)
        if r.returncode: raise RuntimeError(r.stderr[-2000:])
        subprocess.check_call([B._nvcc(), "-shared", "-o", so, obj] + objs + ["-gencode", "arch=compute_100a,code=sm_100a", "-ccbin", "/usr/bin/g++"])
        import re
        txt = open(log).read()
        m = re.search(r"fb_profile_spec_kernelI4SpecIJLi13ELi16ELi8ELi1EEELb1ELi4E.*?\n.*?\n\s*(\d+) bytes stack frame, (\d+) bytes spill stores.*?\n.*?Used (\d+) registers", txt, re.S)
        return tag, (m.groups() if m else None)
    with cf.ThreadPoolExecutor(max_workers=8) as ex:
        for tag, info in ex.map(one, specs):
            print(tag, "stack/spill/regs:", info, flush=True)


def run(tags, n=2_000_000):
    for tag in tags:
        so = os.path.join(VDIR, f"lib_{tag}.so") if tag != "default" else ""
        code = f"""
import sys, numpy as np, torch
sys.path.insert(0, {ROOT!r})
from fairify_b200 import engine
from fairify_b200.domains import DOMAINS
z = np.load({os.path.join(ROOT, 'tests', 'golden', 'weights.npz')!r})
w = [z[f'AC-1/w{{l}}'] for l in range(3)]; b = [z[f'AC-1/b{{l}}'] for l in range(3)]
m = engine.Model(w, b); D = DOMAINS['AC']; g = engine.Grid(D.columns, D.ranges, 4)
boxes = g.boxes(0, {n})
ref = None
for it in range(3):
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record(); c = m.profile(boxes, S=1000, seed=7); e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
pk = max(engine.fp64_peak(4096)[0], engine.fp64_peak(4096)[0])
tf = {n} * 1000 * (2 * 344 + 25) / (ms * 1e-3) / 1e12
print({tag!r}, 'ms', round(ms, 2), 'TF', round(tf, 2), 'frac', round(tf / pk, 3), 'chk', int(c.sum().item()))
"""
        env = dict(os.environ)
        if so: env["FAIRIFY_B200_LIB"] = so
        r = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True)
        print(r.stdout.strip() or r.stderr[-1500:], flush=True)


if __name__ == "__main__":
    (build if sys.argv[1] == "build" else run)(sys.argv[2:])
