"""Time the experimental projector variants against the defaults (each in its own process, so a faulting kernel
cannot take the others down).  Usage: python scripts/variants_probe.py [N Nx Na reps]   (default 512 512 180 5)"""
import os, subprocess, sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CHILD = r"""
import sys, os
sys.path.insert(0, %r)
import numpy as np
import tomo_tv_b200 as ttv
from tomo_tv_b200.phantom import phantom3d
opt, val, kern, n, nx, na, reps = sys.argv[1], int(sys.argv[2]), sys.argv[3], *(int(v) for v in sys.argv[4:8])
ang = np.deg2rad(np.linspace(-90, 90, na, endpoint=False)).astype(np.float32)
eng = ttv.tomoengine(nx, n, ang)
eng.set_option(opt, val)
eng.initialize_initial_volume(); eng.set_recon_all(phantom3d(nx, n))
eng.time_kernel(kern, 1)
ms = eng.time_kernel(kern, reps)
V, S, samples = eng.sizes()
print(f"{opt}={val} {kern:8s} {ms:8.3f} ms  {samples / ms / 1e9:7.2f} Tsamples/s", flush=True)
""" % ROOT

_a = sys.argv[1:5]
n, nx, na, reps = _a + ["512", "512", "180", "5"][len(_a):]
for opt, val, kern in (("fp_variant", 0, "fp_sirt"), ("fp_variant", 1, "fp_sirt"), ("fp_variant", 8, "fp_sirt"),
                       ("bp_variant", 4, "bp_sirt"), ("bp_variant", 5, "bp_sirt"), ("bp_variant", 8, "bp_sirt"), ("bp_variant", 9, "bp_sirt")):
    try:
        r = subprocess.run([sys.executable, "-c", CHILD, opt, str(val), kern, n, nx, na, reps], capture_output=True, text=True, timeout=300)
        print(r.stdout.strip() or f"{opt}={val}: exit {r.returncode}: {r.stderr.strip()[-300:]}", flush=True)
    except subprocess.TimeoutExpired:
        print(f"{opt}={val}: timed out", flush=True)
