"""Time single kernels (CUDA events, engine stream). Usage: python scripts/kernel_times.py [N Nx Na reps] [opt=val:kern ...]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import tomo_tv_b200 as ttv
from tomo_tv_b200.phantom import phantom3d

args = [a for a in sys.argv[1:] if ":" not in a]
n, nx, na, reps = (int(v) for v in (args + ["512", "512", "180", "5"][len(args):])[:4])
runs = [a for a in sys.argv[1:] if ":" in a] or ["fp_variant=1:fp_sirt", "fp_variant=2:fp_sirt", "bp_variant=1:bp_sirt", "bp_variant=2:bp_sirt"]
ang = np.deg2rad(np.linspace(-90, 90, na, endpoint=False)).astype(np.float32)
eng = ttv.tomoengine(nx, n, ang)
eng.initialize_initial_volume(); eng.set_recon_all(phantom3d(nx, n))
V, S, samples = eng.sizes()
for r in runs:
    ov, kern = r.split(":")
    if ov:
        opt, val = ov.split("=")
        eng.set_option(opt, int(val))
    eng.time_kernel(kern, 1)
    ms = eng.time_kernel(kern, reps)
    print(f"{ov:14s} {kern:10s} {ms:8.3f} ms  {samples / ms / 1e9:7.2f} Tsamples/s", flush=True)
