"""One launch of each projector variant (for ncu --set full). Usage: python scripts/ncu_variants.py [N Nx Na] [opt=val:kern ...]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import tomo_tv_b200 as ttv
from tomo_tv_b200.phantom import phantom3d

args = [a for a in sys.argv[1:] if "=" not in a]
n, nx, na = (int(v) for v in (args + ["512", "128", "180"][len(args):])[:3])
runs = [a for a in sys.argv[1:] if "=" in a] or ["fp_variant=1:fp_sirt", "fp_variant=2:fp_sirt", "bp_variant=1:bp_sirt", "bp_variant=2:bp_sirt"]
ang = np.deg2rad(np.linspace(-90, 90, na, endpoint=False)).astype(np.float32)
eng = ttv.tomoengine(nx, n, ang)
eng.initialize_initial_volume(); eng.set_recon_all(phantom3d(nx, n))
for r in runs:
    ov, kern = r.split(":")
    opt, val = ov.split("=")
    eng.set_option(opt, int(val))
    ms = eng.time_kernel(kern, 1)
    print(f"{opt}={val} {kern} {ms:.3f} ms", flush=True)
