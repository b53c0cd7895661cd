"""Projector timing probe: time the FP / BP kernels alone (CUDA events on the engine stream) for
vertical-only, horizontal-only and mixed angle sets.  Usage: python scripts/fp_probe.py [N] [Nx] [reps]"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import tomo_tv_b200 as ttv

n = int(sys.argv[1]) if len(sys.argv) > 1 else 512
nx = int(sys.argv[2]) if len(sys.argv) > 2 else 128
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 3
sets = {
    "vertical(-44..44)": np.linspace(-44, 44, 90),
    "horizontal(46..134)": np.linspace(46, 134, 90),
    "mixed(-90..90,180)": np.linspace(-90, 90, 180, endpoint=False),
}
rng = np.random.default_rng(0)
for name, deg in sets.items():
    eng = ttv.tomoengine(nx, n, np.deg2rad(deg).astype(np.float32))
    eng.set_recon_all(rng.random((nx, n, n), dtype=np.float32))
    V, S, samples = eng.sizes()
    for k in ("fp", "fp_sirt", "bp", "bp_sirt"):
        ms = eng.time_kernel(k, reps)
        print(f"N={n} Nx={nx} {name:22s} {k:8s} {ms:9.3f} ms  {samples / ms / 1e6:9.1f} Gsamples/s", flush=True)
    del eng
