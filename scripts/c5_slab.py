"""BASELINE.json configs[4] (2048^2 x 2048 slices x 180 angles across 8 B200), one rank's share on ONE GPU:
a 256-slice slab of the 2048^3 volume (V = 4 GiB, 2^30 voxels: byte offsets beyond 2^32, FISTA-TV holds
10 V = 40 GiB) run through the same FISTA-TV iteration as bench.py.

Size-independent checks (no oracle at this size):
  * the phantom slab repeats with period 64 slices, so every slice-independent result (SIRT step on yk)
    must repeat bit-exactly across the slab, including the slices whose addresses lie beyond 2^31 elements;
  * the TV-FGP output (10 sweeps) must repeat on all slices further than 24 from the slab's ends.
Prints one JSON line (iteration time of the slab, which is the 8-GPU iteration time of c5 less the halo wait).
Usage: python scripts/c5_slab.py [Nx N Na steps]"""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import tomo_tv_b200 as ttv
from tomo_tv_b200.phantom import phantom3d

nx, n, na, steps = (int(v) for v in sys.argv[1:5]) if len(sys.argv) > 4 else (256, 2048, 180, 3)
per = 64
ang = np.deg2rad(np.linspace(-90, 90, na, endpoint=False)).astype(np.float32)
eng = ttv.tomoengine(nx, n, ang, device=0)
base = phantom3d(per, n)                                   # [64, n, n]
vol = np.concatenate([base] * (nx // per), axis=0)
eng.initialize_initial_volume()
eng.set_original_volume_all(vol)
eng.create_projections()
del vol
eng.initialize_fista(); eng.initialize_FP(); eng.restart_recon()
res = {"workload": f"c5 slab {nx} x {n}^2 x {na} angles on 1 GPU", "checks": {}}


def periodic(a, margin=0):
    """max |a[s] - a[s + per]| over the slab, skipping `margin` slices at both slab ends."""
    worst = 0.0
    for s in range(margin, nx - per - margin):
        worst = max(worst, float(np.max(np.abs(a[s] - a[s + per]))))
    return worst


t = 1.0
eng.SIRT(1)
eng.tv_fgp(10, 0.1)
tk = 0.5 * (1 + np.sqrt(1 + 4 * t * t)); eng.fista_momentum((t - 1) / tk); t = tk
rec = eng.get_recon_all()                                   # recon <- yk = SIRT step result (slice independent)
res["checks"]["sirt_periodic_maxabs"] = periodic(rec)
res["checks"]["recon_max"] = float(rec.max())
eng.tv_fgp(10, 0.1)                                         # on a non-trivial recon; the slab edges (non-periodic
rec = eng.get_recon_all()                                   # boundary rules) reach <= 2 slices per sweep inwards
res["checks"]["fgp_periodic_interior_maxabs"] = periodic(rec, 24)
res["checks"]["fgp_changed_recon"] = bool(float(rec.max()) != res["checks"]["recon_max"])
del rec
eng.sync()
eng.timer_start()
for _ in range(steps):
    eng.SIRT(1); eng.tv_fgp(10, 0.1)
    tk = 0.5 * (1 + np.sqrt(1 + 4 * t * t)); eng.fista_momentum((t - 1) / tk); t = tk
ms = eng.timer_stop() / steps
V, S, samples = eng.sizes()
res["ms_per_iteration"] = ms
res["iterations_per_s"] = 1e3 / ms
res["algorithmic_GBps"] = (74 * V + 3 * S) / ms / 1e6
res["dd"] = eng.data_distance()
ok = res["checks"]["sirt_periodic_maxabs"] == 0.0 and res["checks"]["fgp_periodic_interior_maxabs"] == 0.0 and res["checks"]["recon_max"] > 0
res["pass"] = bool(ok)
print(json.dumps(res))
sys.exit(0 if ok else 1)
