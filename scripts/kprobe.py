"""Time named kernel groups alone (CUDA events on the engine stream, ttv_time_kernel).
Usage: python scripts/kprobe.py N Nx Na reps name [name ...]"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import tomo_tv_b200 as ttv

n, nx, na, reps = (int(v) for v in sys.argv[1:5])
names = sys.argv[5:]
rng = np.random.default_rng(0)
ang = np.deg2rad(np.linspace(-90, 90, na, endpoint=False)).astype(np.float32)
eng = ttv.tomoengine(nx, n, ang)
from tomo_tv_b200.phantom import phantom3d
eng.set_recon_all(phantom3d(nx, n) + 0.05 * rng.standard_normal((nx, n, n)).astype(np.float32))
V, S, samples = eng.sizes()
for k in names:
    ms = eng.time_kernel(k, reps)
    print(f"N={n} Nx={nx} Na={na} {k:14s} {ms:9.4f} ms   V/ms={V / ms / 1e6:9.1f} GB/s-per-V", flush=True)
