"""Sharded-vs-whole parity on N GPUs (run under torchrun, one rank per GPU, NCCL):
every rank owns an x-slab; the result must match a whole-volume engine run on rank 0's GPU.
Covers the collectives of SURVEY 8e: FGP halo chain, TV-GD halo ring + sum-of-squares all-reduce,
TV value ring, data-distance / KL-cost all-reduce.
Usage: torchrun --nproc-per-node 2 scripts/mgpu_check.py [Nx N Na]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.distributed as dist
import tomo_tv_b200 as ttv
from tomo_tv_b200.phantom import phantom3d, slab_bounds

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
nx, n, na = (int(v) for v in sys.argv[1:4]) if len(sys.argv) > 3 else (70, 96, 20)
ang = np.deg2rad(np.linspace(-80, 80, na)).astype(np.float32)
rng = np.random.default_rng(5)
vol = phantom3d(nx, n) + 0.02 * rng.standard_normal((nx, n, n)).astype(np.float32)


def run(eng, v):
    res = {}
    eng.initialize_initial_volume(); eng.set_original_volume_all(v); eng.create_projections()
    eng.initialize_recon_copy()
    eng.initialize_fista(); eng.initialize_FP(); eng.restart_recon()
    t0 = 1.0
    for _ in range(3):
        eng.SIRT(1)
        res.setdefault("tv_in", []).append(eng.tv_fgp(4, 0.1))
        tk = 0.5 * (1 + np.sqrt(1 + 4 * t0 * t0)); eng.fista_momentum((t0 - 1) / tk); t0 = tk
    res["dd"] = eng.data_distance(); res["tv"] = eng.tv(); res["rmse"] = eng.rmse()
    res["fista"] = eng.get_recon_all()
    eng.remove_momentum()
    eng.copy_recon()
    res["tvgd_in"] = eng.tv_gd(3, 0.5)
    res["dg"] = eng.matrix_2norm()
    res["tvgd"] = eng.get_recon_all()
    eng.tv_fgp(1, 0.1); res["fgp1"] = eng.get_recon_all()
    eng.tv_fgp(0, 0.1); res["tvmap"] = eng.get_recon_all()
    eng.initialize_poisson_ML()
    res["kl"] = [eng.poisson_ML(0.5) for _ in range(2)]
    res["klrec"] = eng.get_recon_all()
    return res


x0, cnt = slab_bounds(nx, world, rank)
eng = ttv.tomoengine(cnt, n, ang, device=local, x_offset=x0, nslice_global=nx)
eng.setup_sharding(dist, rank, world, p2p=os.environ.get("TTV_P2P", "1") != "0")
mine = run(eng, vol[x0:x0 + cnt])
gathered = [None] * world
dist.all_gather_object(gathered, {k: v for k, v in mine.items()})
ok = True
if rank == 0:
    whole = run(ttv.tomoengine(nx, n, ang, device=local), vol)
    for k, w in whole.items():
        if isinstance(w, np.ndarray):
            g = np.concatenate([d[k] for d in gathered])
            err = float(np.linalg.norm((g - w).ravel()) / max(np.linalg.norm(w.ravel()), 1e-30))
        else:
            g = np.asarray(gathered[0][k], np.float64); w = np.asarray(w, np.float64)
            err = float(np.max(np.abs(g - w) / np.maximum(np.abs(w), 1e-30)))
            for d in gathered[1:]:
                assert np.allclose(np.asarray(d[k]), g), f"{k}: ranks disagree on a global scalar"
        flag = err <= 2e-6
        ok &= flag
        print(f"mgpu[{world}] {k:8s} rel err {err:.3e} {'ok' if flag else 'FAIL'}", flush=True)
    print("MGPU_CHECK", "PASS" if ok else "FAIL", flush=True)
dist.barrier()
del eng

# ---- the reference-facing layer under torchrun: TomoGPU on the sharded multigpuengine (config c3: SIRT + ASD-POCS)
tmp = ttv.tomoengine(nx, n, ang, device=local)
tmp.initialize_initial_volume(); tmp.set_original_volume_all(vol); tmp.create_projections()
tilt = tmp.get_projections().reshape(nx, na, n).transpose(0, 2, 1)       # [Nslice, Nray, Nangles]
del tmp
deg = np.rad2deg(ang.astype(np.float64))


def api_run(t):
    out = {}
    t.sirt(3, show_convergence=False); out["api_sirt"] = t.get_recon().copy(); out["api_sirt_cost"] = t.cost.copy()
    t.asd_pocs(3, show_convergence=False); out["api_asd"] = t.get_recon().copy(); out["api_asd_tv"] = t.tv_vec.copy()
    return out


sh = ttv.TomoGPU(deg, tilt, verbose=False)
assert isinstance(sh.tomo, ttv.multigpuengine) and sh.tomo.is_multi_gpu_enabled() and len(sh.tomo.get_gpu_ids()) == world
got = api_run(sh)
if rank == 0:
    want = api_run(ttv.TomoGPU(deg, tilt, gpu_id=local, verbose=False))
    for k, w in want.items():
        err = float(np.linalg.norm((got[k] - w).ravel()) / max(np.linalg.norm(w.ravel()), 1e-30))
        flag = err <= 2e-6
        ok &= flag
        print(f"mgpu[{world}] {k:12s} rel err {err:.3e} {'ok' if flag else 'FAIL'}", flush=True)
    print("MGPU_API_CHECK", "PASS" if ok else "FAIL", flush=True)
dist.barrier()
del sh
dist.destroy_process_group()
sys.exit(0 if ok else 1)
