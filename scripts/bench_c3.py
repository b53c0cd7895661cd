#!/usr/bin/env python
"""BASELINE.json config c3: 512^3 x 180 angles, SIRT and ASD-POCS, volume sharded along the tilt axis over N GPUs.
Run alone (N = 1) or under torchrun (one rank per GPU).  Prints one JSON line per workload on rank 0:
iterations/s (device-timed, max over ranks), per-kernel-group times, and sha256 of the reconstruction in global slice
order (equal for every N: all kernels on the path compute a voxel with the same operations whatever the slab thickness).
  python scripts/bench_c3.py [--size 512] [--angles 180] [--sirt-steps 10] [--asd-steps 3]
ASD-POCS iteration = the loop body of tomofusion/examples/sim_ASD.py:66-94 (SART pass + data distance + 10 TV-GD steps)."""
import argparse, hashlib, json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--size", type=int, default=512); ap.add_argument("--angles", type=int, default=180)
    ap.add_argument("--sirt-steps", type=int, default=10); ap.add_argument("--asd-steps", type=int, default=3)
    a = ap.parse_args()
    if os.environ.get("NCCL_DEBUG", "").upper() in ("VERSION", ""):
        os.environ["NCCL_DEBUG"] = "WARN"
    import torch
    import tomo_tv_b200 as ttv
    from tomo_tv_b200.phantom import phantom3d, tilt_angles_deg, slab_bounds
    world, rank, local = int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    n, na = a.size, a.angles
    x0, nloc = slab_bounds(n, world, rank)
    eng = ttv.tomoengine(nloc, n, np.deg2rad(tilt_angles_deg(na)).astype(np.float32), device=local, x_offset=x0, nslice_global=n)
    if world > 1:
        eng.setup_sharding(dist, rank, world)
    eng.initialize_initial_volume(); eng.set_original_volume_all(phantom3d(n, n, x0=x0, count=nloc)); eng.create_projections()

    def allmax(v):
        if dist is None: return v
        t = torch.tensor([v], dtype=torch.float64, device="cuda"); dist.all_reduce(t, op=dist.ReduceOp.MAX); return float(t.item())

    def barrier():
        if dist is not None: dist.barrier()
        torch.cuda.synchronize()

    def sha():
        rec = eng.get_recon_all()
        dg = [hashlib.sha256(rec[i].tobytes()).digest() for i in range(nloc)]
        if dist is not None:
            allg = [None] * world
            dist.all_gather_object(allg, (x0, dg))
            dg = [d for _, ds in sorted(allg, key=lambda t: t[0]) for d in ds]
        return hashlib.sha256(b"".join(dg)).hexdigest()

    def timed(name, setup, step, steps, state):
        setup()
        step(state)                                         # warm-up: weights, allocations
        barrier(); eng.profile(True); eng.timer_start()
        for _ in range(steps):
            step(state)
        ms = allmax(eng.timer_stop()) / steps
        prof = eng.profile_report(); eng.profile(False); barrier()
        line = {"workload": name, "size": n, "angles": na, "n_gpus": world, "ms_per_step": ms, "steps_per_s": 1e3 / ms, "steps": steps,
                "kernels": {g: {"ms_per_step": t / steps, "launches_per_step": c / steps} for g, (t, c) in prof.items()}, "recon_sha256": sha()}
        if rank == 0:
            print(json.dumps(line), flush=True)

    timed("sirt", lambda: (eng.initialize_SIRT(), eng.initialize_FP(), eng.restart_recon()), lambda st: eng.SIRT(1), a.sirt_steps, {})

    def asd_setup():
        eng.initialize_SART("sequential"); eng.initialize_FP(); eng.initialize_recon_copy(); eng.restart_recon()

    def asd_step(st):
        eng.copy_recon(); eng.SART(st["beta"], 1); st["beta"] *= 0.9985
        dp = eng.matrix_2norm()
        if st["dPOCS"] is None:
            st["dPOCS"] = dp * 0.2
        dd = eng.data_distance()
        eng.copy_recon(); eng.tv_gd(10, st["dPOCS"])
        dg = eng.matrix_2norm()
        if dg > dp * 0.95 and dd > 0.025:
            st["dPOCS"] *= 0.95
    timed("asd_pocs", asd_setup, asd_step, a.asd_steps, {"beta": 0.25, "dPOCS": None})
    del eng
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
