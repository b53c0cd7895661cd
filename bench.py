#!/usr/bin/env python
"""Headline benchmark: FISTA-TV iterations/s at 512^3 x 180 angles (BASELINE.json), 1..8 B200s.

One step = one FISTA-TV iteration of the reference loop (tomofusion/gpu/reconstructor.py:137-148 with
show_convergence=False): SIRT(1) on yk  +  tv_fgp(nTViter=10, lambda=0.1) on recon  +
fista_momentum(beta_k), issued through the C ABI exactly as TomoGPU.fista issues them.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--size 512] [--angles 180]
  python bench.py --impl reference ...      times the CPU restatement of the same iteration (oracle/,
                                            OpenMP on all host cores) on a bounded slab of the workload

N > 1: launched by torchrun, one process per GPU; the volume is sharded into contiguous x-slabs
(strong scaling: the 512^3 job is fixed), TV halos and scalar norms go over NCCL.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

def traffic_512():
    """dram__bytes_read.sum + dram__bytes_write.sum per launch from the committed `ncu --set full` capture of this round's
    default kernels at 512^3 x 180 on one B200 (profiles/traffic_512.json, written by scripts/ncu_traffic.py from the
    .ncu-rep); no capture -> no number (traffic: null)."""
    try:
        return json.load(open(os.path.join(ROOT, "profiles", "traffic_512.json")))
    except Exception:
        return {}


METRIC = "fista_tv_iters_per_sec"
UNIT = "iterations/s"
LAMBDA, NTV = 0.1, 10


def peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))), "measured"
    except Exception:
        return {"hbm_gbs": 6650.0}, "fallback"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm = [float(r[0]) for r in self.rows if len(r) >= 7 and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) >= 7 and r[1].replace(".", "").isdigit()]
        pw = [float(r[2]) for r in self.rows if len(r) >= 7 and r[2].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in self.rows if len(r) >= 7 for i in range(4) if r[3 + i].lower().startswith("active")})
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(pw) if pw else None, "samples": len(sm), "reasons": reasons}


def fista_iteration(eng, state):
    """One FISTA-TV iteration through the reference-facing calls."""
    eng.SIRT(1)
    eng.tv_fgp(NTV, LAMBDA)
    tk = 0.5 * (1.0 + np.sqrt(1.0 + 4.0 * state["t"] ** 2))
    eng.fista_momentum((state["t"] - 1.0) / tk)
    state["t"] = tk


# ------------------------------------------------------------------------------------------------
def cpu_fista_sample(n, na, nslab, steps, warmup):
    """The same iteration on the CPU oracle (test infrastructure, timed here as the reported CPU
    baseline) for a slab of `nslab` slices of the n^2 x na problem.  Returns seconds per iteration."""
    import oracle as orc
    from tomo_tv_b200.phantom import phantom3d, tilt_angles_deg
    ang = np.deg2rad(tilt_angles_deg(na)).astype(np.float32)
    o = orc.Oracle(nslab, n, ang, quant=1)
    o.recon[:] = phantom3d(nslab, n)
    o.forward_projection()
    o.set_tilt_series(o.g)
    o.initialize_fista(); o.restart_recon()
    st = {"t": 1.0}
    for _ in range(warmup):
        fista_iteration(o, st)
    t0 = time.perf_counter()
    for _ in range(steps):
        fista_iteration(o, st)
    return (time.perf_counter() - t0) / max(steps, 1), orc.lib().orc_num_threads()


def ref_gpu_fista_sample(n, na, nslab, steps):
    """The UNMODIFIED reference GPU engine (oracle/_ref/tomoengine*.so: tomo_TV's pybind11 class + its ASTRA fork,
    built by oracle/build_ref.sh; test infrastructure) on a slab of `nslab` slices of the same problem.  The
    reference loops over slices on the host (tomoengine.cpp:189-205), so its cost is linear in the slice count.
    Returns seconds per iteration of the slab, or None when the extension is not available."""
    ref_dir = os.path.join(ROOT, "oracle", "_ref")
    if not os.path.isdir(ref_dir):
        return None
    sys.path.insert(0, ref_dir)
    try:
        import tomoengine as ref_mod
    except Exception:
        return None
    import oracle as orc  # noqa: F401  (only to keep the leg inside the oracle/ boundary explicit)
    from tomo_tv_b200.phantom import phantom3d, tilt_angles_deg
    import tomo_tv_b200 as ttv
    ang = np.deg2rad(tilt_angles_deg(na)).astype(np.float32)
    x0 = (n - nslab) // 2
    mk = ttv.tomoengine(nslab, n, ang)
    mk.initialize_initial_volume(); mk.set_original_volume_all(phantom3d(n, n, x0=x0, count=nslab)); mk.create_projections()
    b = mk.get_projections(); del mk
    r = ref_mod.tomoengine(nslab, n, ang)
    r.set_tilt_series(b)
    r.initialize_fista(); r.initialize_FP(); r.restart_recon()
    st = {"t": 1.0}
    fista_iteration(r, st)
    t0 = time.perf_counter()
    for _ in range(steps):
        fista_iteration(r, st)
    return (time.perf_counter() - t0) / max(steps, 1)


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    # torchrun exports OMP_NUM_THREADS=1 to its workers: this arm is the CPU baseline "with all the host threads it can
    # use", so the thread count is set explicitly before the OpenMP runtime of oracle/liboracle.so starts
    os.environ["OMP_NUM_THREADS"] = str(cores)
    nslab = min(args.size, max(64, cores))               # >= 64 slices per step: a bounded SAMPLE of the job, scaled below
    sec, threads = cpu_fista_sample(args.size, args.angles, nslab, args.steps, args.warmup)
    value = (1.0 / sec) * (nslab / args.size)          # iterations/s of the FULL job at this per-slice rate
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": sec * 1e3 * (args.size / nslab), "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"FISTA-TV {args.size}^3 x {args.angles} angles, nTViter={NTV}, lambda={LAMBDA}",
                   "sampled": f"{nslab} of {args.size} slices timed per step; value = per-slice rate x {args.size}/{nslab} (slices are independent on the CPU path)",
                   "note": "reference GPU path needs CUDA; its CPU package (tomofusion/cpu) has no TV-FGP/FISTA, so the CPU arm is the "
                           "oracle restatement of the same iteration (oracle/oracle.c, OpenMP over slices)"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "port",
                         "sample": f"{nslab} of {args.size} slices of the {args.size}^2 x {args.angles} problem per step, scaled by {args.size}/{nslab}"},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------
def run_ours(args):
    # stdout carries exactly one JSON line: NCCL's "NCCL version ..." banner (NCCL_DEBUG=VERSION on the GPU
    # boxes) is printed to stdout by rank 0, so keep only warnings unless the user asked for INFO/TRACE
    if os.environ.get("NCCL_DEBUG", "").upper() in ("VERSION", ""):
        os.environ["NCCL_DEBUG"] = "WARN"
    import torch
    import tomo_tv_b200 as ttv
    from tomo_tv_b200.phantom import phantom3d, tilt_angles_deg, slab_bounds

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (tomo_tv_b200 has no CPU fallback)")
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def allmax(v):
        if dist is None:
            return v
        t = torch.tensor([v], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def allsum(v):
        if dist is None:
            return v
        t = torch.tensor([v], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    n, na = args.size, args.angles
    x0, nloc = slab_bounds(n, world, rank)
    ang = np.deg2rad(tilt_angles_deg(na)).astype(np.float32)
    eng = ttv.tomoengine(nloc, n, ang, device=local, x_offset=x0, nslice_global=n)
    if world > 1:
        eng.setup_sharding(dist, rank, world, p2p=os.environ.get("TTV_P2P", "1") != "0")

    # synthetic data: phantom slab -> tilt series by the engine's own forward projector
    if args.phantom_period > 0:
        # c5-sized runs (2048^3): evaluating every slice of the phantom on the host takes ~1 s per slice, so a
        # block of `period` slices from the middle of this rank's slab is tiled along x instead
        per = min(args.phantom_period, nloc)
        blk = phantom3d(n, n, x0=x0 + (nloc - per) // 2, count=per)
        vol = np.concatenate([blk] * ((nloc + per - 1) // per), axis=0)[:nloc]
        del blk
    else:
        vol = phantom3d(n, n, x0=x0, count=nloc)
    eng.initialize_initial_volume()
    eng.set_original_volume_all(vol)
    eng.create_projections()
    # pinned host buffers for the end-to-end leg
    b_host = torch.empty((nloc, na * n), dtype=torch.float32, pin_memory=True)
    b_np = b_host.numpy()
    b_np[:] = eng.get_projections()
    out_host = torch.empty((nloc, n, n), dtype=torch.float32, pin_memory=True)
    out_np = out_host.numpy()
    del vol

    V, S, samples = eng.sizes()           # bytes of the local volume / sinogram slab
    eng.initialize_fista(); eng.initialize_FP(); eng.restart_recon()
    st = {"t": 1.0}
    barrier()                                   # ranks finish their host-side set-up seconds apart at large sizes
    for _ in range(max(args.warmup, 3)):
        fista_iteration(eng, st)

    # ---------------- device-resident timed region ----------------
    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()                          # before the barrier: spawning nvidia-smi on rank 0 must not skew the ranks' start
    barrier()
    eng.profile(True)
    l0 = eng.launch_count()
    eng.timer_start()
    for _ in range(args.steps):
        fista_iteration(eng, st)
    ms = eng.timer_stop()
    barrier()
    launches = eng.launch_count() - l0
    prof = eng.profile_report()
    eng.profile(False)
    clk = clocks.stop() if rank == 0 else None
    ms = allmax(ms)
    launches = int(allsum(launches))
    ms_per_step = ms / args.steps
    value = 1e3 / ms_per_step

    # ---------------- end-to-end leg: host buffers in, host volume out ----------------
    barrier()
    t0 = time.perf_counter()
    eng.set_tilt_series(b_np)                                   # H2D from pinned memory (+ layout pack on device)
    eng.restart_recon()
    st2 = {"t": 1.0}
    for _ in range(args.steps):
        fista_iteration(eng, st2)                               # tv_fgp returns the TV scalar to the host every step
    eng.get_recon_all_into(out_np)                              # D2H into pinned memory
    torch.cuda.synchronize()
    e2e_s = allmax(time.perf_counter() - t0)
    barrier()
    e2e_value = args.steps / e2e_s
    # the reconstruction the end-to-end leg produced, as one hash that does not depend on how many GPUs it was sharded
    # over: sha256 over the per-slice sha256 digests in global slice order (every kernel on the path computes a voxel
    # with the same operations whatever the slab thickness, so the hashes of the 1-, 2-, 4- and 8-GPU runs are equal)
    import hashlib
    digests = [hashlib.sha256(out_np[i].tobytes()).digest() for i in range(nloc)]
    if dist is not None:
        allg = [None] * world
        dist.all_gather_object(allg, (x0, digests))
        digests = [d for _, ds in sorted(allg, key=lambda t: t[0]) for d in ds]
    recon_sha = hashlib.sha256(b"".join(digests)).hexdigest()
    recon_tv = eng.tv()                                         # TV of the final volume (all-reduced scalar)
    # Under momentum the reference discards the TV step (tomoengine.cpp:382: recon is overwritten by the momentum copy), so the
    # hash above covers the projector / momentum path only.  The TV-FGP path with its ghost hand-off gets its own hash: one more
    # tv_fgp on the final volume (outside every timed region), hashed the same way.
    eng.tv_fgp(NTV, LAMBDA)
    eng.get_recon_all_into(out_np)
    digests = [hashlib.sha256(out_np[i].tobytes()).digest() for i in range(nloc)]
    if dist is not None:
        allg = [None] * world
        dist.all_gather_object(allg, (x0, digests))
        digests = [d for _, ds in sorted(allg, key=lambda t: t[0]) for d in ds]
    fgp_sha = hashlib.sha256(b"".join(digests)).hexdigest()

    # ---------------- roofline of the dominant kernel ----------------
    pk, pk_kind = peaks()
    nvox = nloc * n * n
    alg_bytes = {                                               # algorithmic bytes per launch
        "fp": V + 2 * S,                                        # read x, read b, write residual (fused SIRT residual)
        "bp": S + 2 * V,                                        # read residual, read x, write x (fused SIRT update)
        "fgp_first": 4 * V, "fgp_sweep": 7 * V, "fgp_final": 5 * V, "tv": V, "momentum": 3 * V,
        # two sweeps per launch (ttv_fgp2.cuh): SURVEY 8d's per-sweep figures x the two sweeps one launch processes
        # (FIRST 4V + SWEEP 7V, SWEEP 7V + SWEEP 7V, SWEEP 7V + FINAL 5V); the launch MOVES about half of that (see `traffic`)
        "fgp2_first": 11 * V, "fgp2_sweep": 14 * V, "fgp2_final": 12 * V,
    }
    groups = {}
    for g, (tot, cnt) in prof.items():
        per = tot / cnt
        ent = {"ms_per_launch": per, "launches_per_step": cnt / args.steps, "share_of_step": tot / ms if ms > 0 else None}
        if g in alg_bytes:
            ent["algorithmic_bytes"] = alg_bytes[g]
            ent["achieved_gbs"] = alg_bytes[g] / (per * 1e-3) / 1e9
            ent["frac_of_hbm_peak"] = ent["achieved_gbs"] / pk["hbm_gbs"]
        if g in ("fp", "bp"):
            ent["samples_per_s"] = samples / (per * 1e-3)
        groups[g] = ent
    # `roofline`: the dominant HBM-bound kernel (fgp_tile_kernel, sweep mode: 8 launches per step).  The projector
    # pair takes more of the step but is bound by shared-memory bandwidth, not HBM: it gets its own
    # object against the LDS peak (SMs x 128 B/clk x measured SM clock); 8 B of shared memory per bilinear sample.
    hbm_groups = [g for g in prof if g in alg_bytes and g not in ("fp", "bp")]
    dom = max(hbm_groups, key=lambda g: prof[g][0]) if hbm_groups else None
    roof = None
    if dom:
        roof = {"kernel": {"fgp_sweep": "fgp_tile_kernel<SWEEP>", "fgp_first": "fgp_tile_kernel<FIRST>", "fgp_final": "fgp_tile_kernel<FINAL>",
                           "fgp2_sweep": "fgp2_kernel<SWEEP,SWEEP>", "fgp2_first": "fgp2_kernel<FIRST,SWEEP>", "fgp2_final": "fgp2_kernel<SWEEP,FINAL>"}.get(dom, dom),
                "bound": "hbm", "achieved": groups[dom]["achieved_gbs"], "peak": pk["hbm_gbs"], "unit": "GB/s",
                "frac": groups[dom]["frac_of_hbm_peak"], "traffic": traffic_512().get(dom) if (n, na, world) == (512, 180, 1) else None, "peak_kind": pk_kind,
                "algorithmic_bytes_per_launch": alg_bytes[dom], "share_of_step": groups[dom]["share_of_step"],
                "note": ("two TV-FGP sweeps per launch: `achieved` counts the algorithmic bytes of both sweeps (2 x SURVEY 8d), so frac > 1 means the "
                         "temporal fusion beats the per-sweep HBM roofline; `traffic` / duration is what HBM actually delivers" if dom.startswith("fgp2") else None),
                "traffic_note": "dram__bytes_read.sum + dram__bytes_write.sum per launch from profiles/traffic_512.json (ncu --set full, 512^3 x 180, 1 GPU); null without a capture"}
    if roof and roof.get("traffic"):
        roof["traffic_gbs"] = roof["traffic"] / (groups[dom]["ms_per_launch"] * 1e-3) / 1e9
        roof["traffic_frac_of_peak"] = roof["traffic_gbs"] / pk["hbm_gbs"]
    sm_mhz = (clk or {}).get("sm_mhz") or pk.get("sm_max_mhz") or 1965.0
    try:
        n_sm = torch.cuda.get_device_properties(local).multi_processor_count
    except Exception:
        n_sm = 148
    lds_peak = n_sm * 128 * sm_mhz * 1e6 / 1e9
    # shared-memory bytes per bilinear sample of the default kernels: register-file deliveries (LDS wavefronts
    # x 128 B) + TMA fill, measured with ncu (profiles/r2d_*): fpt_kernel 1345 + 316 wavefronts per 32768 samples,
    # bpt_kernel 897 + 168 per 32768; the classic kernels (thin slabs) read 8 B per sample + fill
    lds_model = {"fp": (1345 + 316) * 128 / 32768.0 if nloc >= 128 else 8.0 + 1.25, "bp": (897 + 168) * 128 / 32768.0 if nloc >= 64 else 8.0 + 0.4}
    proj_roof = {}
    for g in ("fp", "bp"):
        if g in groups:
            ach = samples * lds_model[g] / (groups[g]["ms_per_launch"] * 1e-3) / 1e9
            proj_roof[g] = {"bound": "shared-memory", "achieved": ach, "peak": lds_peak, "unit": "GB/s", "frac": ach / lds_peak,
                            "samples_per_s": groups[g]["samples_per_s"], "share_of_step": groups[g]["share_of_step"],
                            "smem_bytes_per_sample": lds_model[g], "peak_def": f"{n_sm} SMs x 128 B/clk x {sm_mhz:.0f} MHz"}
    iter_bytes = world * (74 * V + 3 * S)
    iter_roof = {"algorithmic_bytes_per_iteration": iter_bytes, "achieved_gbs": iter_bytes / (ms_per_step * 1e-3) / 1e9,
                 "frac_of_hbm_peak_x_gpus": iter_bytes / (ms_per_step * 1e-3) / 1e9 / (pk["hbm_gbs"] * world)}

    line = None
    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"FISTA-TV {n}^3 x {na} angles, nTViter={NTV}, lambda={LAMBDA}", "parallelism": f"x-slab x{world}",
                       "l2": "working set (10 volumes of %d MiB per GPU) exceeds the 126 MB L2" % (V >> 20),
                       "e2e_def": "set_tilt_series(host) + K iterations (TV scalar read back each) + get_recon_all(host), K = steps"},
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(S / args.steps),
                    "d2h_bytes_per_step": int(V / args.steps) + 4},
            "gpu_launches": launches,
            "recon_sha256": recon_sha, "recon_tv": recon_tv, "tv_fgp_sha256": fgp_sha,
            "clocks": clk,
            "roofline": roof,
            "projector_roofline": proj_roof,
            "kernels": groups,
            "iteration_roofline": iter_roof,
        }
    if rank == 0 and world == 1 and not args.no_cpu:
        cores = os.cpu_count() or 1
        nslab = min(n, max(64, cores))
        sec, threads = cpu_fista_sample(n, na, nslab, 2, 1)
        line["cpu_baseline"] = {"value": (1.0 / sec) * (nslab / n), "unit": UNIT, "cores": threads, "kind": "port",
                                "sample": f"{nslab} of {n} slices per step (2 timed iterations), scaled by {n}/{nslab}"}
    if rank == 0 and world == 1 and not args.no_cpu:
        nsl = min(32, n)
        try:
            sec = ref_gpu_fista_sample(n, na, nsl, 2)
        except Exception as ex:  # the reference aborts the process on CUDA errors; anything catchable is reported
            sec = None
            line["reference_gpu"] = {"unavailable": str(ex)[:200]}
        if sec is not None:
            line["reference_gpu"] = {"value": (1.0 / sec) * (nsl / n), "unit": UNIT, "kind": "reference (oracle/_ref tomoengine + ASTRA fork, same B200)",
                                     "sample": f"{nsl} central slices of {n} per step (2 timed iterations), scaled by {n}/{nsl}: the reference loops slices on the host"}
    if rank == 0:
        print(json.dumps(line), flush=True)
    del eng
    if dist is not None:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--size", type=int, default=512)
    ap.add_argument("--angles", type=int, default=180)
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--max-seconds", type=float, default=900.0, help="watchdog: hard-exit the process after this long")
    ap.add_argument("--phantom-period", type=int, default=0, help="tile a block of this many phantom slices along x (large sizes)")
    args = ap.parse_args()

    def _watchdog():
        # a rank stuck in a collective or a device wait after a peer died must not hold the box until the
        # launcher's own limit: leave hard (no Python clean-up, which could block on the same device)
        time.sleep(args.max_seconds)
        sys.stderr.write(f"bench.py: watchdog fired after {args.max_seconds:.0f} s, exiting\n"); sys.stderr.flush()
        os._exit(3)
    threading.Thread(target=_watchdog, daemon=True).start()
    try:
        if args.impl == "reference":
            run_reference(args)
        else:
            run_ours(args)
    except BaseException:
        # under torchrun a rank that fails must go away at once: interpreter shutdown can block on a device
        # that still runs a peer wait / NCCL kernel, and the launcher only tears the other ranks down (they are
        # waiting for this one) when it sees this process exit
        import traceback
        traceback.print_exc()
        sys.stderr.flush(); sys.stdout.flush()
        if int(os.environ.get("WORLD_SIZE", "1")) > 1:
            os._exit(1)
        raise


if __name__ == "__main__":
    main()
