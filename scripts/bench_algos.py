"""Secondary workloads of the hot path on ONE GPU (bench.py stays the headline FISTA-TV line): iterations/s and the
per-kernel-group table for SIRT, one SART pass, KL/EM, CGLS, one ASD-POCS iteration (SART + TV-GD x10) and FBP, with
the algorithmic bytes of SURVEY.md 8(d) next to the measured time.  Prints one JSON line per workload.
Usage: python scripts/bench_algos.py [--size 512] [--angles 180] [--steps 5] [--only sirt,asd_pocs]"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np


def workloads(V, S):
    """name -> (setup(eng), step(eng, state), algorithmic bytes per step)"""
    def asd_setup(e):
        e.initialize_SART("sequential"); e.initialize_FP(); e.initialize_recon_copy(); e.restart_recon()

    def asd_step(e, st):                                     # tomofusion/examples/sim_ASD.py:66-94
        e.copy_recon(); e.SART(st["beta"], 1); st["beta"] *= 0.9985
        dp = e.matrix_2norm()
        if st["dPOCS"] is None:
            st["dPOCS"] = dp * 0.2
        dd = e.data_distance()
        e.copy_recon(); e.tv_gd(10, st["dPOCS"])
        dg = e.matrix_2norm()
        if dg > dp * 0.95 and dd > 0.025:
            st["dPOCS"] *= 0.95

    return {
        "sirt": (lambda e: (e.initialize_SIRT(), e.initialize_FP(), e.restart_recon()), lambda e, st: e.SIRT(1), 3 * V + 3 * S),
        "sart_pass": (lambda e: (e.initialize_SART("sequential"), e.initialize_FP(), e.restart_recon()), lambda e, st: e.SART(0.5, 1), 2 * V + 2 * S),
        "kl": (lambda e: (e.initialize_poisson_ML(), e.initialize_FP(), e.restart_recon()), lambda e, st: e.poisson_ML(0.1), 3 * V + 3 * S),
        "cgls": (lambda e: (e.initialize_CGLS(), e.initialize_FP(), e.restart_recon()), lambda e, st: e.CGLS(1), None),
        "asd_pocs": (asd_setup, asd_step, 62 * V + 5 * S),
        "fbp": (lambda e: (e.initialize_FBP("ram-lak"), e.restart_recon()), lambda e, st: e.FBP(True), None),
    }


def run(eng, name, spec, steps, hbm_gbs):
    setup, step, alg = spec
    setup(eng)
    st = {"beta": 0.25, "dPOCS": None}
    step(eng, st)                                            # warm-up (allocations, weights)
    eng.sync(); eng.profile(True); l0 = eng.launch_count(); eng.timer_start()
    for _ in range(steps):
        step(eng, st)
    ms = eng.timer_stop() / steps
    prof = eng.profile_report(); eng.profile(False)
    line = {"workload": name, "ms_per_step": ms, "steps_per_s": 1e3 / ms, "gpu_launches_per_step": (eng.launch_count() - l0) / steps,
            "kernels": {g: {"ms_per_launch": t / c, "launches_per_step": c / steps, "share_of_step": t / (ms * steps)} for g, (t, c) in prof.items()}}
    if alg:
        line["algorithmic_bytes_per_step"] = alg
        line["frac_of_hbm_peak"] = alg / (ms * 1e-3) / 1e9 / hbm_gbs
    return line


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--size", type=int, default=512); ap.add_argument("--angles", type=int, default=180)
    ap.add_argument("--steps", type=int, default=5); ap.add_argument("--only", default="")
    a = ap.parse_args()
    import tomo_tv_b200 as ttv
    from tomo_tv_b200.phantom import phantom3d, tilt_angles_deg
    try:
        hbm = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")))["hbm_gbs"]
    except Exception:
        hbm = 6650.0
    n, na = a.size, a.angles
    eng = ttv.tomoengine(n, n, np.deg2rad(tilt_angles_deg(na)).astype(np.float32))
    eng.initialize_initial_volume(); eng.set_original_volume_all(phantom3d(n, n)); eng.create_projections()
    V, S, _ = eng.sizes()
    todo = workloads(V, S)
    for name in ([s for s in a.only.split(",") if s] or list(todo)):
        print(json.dumps(dict(run(eng, name, todo[name], a.steps, hbm), size=n, angles=na)), flush=True)


if __name__ == "__main__":
    main()
