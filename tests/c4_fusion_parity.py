#!/usr/bin/env python
"""BASELINE.json config c4 (ChemicalTomo data fusion: ADF + chemical maps, KL/EM Poisson term) at a chosen size on one B200,
run through (a) the UNMODIFIED reference fusion engine (oracle/_ref/multimodal*.so, test infrastructure) and (b)
tomo_tv_b200.multimodal on identical inputs: relative L2 of the 4-D reconstruction after the data-fusion scenario of
tests/scenarios_fusion.py (Poisson-ML x4, rescale, 2 x [SIRT data fusion x3 + 4-D TV-FGP x5]) and both engines' wall times.
Inputs: disjoint-support element phantoms projected with this repo's forward projector (synthetic, seeded noise).
    python tests/c4_fusion_parity.py [--size 256] [--out gpurun_out/c4_fusion.json] [--no-ref]
"""
import argparse, json, os, sys, time
import numpy as np
HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT); sys.path.insert(0, HERE); sys.path.insert(0, os.path.join(ROOT, "oracle", "_ref"))
import scenarios_fusion as F


def make_inputs(ttv, n, els, hang, cang, seed=7):
    nel = len(els)
    rng = np.random.default_rng(seed + n)
    zz, yy, xx = np.meshgrid(np.arange(n), np.arange(n), np.arange(n), indexing="ij")
    vols = np.zeros((nel, n, n, n), np.float32)
    for e in range(nel):
        c = n / 2 + (n / 5) * np.array([np.cos(2.1 * e), np.sin(2.1 * e)])
        r2 = ((yy - c[0]) ** 2 + (xx - c[1]) ** 2 + 0.5 * (zz - n / 2) ** 2)
        vols[e] = (r2 < (n / 6.5) ** 2) * (0.5 + 0.5 * e / nel)
    w = F.weights(els)
    ha, ca = np.deg2rad(hang).astype(np.float32), np.deg2rad(cang).astype(np.float32)

    def project(vol, ang):
        e = ttv.tomoengine(n, n, ang)
        e.initialize_initial_volume(); e.set_original_volume_all(vol); e.create_projections()
        return e.get_projections()
    bh = project(sum(w[e] * vols[e] ** 1.6 for e in range(nel)).astype(np.float32), ha)
    bh = (bh / bh.max()).astype(np.float32)
    parts = []
    for e in range(nel):
        b = project(vols[e], ca) * rng.uniform(0.8, 1.2, (n, ca.size * n)).astype(np.float32)
        parts.append((b / b.max()).astype(np.float32))
    return {"bh": bh, "bchem": np.ascontiguousarray(np.concatenate(parts, axis=1)), "sigma": F.sigma_triplets(n, els), "hang": ha, "cang": ca}


_volume4d_slices = F.volume4d


def volume4d_fast(mm, nel, nx):
    if hasattr(mm, "get_recon_all"):
        return np.stack([mm.get_recon_all(e) for e in range(nel)])
    return _volume4d_slices(mm, nel, nx)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--size", type=int, default=256); ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "c4_fusion.json"))
    ap.add_argument("--no-ref", action="store_true")
    a = ap.parse_args()
    import tomo_tv_b200 as ttv
    n = a.size
    els = ["C", "Zn", "Au"]
    hang, cang = np.arange(-70.0, 71.0, 2.0), np.arange(-70.0, 71.0, 14.0)        # 71 ADF and 11 chemical tilt angles
    inp = make_inputs(ttv, n, els, hang, cang)
    nel = len(els)
    F.volume4d = volume4d_fast

    def run(mm):
        t = time.perf_counter()
        out = F._fusion(mm, inp, nel, n, 1.6, niter=2)
        return out, time.perf_counter() - t
    ours = ttv.multimodal(n, n, nel, inp["hang"], inp["cang"])
    run(ours)                                                   # warm-up (plans, weights, allocations)
    got, t_ours = run(ours)
    # one data-fusion iteration on its own (SIRT data fusion x3 + 4-D TV-FGP x5), device-timed
    ours.timer_start()
    for _ in range(5):
        ours.sirt_data_fusion(10.0, 0.05, 3); ours.tv_fgp_4D(5, 1e-4)
    ms_iter = ours.timer_stop() / 5
    res = {"config": f"ChemicalTomo data fusion, {nel} elements, {n}^3, {hang.size} ADF + {cang.size} chemical angles", "ours_scenario_s": t_ours,
           "ours_fusion_iteration_ms": ms_iter, "ours_fusion_iterations_per_s": 1e3 / ms_iter}
    if not a.no_ref:
        import multimodal as ref_mod
        ref = ref_mod.multimodal(n, n, nel, inp["hang"], inp["cang"])
        want, t_ref = run(ref)
        errs = {k: float(F.rel_l2(got[k], want[k])) for k in want}
        res.update(reference_scenario_s=t_ref, speedup=t_ref / t_ours, rel_l2=errs, tolerance=1e-4, ok=bool(max(errs.values()) <= 1e-4))
    os.makedirs(os.path.dirname(a.out), exist_ok=True)
    json.dump(res, open(a.out, "w"), indent=1)
    print(json.dumps(res))


if __name__ == "__main__":
    main()
