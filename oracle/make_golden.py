#!/usr/bin/env python
"""TEST INFRASTRUCTURE -- generates tests/golden/*.npz by running the UNMODIFIED reference
(`oracle/_ref/tomoengine*.so`: tomo_TV's pybind11 engine + its ASTRA fork, built by
oracle/build_ref.sh from /root/reference) on a B200 through the scenarios of tests/scenarios.py.

Run on the GPU box:   python oracle/make_golden.py [--out gpurun_out/golden]
then copy gpurun_out/golden/*.npz into tests/golden/ and commit them.

For every scenario it also runs the CPU oracle (oracle.Oracle) and the CUDA engine
(tomo_tv_b200.tomoengine) on the same inputs and prints the three-way relative-L2 table
(reference vs oracle, reference vs CUDA, oracle vs CUDA) -> <out>/parity_table.txt.
"""
import argparse
import os
import sys
import time
import traceback

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, os.path.join(HERE, "_ref"))

import scenarios as S  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "golden"))
    ap.add_argument("--only", default="")
    ap.add_argument("--geoms", default="", help="comma-separated geometry names (default: all)")
    ap.add_argument("--spread", type=int, default=1, help="run the reference twice on the TV-GD scenarios and store the spread")
    args = ap.parse_args()
    os.makedirs(args.out, exist_ok=True)

    import oracle as orc
    import tomoengine as ref_mod                      # the reference extension
    import tomo_tv_b200 as ours

    lines = []

    def say(s):
        print(s, flush=True)
        lines.append(s)

    for geom, (nx, n, ang_deg) in S.GEOMS.items():
        if args.geoms and geom not in args.geoms.split(","):
            continue
        ang = S.angles_rad(geom)
        inp = S.make_inputs(geom)
        # measured data: oracle forward projection of the phantom (deterministic CPU code), stored
        o0 = orc.Oracle(nx, n, ang)
        o0.recon[:] = inp["vol"]
        o0.forward_projection()
        inp["b"] = o0.g.copy()
        # __cosf/__sinf of the projector angles as the B200 evaluates them (the reference kernels
        # use these intrinsics, par_fp.cu:64-65); stored so the CPU oracle can be checked offline
        probe = ours.tomoengine(nx, n, ang)
        cosv, sinv = probe.get_trig()
        del probe
        np.savez_compressed(os.path.join(args.out, f"{geom}_inputs.npz"), angles=ang, cosv=cosv, sinv=sinv, **S.digest(geom, inp))
        for scn, fn in S.scenarios_for(geom).items():
            if args.only and scn not in args.only.split(","):
                continue
            t0 = time.time()
            try:
                r_eng = ref_mod.tomoengine(nx, n, ang)
                want = fn(r_eng, inp, nx)
                del r_eng
            except Exception:
                say(f"{geom}/{scn}: REFERENCE FAILED\n{traceback.format_exc()}")
                continue
            np.savez_compressed(os.path.join(args.out, f"{geom}_{scn}.npz"), **S.digest(geom, want))
            t1 = time.time()
            if scn in ("asd", "tv", "gd10") and args.spread:
                # the reference's TV-GD kernel reads out of bounds (tv_gd.cu:73,76,79): a second run of the unmodified
                # reference on the same inputs measures its own run-to-run spread (the floor of any parity claim)
                r_eng = ref_mod.tomoengine(nx, n, ang)
                again = fn(r_eng, inp, nx)
                del r_eng
                sp = {k: np.float64(S.rel_l2(again[k], want[k])) for k in want}
                if scn == "asd":
                    # sensitivity of the UNMODIFIED reference to its input: the measured data moved by one ulp (random sign per
                    # value); the change of its ASD-POCS outputs is the floor below which no two float32 implementations agree
                    rng = np.random.default_rng(99)
                    b1 = np.nextafter(inp["b"], np.where(rng.random(inp["b"].shape) < 0.5, -np.inf, np.inf).astype(np.float32)).astype(np.float32)
                    r_eng = ref_mod.tomoengine(nx, n, ang)
                    pert = fn(r_eng, dict(inp, b=b1), nx)
                    del r_eng
                    for k in want:
                        sp[k + "_sens"] = np.float64(S.rel_l2(pert[k], want[k]))
                    say(f"{geom}/{scn}: reference with inputs moved by one ulp  " + "  ".join(f"{k}={sp[k + '_sens']:.3e}" for k in sorted(want)))
                np.savez_compressed(os.path.join(args.out, f"spread_{geom}_{scn}.npz"), **sp)
                say(f"{geom}/{scn}: reference run-to-run spread  " + "  ".join(f"{k}={v:.3e}" for k, v in sorted(sp.items())))
                t1 = time.time()
            res = {}
            for label, mk in (("oracle", lambda: orc.Oracle(nx, n, ang, quant=1, cosv=cosv, sinv=sinv)),
                              ("oracle_exactw", lambda: orc.Oracle(nx, n, ang, quant=0, cosv=cosv, sinv=sinv)),
                              ("cuda", lambda: ours.tomoengine(nx, n, ang))):
                try:
                    res[label] = fn(mk(), inp, nx)
                except Exception:
                    say(f"{geom}/{scn}: {label} FAILED\n{traceback.format_exc()}")
            for k in sorted(want):
                row = f"{geom}/{scn}/{k:8s}"
                for label in ("oracle", "oracle_exactw", "cuda"):
                    if label in res and k in res[label]:
                        row += f"  ref~{label}={S.rel_l2(res[label][k], want[k]):.3e}"
                if "oracle" in res and "cuda" in res and k in res["oracle"] and k in res["cuda"]:
                    row += f"  oracle~cuda={S.rel_l2(res['cuda'][k], res['oracle'][k]):.3e}"
                say(row)
                if (scn, k) in S.GD_MASKED:
                    # the same comparison away from the reference kernel's out-of-bounds reads (scenarios.gd_mask)
                    m = S.gd_mask(np.shape(want[k]), S.GD_MASKED[(scn, k)])
                    row = f"{geom}/{scn}/{k:8s}  [masked, {100.0 * m.mean():.1f}% of the voxels]"
                    for label in ("oracle", "cuda"):
                        if label in res and k in res[label]:
                            row += f"  ref~{label}={S.rel_l2(np.asarray(res[label][k])[m], np.asarray(want[k])[m]):.3e}"
                    say(row)
            say(f"# {geom}/{scn}: reference {t1 - t0:.2f}s, others {time.time() - t1:.2f}s")
    with open(os.path.join(args.out, "parity_table.txt"), "w") as f:
        f.write("\n".join(lines) + "\n")


if __name__ == "__main__":
    main()
