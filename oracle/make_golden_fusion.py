#!/usr/bin/env python
"""TEST INFRASTRUCTURE -- golden vectors of the multi-modal fusion path from the UNMODIFIED reference
(`oracle/_ref/multimodal*.so`: tomofusion/chemistry/utils/multimodal.cpp + its ASTRA fork, built by
oracle/build_ref.sh) run on a B200 through tests/scenarios_fusion.py.

Run on the GPU box:   python oracle/make_golden_fusion.py [--out gpurun_out/golden]
then copy <out>/fusion_*.npz into tests/golden/ and commit them.  Also runs the CPU oracle (OracleMM)
and the CUDA engine on the same inputs and writes the three-way table <out>/parity_table_fusion.txt.
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

import scenarios_fusion as F  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "golden"))
    args = ap.parse_args()
    os.makedirs(args.out, exist_ok=True)
    import oracle as orc
    import multimodal as ref_mod                      # the reference extension
    import tomo_tv_b200 as ours

    lines = []

    def say(s):
        print(s, flush=True)
        lines.append(s)

    for geom, (n, els, _, _) in F.FGEOMS.items():
        nel = len(els)
        inp = F.make_inputs(geom, orc)
        ha, ca = inp["hang"], inp["cang"]
        probe = ours.multimodal(n, n, nel, ha, ca)
        del probe
        ph, pc = ours.tomoengine(n, n, ha), ours.tomoengine(n, n, ca)
        trig_h, trig_c = ph.get_trig(), pc.get_trig()
        del ph, pc
        np.savez_compressed(os.path.join(args.out, f"fusion_{geom}_inputs.npz"), cos_h=trig_h[0], sin_h=trig_h[1],
                            cos_c=trig_c[0], sin_c=trig_c[1], **inp)
        for scn, fn in F.FSCENARIOS.items():
            t0 = time.time()
            try:
                r = ref_mod.multimodal(n, n, nel, ha, ca)
                want = fn(r, inp, nel, n)
                del r
            except Exception:
                say(f"{geom}/{scn}: REFERENCE FAILED\n{traceback.format_exc()}")
                continue
            np.savez_compressed(os.path.join(args.out, f"fusion_{geom}_{scn}.npz"), **want)
            t1 = time.time()
            res = {}
            for label, mk in (("oracle", lambda: orc.OracleMM(n, n, nel, ha, ca, 1, trig_h, trig_c)),
                              ("cuda", lambda: ours.multimodal(n, n, nel, ha, ca))):
                try:
                    res[label] = fn(mk(), inp, nel, n)
                except Exception:
                    say(f"{geom}/{scn}: {label} FAILED\n{traceback.format_exc()}")
            for k in sorted(want):
                row = f"fusion {geom}/{scn}/{k:12s}"
                for label in ("oracle", "cuda"):
                    if label in res and k in res[label]:
                        row += f"  ref~{label}={F.rel_l2(res[label][k], want[k]):.3e}"
                if len(res) == 2 and k in res["oracle"] and k in res["cuda"]:
                    row += f"  oracle~cuda={F.rel_l2(res['cuda'][k], res['oracle'][k]):.3e}"
                say(row)
            say(f"# fusion {geom}/{scn}: reference {t1 - t0:.2f}s, others {time.time() - t1:.2f}s")
    with open(os.path.join(args.out, "parity_table_fusion.txt"), "w") as f:
        f.write("\n".join(lines) + "\n")


if __name__ == "__main__":
    main()
