"""Golden vectors for BASELINE.json configs[0]: 2-D Shepp-Logan 256^2 x 60 angles on the reference's CPU
package (tomofusion/cpu: Eigen sparse system matrix + ART / SIRT, OpenMP over slices).

Test infrastructure.  Runs IN THE BUILD CONTAINER only (needs /root/reference): the reference's own
`parallelRay` (tomofusion/cpu/utils/pytvlib.py:8-121, executed from the reference file -- its module imports
skimage / h5py, which this image lacks, so the two functions are pulled out of the parsed source) builds the
measurement matrix, and the reference's compiled engine oracle/_ref/ctvlib*.so (oracle/build_ref.sh) creates
the projections and reconstructs, following tomofusion/cpu/sim_tomo.py:14-19,38-60 (beta0 = 0.5, x0.995 per
iteration, 100 iterations of ART; SIRT with beta = 1/lipschits).  Writes tests/golden/cpuref_c1.npz.
"""
import ast
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle", "_ref"))
REF_PY = "/root/reference/tomofusion/cpu/utils/pytvlib.py"


def reference_functions(names):
    src = open(REF_PY).read()
    tree = ast.parse(src)
    keep = [n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name in names]
    ns = {"np": np, "tqdm": (lambda it: it)}
    exec(compile(ast.Module(body=keep, type_ignores=[]), REF_PY, "exec"), ns)
    return [ns[n] for n in names]


def main():
    import ctvlib                                   # the reference's CPU engine, unmodified
    from tomo_tv_b200.phantom import shepp_logan_2d
    parallelRay, _ = reference_functions(["parallelRay", "rmepsilon"])
    n, na, niter = 256, 60, 100
    angles = np.linspace(-90.0, 90.0, na, endpoint=False)
    t0 = time.perf_counter()
    A = parallelRay(n, angles)
    t_matrix = time.perf_counter() - t0
    vol = shepp_logan_2d(n)
    out = {"angles_deg": angles, "n": n, "t_matrix_s": t_matrix}
    for alg in ("ART", "SIRT"):
        tomo = ctvlib.ctvlib(1, n, na)
        tomo.load_A(A)
        if alg == "ART":
            tomo.row_inner_product()
        tomo.initialize_original_volume()
        tomo.set_original_volume(vol, 0)
        tomo.create_projections()
        beta = 0.5 if alg == "ART" else 1.0 / tomo.lipschits()
        dd, rm = np.zeros(niter), np.zeros(niter)
        t0 = time.perf_counter()
        for i in range(niter):
            if alg == "ART":
                tomo.ART(beta); beta *= 0.995
            else:
                tomo.SIRT(beta)
            dd[i] = tomo.data_distance(); rm[i] = tomo.rmse()
        sec = time.perf_counter() - t0
        out[f"{alg}_recon"] = np.asarray(tomo.get_recon(0), np.float32)
        out[f"{alg}_dd"], out[f"{alg}_rmse"], out[f"{alg}_seconds"] = dd, rm, sec
        if alg == "ART":
            out["b"] = np.asarray(tomo.get_projections(), np.float32)       # [1, Na*N]
        print(f"{alg}: {niter} iterations in {sec:.2f} s ({niter / sec:.1f} it/s, 1 slice -> 1 core), rmse {rm[-1]:.4f}, dd {dd[-1]:.4f}")
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "cpuref_c1.npz"), **out)


if __name__ == "__main__":
    main()
