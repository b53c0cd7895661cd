"""BASELINE.json configs[0]: 2-D Shepp-Logan 256^2 x 60 angles, reconstructed by the reference's own CPU
package (tomofusion/cpu: Siddon system matrix from parallelRay + ART / SIRT in ctvlib.cpp).  The golden file
tests/golden/cpuref_c1.npz was produced by oracle/make_golden_c1.py from the unmodified reference.

The CPU package discretises the projector differently (exact ray / pixel intersection lengths) from the GPU
path this repo replaces (ASTRA's Joseph kernels with 8-bit texture weights), and ART (row action) is not SART
(angle-block action), so agreement is bounded by the discretisation, not by float32: the sinograms differ by
0.76 %, the 100-iteration reconstructions by 8.8 %.  The tolerances below are those measured values with
headroom; the tight parity of the GPU path is pinned elsewhere (tests/golden from the reference GPU engine).
"""
import os

import numpy as np
import pytest

from conftest import GOLDEN

N, NA = 256, 60


def _load():
    g = np.load(os.path.join(GOLDEN, "cpuref_c1.npz"))
    from tomo_tv_b200.phantom import shepp_logan_2d
    return g, shepp_logan_2d(N), np.deg2rad(g["angles_deg"]).astype(np.float32)


def _rel(a, b):
    return float(np.linalg.norm((a - b).ravel()) / np.linalg.norm(b.ravel()))


def _sart100(e, b):
    e.set_tilt_series(b); e.initialize_SART("sequential"); e.initialize_FP(); e.restart_recon()
    beta = 0.5
    for _ in range(100):                      # tomofusion/cpu/sim_tomo.py:14-19,44-52: beta0 = 0.5, x0.995 per iteration
        e.SART(beta, 1); beta *= 0.995
    return e.get_recon(0)


def _check_recon(rec, g, vol):
    rmse = float(np.sqrt(np.mean((rec - vol) ** 2)))
    ref_rmse = float(np.sqrt(np.mean((g["ART_recon"] - vol) ** 2)))
    assert abs(ref_rmse - float(g["ART_rmse"][-1])) < 1e-4          # the golden's own bookkeeping
    assert rmse <= 1.15 * ref_rmse, (rmse, ref_rmse)
    assert _rel(rec, g["ART_recon"]) <= 0.12


def test_same_geometry_convention_as_the_cpu_package(orc):
    """Same angles, same [row, column] orientation, same detector order: the oracle's projection of the phantom
    matches the sinogram the reference CPU engine created, up to the Siddon-vs-Joseph discretisation."""
    g, vol, ang = _load()
    o = orc.Oracle(1, N, ang)
    o.recon[0] = vol
    o.forward_projection()
    assert _rel(o.g, g["b"]) <= 1e-2


def test_oracle_sart_reconstructs_like_the_cpu_reference_art(orc):
    g, vol, ang = _load()
    _check_recon(_sart100(orc.Oracle(1, N, ang), g["b"]), g, vol)


@pytest.mark.gpu
def test_cuda_sart_and_sirt_on_the_cpu_reference_sinogram(ttv, orc):
    g, vol, ang = _load()
    eng = ttv.tomoengine(1, N, ang)
    rec = _sart100(eng, g["b"])
    _check_recon(rec, g, vol)
    cosv, sinv = eng.get_trig()
    want = _sart100(orc.Oracle(1, N, ang, quant=1, cosv=cosv, sinv=sinv), g["b"])
    assert _rel(rec, want) <= 1e-4                                   # north_star: reconstructions within 1e-4
    # SIRT (ASTRA's R/C-weighted form) against the CPU package's Landweber SIRT: same fixed point, different
    # preconditioning -- 100 iterations agree to 10 % and ours is at least as close to the phantom
    eng.initialize_SIRT(); eng.restart_recon(); eng.SIRT(100)
    sirt = eng.get_recon(0)
    assert _rel(sirt, g["SIRT_recon"]) <= 0.12
    assert np.sqrt(np.mean((sirt - vol) ** 2)) <= float(g["SIRT_rmse"][-1]) * 1.05
