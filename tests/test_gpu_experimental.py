"""Experimental kernels that have never run on a GPU (written at the end of round 1 without GPU budget).
They are NOT part of the default suite: a faulting kernel would poison the CUDA context of the whole pytest
process.  Run each variant in its own process (scripts/round2_first_call.sh does):
    TTV_EXPERIMENTAL=1 python -m pytest tests/test_gpu_experimental.py -m gpu -k bp9
"""
import os

import numpy as np
import pytest

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(os.environ.get("TTV_EXPERIMENTAL") != "1", reason="experimental kernels: set TTV_EXPERIMENTAL=1")]

SIZES = [(5, 40, 7), (33, 72, 17), (64, 128, 40), (40, 200, 24), (3, 512, 12)]
# name -> (option, value, bit-identical to the default kernel?)
BP_VARIANTS = {"bp8": ("bp_variant", 8, True),      # bp8_kernel FORM 0: same operations in the same order per sample
               "bp9": ("bp_variant", 9, False),     # bp8_kernel FORM 1: parity registers, taps added even-cell-first
               "bp5": ("bp_variant", 5, False)}     # bp_kernel<EPI, REUSE = 1>: parity registers in the default layout
FP_VARIANTS = {"fp1": ("fp_variant", 1, False),     # fp_kernel<EPI, REUSE = 1>
               "fp8": ("fp_variant", 8, False)}     # fp8_kernel: 8 slices per lane + parity registers


def _close(a, c, exact):
    if exact:
        return np.array_equal(a, c)
    return np.linalg.norm((np.asarray(a, np.float64) - c).ravel()) <= 1e-6 * np.linalg.norm(np.asarray(a, np.float64).ravel())


@pytest.mark.parametrize("nx,n,na", SIZES)
@pytest.mark.parametrize("name", list(BP_VARIANTS))
def test_bp_variant_matches_default(ttv, name, nx, n, na):
    """Back projection, the fused SIRT update and the KL step of an experimental back projector against the
    default bp_kernel: bit-identical where the operation order is the same, one rounding per sample otherwise."""
    opt, val, exact = BP_VARIANTS[name]
    rng = np.random.default_rng(nx * 7 + n)
    ang = np.sort(rng.uniform(-np.pi / 2, np.pi / 2, na)).astype(np.float32)
    sino = rng.standard_normal((nx, na * n)).astype(np.float32)
    b = rng.random((nx, na * n), dtype=np.float32) * 3
    outs = []
    for v in (4, val):
        eng = ttv.tomoengine(nx, n, ang)
        eng.set_option(opt, v)
        eng.initialize_initial_volume()
        eng.set_model_projections(sino); eng.back_projection()
        bp = eng.get_recon_all()
        eng.set_tilt_series(b); eng.initialize_SIRT(); eng.initialize_FP(); eng.restart_recon(); eng.SIRT(3)
        sirt = eng.get_recon_all()
        eng.initialize_poisson_ML(); eng.restart_recon()
        eng.set_recon_all(np.full((nx, n, n), 0.3, np.float32))
        cost = [eng.poisson_ML(0.4) for _ in range(2)]
        outs.append((bp, sirt, eng.get_recon_all(), cost))
    for a, c in zip(outs[0][:3], outs[1][:3]):
        assert _close(a, c, exact)
    assert np.allclose(outs[0][3], outs[1][3], rtol=0 if exact else 1e-5)


@pytest.mark.parametrize("nx,n,na", SIZES)
@pytest.mark.parametrize("name", list(FP_VARIANTS))
def test_fp_variant_matches_default(ttv, name, nx, n, na):
    """Forward projection, three SIRT steps and the data distance of an experimental forward projector against the
    default fp_kernel (the parity-register forms add the two taps of a sample even-cell-first: one rounding)."""
    opt, val, exact = FP_VARIANTS[name]
    rng = np.random.default_rng(nx * 11 + n)
    ang = np.sort(rng.uniform(-np.pi / 2, np.pi / 2, na)).astype(np.float32)
    vol = rng.random((nx, n, n), dtype=np.float32)
    b = rng.random((nx, na * n), dtype=np.float32) * 3
    outs = []
    for v in (0, val):
        eng = ttv.tomoengine(nx, n, ang)
        eng.set_option(opt, v)
        eng.initialize_initial_volume()
        eng.set_recon_all(vol); eng.forward_projection()
        g = eng.get_model_projections()
        eng.set_tilt_series(b); eng.initialize_SIRT(); eng.initialize_FP(); eng.restart_recon(); eng.SIRT(3)
        outs.append((g, eng.get_recon_all(), eng.data_distance()))
    for a, c in zip(outs[0][:2], outs[1][:2]):
        assert _close(a, c, exact)
    assert abs(outs[0][2] - outs[1][2]) <= 1e-5 * abs(outs[0][2])
