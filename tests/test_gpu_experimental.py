"""Experimental kernels that have never run on a GPU (written at the end of round 1 without GPU budget).
They are NOT part of the default suite: a faulting kernel would poison the CUDA context of the whole pytest
process.  Run them in their own process:  TTV_EXPERIMENTAL=1 python -m pytest tests/test_gpu_experimental.py -m gpu
"""
import os

import numpy as np
import pytest

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(os.environ.get("TTV_EXPERIMENTAL") != "1", reason="experimental kernels: set TTV_EXPERIMENTAL=1")]


@pytest.mark.parametrize("nx,n,na", [(5, 40, 7), (33, 72, 17), (64, 128, 40), (40, 200, 24), (3, 512, 12)])
def test_bp_variant_8_matches_default(ttv, nx, n, na):
    """bp8_kernel (ttv_bp8.cuh: 8 slices per lane, register reuse along the voxel run).  FORM 0 (variant 8) performs
    the same operations in the same order per sample as bp_kernel: back projection, the fused SIRT update and the
    KL step must be bit-identical.  FORM 1 (variant 9) may differ by one rounding per sample."""
    rng = np.random.default_rng(nx * 7 + n)
    ang = np.sort(rng.uniform(-np.pi / 2, np.pi / 2, na)).astype(np.float32)
    sino = rng.standard_normal((nx, na * n)).astype(np.float32)
    b = rng.random((nx, na * n), dtype=np.float32) * 3
    outs = []
    for variant in (4, 8, 9, 5):
        eng = ttv.tomoengine(nx, n, ang)
        eng.set_option("bp_variant", variant)
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
        assert np.array_equal(a, c)
    assert outs[0][3] == outs[1][3]
    # variant 9 (parity registers) adds the two taps of a sample even-cell-first: one rounding of the partial sum
    # (so does variant 5, the same reuse inside bp_kernel's own 4-slices-per-lane layout)
    for other in outs[2:]:
        for a, c in zip(outs[0][:3], other[:3]):
            assert np.linalg.norm((a - c).ravel()) <= 1e-6 * np.linalg.norm(a.ravel())
        assert np.allclose(outs[0][3], other[3], rtol=1e-5)


@pytest.mark.parametrize("nx,n,na", [(5, 40, 7), (33, 72, 17), (64, 128, 40), (40, 200, 24), (3, 512, 12)])
def test_fp_variants_match_default(ttv, nx, n, na):
    """fp_kernel<EPI, REUSE = 1> (parity registers: a held cell is not loaded again) adds the two taps of a sample
    even-cell-first, so it agrees with the default kernel to one rounding per sample, not bit-exactly."""
    rng = np.random.default_rng(nx * 11 + n)
    ang = np.sort(rng.uniform(-np.pi / 2, np.pi / 2, na)).astype(np.float32)
    vol = rng.random((nx, n, n), dtype=np.float32)
    b = rng.random((nx, na * n), dtype=np.float32) * 3
    outs = []
    for variant in (0, 1, 8):
        eng = ttv.tomoengine(nx, n, ang)
        eng.set_option("fp_variant", variant)
        eng.initialize_initial_volume()
        eng.set_recon_all(vol); eng.forward_projection()
        g = eng.get_model_projections()
        eng.set_tilt_series(b); eng.initialize_SIRT(); eng.initialize_FP(); eng.restart_recon(); eng.SIRT(3)
        outs.append((g, eng.get_recon_all(), eng.data_distance()))
    for other in outs[1:]:                 # variant 1 (same layout, reuse) and variant 8 (fp8_kernel: 8 slices per lane)
        for a, c in zip(outs[0][:2], other[:2]):
            assert np.linalg.norm((a - c).ravel()) <= 1e-6 * np.linalg.norm(a.ravel())
        assert abs(outs[0][2] - other[2]) <= 1e-5 * abs(outs[0][2])
