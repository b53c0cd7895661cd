"""Table-driven projectors (ttv_fpt.cuh / ttv_bpt.cuh, the defaults for thick slabs) against the classic kernels
(ttv_fp.cuh / ttv_bp.cuh, still the defaults for thin slabs and the ones the round-1 goldens were first matched
with): same sample positions, same 8-bit weights and the same accumulation order (even cell first, then odd cell,
lines / angles ascending), so the two families are BIT-IDENTICAL -- which is what makes a volume independent of the
number of GPUs it was sharded over (thin slabs run the classic kernels, thick ones the table-driven ones)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

# (Nx, N, Na): one x-block, ragged x-block counts (odd, < 8, > 8), ragged N (tile overhang), N > 256 windows
SIZES = [(5, 40, 7), (33, 72, 17), (64, 128, 40), (40, 200, 24), (3, 512, 12), (300, 96, 31), (257, 64, 9)]


def _rel(a, c):
    a = np.asarray(a, np.float64)
    return float(np.linalg.norm((a - c).ravel()) / max(np.linalg.norm(a.ravel()), 1e-30))


@pytest.mark.parametrize("nx,n,na", SIZES)
def test_bpt_matches_bp(ttv, nx, n, na):
    """Back projection, the fused SIRT update and the KL step (A/cuda/2d/par_bp.cu:52-87, sirt.cu:259-267,
    tomoengine.cpp:302-313)."""
    rng = np.random.default_rng(nx * 7 + n)
    ang = np.sort(rng.uniform(-np.pi / 2, np.pi / 2, na)).astype(np.float32)
    sino = rng.standard_normal((nx, na * n)).astype(np.float32)
    b = rng.random((nx, na * n), dtype=np.float32) * 3
    outs = []
    for v, xb in ((1, 0), (2, 2), (2, 8)):
        eng = ttv.tomoengine(nx, n, ang)
        eng.set_option("bp_variant", v); eng.set_option("bp_xb", xb); eng.set_option("fp_variant", 1)
        eng.initialize_initial_volume()
        eng.set_model_projections(sino); eng.back_projection()
        bp = eng.get_recon_all()
        eng.set_tilt_series(b); eng.initialize_SIRT(); eng.initialize_FP(); eng.restart_recon(); eng.SIRT(3)
        sirt = eng.get_recon_all()
        eng.initialize_poisson_ML(); eng.restart_recon()
        eng.set_recon_all(np.full((nx, n, n), 0.3, np.float32))
        cost = [eng.poisson_ML(0.4) for _ in range(2)]
        outs.append((bp, sirt, eng.get_recon_all(), cost))
    for other in outs[1:]:
        for a, c in zip(outs[0][:3], other[:3]):
            assert np.array_equal(a, c)
        assert np.allclose(outs[0][3], other[3], rtol=1e-5)


@pytest.mark.parametrize("nx,n,na", SIZES)
def test_fpt_matches_fp(ttv, nx, n, na):
    """Forward projection, three SIRT steps, the data distance and the KL cost (A/cuda/2d/par_fp.cu:53-200,
    sirt.cu:246-257, tomoengine.cpp:301-311,410-413)."""
    rng = np.random.default_rng(nx * 11 + n)
    ang = np.sort(rng.uniform(-np.pi / 2, np.pi / 2, na)).astype(np.float32)
    vol = rng.random((nx, n, n), dtype=np.float32)
    b = rng.random((nx, na * n), dtype=np.float32) * 3
    outs = []
    for v, xb in ((1, 0), (2, 2), (2, 1), (2, 3)):   # classic; table-driven with 2 x-blocks per CTA; with 1 (thin slabs); both in one launch
        eng = ttv.tomoengine(nx, n, ang)
        eng.set_option("fp_variant", v); eng.set_option("fp_xb", xb); eng.set_option("bp_variant", 1)
        eng.initialize_initial_volume()
        eng.set_recon_all(vol); eng.forward_projection()
        g = eng.get_model_projections()
        eng.set_tilt_series(b); eng.initialize_SIRT(); eng.initialize_FP(); eng.restart_recon(); eng.SIRT(3)
        rec = eng.get_recon_all()
        dd = eng.data_distance()
        eng.initialize_poisson_ML(); eng.set_recon_all(np.full((nx, n, n), 0.3, np.float32))
        outs.append((g, rec, dd, eng.poisson_ML(0.4)))
    for other in outs[1:]:
        for a, c in zip(outs[0][:2], other[:2]):
            assert np.array_equal(a, c)
        assert abs(outs[0][2] - other[2]) <= 1e-5 * abs(outs[0][2])
        assert abs(outs[0][3] - other[3]) <= 1e-5 * abs(outs[0][3])


def test_default_choice_by_slab_thickness(ttv):
    """Default = table-driven kernels from 2 x-blocks on (BP: 2 x-blocks per CTA below 7 x-blocks, 8 above), classic for one;
    both give the same reconstruction bit for bit."""
    rng = np.random.default_rng(3)
    nx, n, na = 200, 64, 20
    ang = np.sort(rng.uniform(-np.pi / 2, np.pi / 2, na)).astype(np.float32)
    b = rng.random((nx, na * n), dtype=np.float32)
    recs = []
    for fv, bv in ((0, 0), (1, 1)):
        eng = ttv.tomoengine(nx, n, ang)
        eng.set_option("fp_variant", fv); eng.set_option("bp_variant", bv)
        eng.initialize_initial_volume(); eng.set_tilt_series(b); eng.initialize_SIRT(); eng.initialize_FP(); eng.restart_recon()
        eng.SIRT(4)
        recs.append(eng.get_recon_all())
    assert np.array_equal(recs[1], recs[0])
    with pytest.raises(Exception):
        ttv.tomoengine(nx, n, ang).set_option("bp_variant", 9)


@pytest.mark.parametrize("nx", [64, 96, 128, 256])
def test_fp_planned_mixed_launch_at_slab_sizes(ttv, nx):
    """The launch plan the engine picks by itself (fp_xb = 0) on the slabs of a 512^3 x 180 volume sharded over 8, 4 and 2 GPUs (and a three-x-block slab):
    384 single-x-block items -> 148 pair CTAs + 88 short ones, 768 pair items -> 740 + 56 short ones (fptm_kernel,
    Engine::fp_mix_plan).  Same sinogram, bit for bit, as the launch with one CTA size."""
    n, na = 512, 180
    rng = np.random.default_rng(nx)
    ang = np.deg2rad(np.linspace(-90, 90, na, endpoint=False)).astype(np.float32)
    vol = rng.random((nx, n, n), dtype=np.float32)
    outs = []
    for xb in (0, 1):
        eng = ttv.tomoengine(nx, n, ang)
        eng.set_option("fp_xb", xb)
        eng.initialize_initial_volume(); eng.set_recon_all(vol); eng.forward_projection()
        outs.append(eng.get_model_projections())
        del eng
    assert np.array_equal(outs[0], outs[1])
