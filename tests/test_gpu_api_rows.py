"""Small API rows of SURVEY.md 8(f) that had no test in round 1 (missing 3/4/6):
update_projection_angles, poisson_noise, and the 4-D TV-GD / RMSE of the fusion engine."""
import numpy as np
import pytest

import scenarios_fusion as F

pytestmark = pytest.mark.gpu


def _rel(a, c):
    a = np.asarray(a, np.float64)
    return float(np.linalg.norm((a - c).ravel()) / max(np.linalg.norm(a.ravel()), 1e-30))


@pytest.mark.parametrize("nx,n", [(20, 48), (70, 64)])
def test_update_projection_angles_equals_fresh_engine(ttv, nx, n):
    """T/gpu/utils/tomoengine.cpp:130-149: new angles replace the geometry and resize b, g; recon is kept.
    Everything derived from the angles (FP plan, __constant__ BP table, R / C weights, Lipschitz constant, SART
    state) must be rebuilt: FP, SIRT, KL and FISTA's L after the update equal a fresh engine with the new angles."""
    rng = np.random.default_rng(nx)
    a1 = np.deg2rad(np.linspace(-70, 70, 15)).astype(np.float32)
    a2 = np.deg2rad(np.linspace(-88, 86, 23)).astype(np.float32)
    vol = rng.random((nx, n, n), dtype=np.float32)
    b1 = rng.random((nx, a1.size * n), dtype=np.float32)
    b2 = rng.random((nx, a2.size * n), dtype=np.float32) + 0.1

    eng = ttv.tomoengine(nx, n, a1)
    eng.initialize_initial_volume(); eng.set_tilt_series(b1)
    eng.initialize_SIRT(); eng.initialize_FP(); eng.initialize_fista(); eng.restart_recon()
    eng.SIRT(2)                                               # weights, plan and constants of the OLD angles are live
    eng.set_recon_all(vol)
    eng.update_projection_angles(a2)
    assert np.array_equal(eng.get_recon_all(), vol)           # recon preserved
    assert eng.get_projections().shape == (nx, a2.size * n)   # b, g resized

    fresh = ttv.tomoengine(nx, n, a2)
    fresh.initialize_initial_volume(); fresh.set_recon_all(vol)
    outs = []
    for e in (eng, fresh):
        e.set_tilt_series(b2)
        e.forward_projection()
        g = e.get_model_projections()
        e.initialize_SIRT(); e.initialize_FP(); e.initialize_fista()
        L = e.get_lipschitz()
        e.SIRT(3)
        rec = e.get_recon_all()
        dd = e.data_distance()
        e.initialize_poisson_ML(); e.set_recon_all(np.full((nx, n, n), 0.3, np.float32))
        kl = e.poisson_ML(0.4)
        e.initialize_SART("sequential"); e.set_recon_all(vol); e.SART(0.5, 1)
        outs.append((g, rec, e.get_recon_all(), L, dd, kl))
    for x, y in zip(outs[0][:3], outs[1][:3]):
        assert np.array_equal(x, y)
    assert outs[0][3:] == outs[1][3:]

    # the constructor without angles + update_projection_angles is the reference's two-step construction (tomoengine.cpp:51-60)
    late = ttv.tomoengine(nx, n)
    late.update_projection_angles(a2)
    late.initialize_initial_volume(); late.set_recon_all(vol); late.set_tilt_series(b2); late.forward_projection()
    assert np.array_equal(late.get_model_projections(), outs[1][0])


def test_poisson_noise_statistics(ttv):
    """T/gpu/utils/tomoengine.cpp:471-484: b <- Poisson(b / sum(b) * Nc * size) / (Nc * size) * sum(b).
    On a constant sinogram every ray has mean Nc counts: the result keeps the mean and has variance c^2 / Nc;
    the total is preserved to sampling error and a second engine draws the same numbers (default-seeded engine)."""
    nx, n, na, Nc, c = 64, 64, 30, 50, 3.0
    ang = np.deg2rad(np.linspace(-75, 75, na)).astype(np.float32)
    outs = []
    for _ in range(2):
        eng = ttv.tomoengine(nx, n, ang)
        eng.set_tilt_series(np.full((nx, na * n), c, np.float32))
        eng.poisson_noise(Nc)
        outs.append(eng.get_projections())
    b = outs[0].astype(np.float64)
    assert np.array_equal(outs[0], outs[1])
    m = b.size
    assert abs(b.mean() - c) <= 5 * c / np.sqrt(Nc * m)
    assert abs(b.var() / (c * c / Nc) - 1) <= 5 * np.sqrt(2.0 / m) + 1e-3
    counts = b * Nc / c
    assert np.abs(counts - np.rint(counts)).max() <= 1e-3         # integer counts, rescaled
    # a non-uniform sinogram: noise scales with the signal (variance of ray i = b_i * sum / (Nc * size))
    ramp = np.tile(np.linspace(0.5, 4.0, na * n, dtype=np.float32), (nx, 1))
    eng = ttv.tomoengine(nx, n, ang)
    eng.set_tilt_series(ramp); eng.poisson_noise(Nc)
    nb = eng.get_projections().astype(np.float64)
    scale = ramp.astype(np.float64).sum() / (Nc * ramp.size)
    z = (nb - ramp) / np.sqrt(ramp * scale)
    assert abs(z.mean()) <= 0.02 and abs(z.std() - 1) <= 0.02


@pytest.mark.parametrize("geom", list(F.FGEOMS))
def test_fusion_tv_gd_4d_and_rmse_4d(ttv, orc, geom):
    """cuda_tv_gd_4D (T/chemistry/utils/regularizers/tv_gd.cu:208: per-element TV-GD, TV values summed) and
    cuda_rmse_4D (T/chemistry/utils/container/matrix_ops.cu:568-615: sqrt(mean((recon - original)^2)) per element)."""
    n, els, _, _ = F.FGEOMS[geom]
    nel = len(els)
    inp = F.make_inputs(geom, orc)
    ha, ca = inp["hang"], inp["cang"]
    th, tc = ttv.tomoengine(n, n, ha).get_trig(), ttv.tomoengine(n, n, ca).get_trig()
    mm = ttv.multimodal(n, n, nel, ha, ca)
    o = orc.OracleMM(n, n, nel, ha, ca, 1, th, tc)
    rng = np.random.default_rng(11)
    x0 = rng.uniform(0, 1, (nel, n, n, n)).astype(np.float32)
    x0[:, :, n // 4: n // 2, n // 3:] += 1.0                  # a step, so that TV-GD has something to flatten
    gt = rng.uniform(0, 1, (nel, n, n, n)).astype(np.float32)
    mm.initialize_initial_volume()
    for e in range(nel):
        mm.set_recon_all(x0[e], e)
        for s in range(n):
            mm.set_original_volume(gt[e, s], e, s)
    o.X[:] = x0
    want_rmse = np.sqrt(((x0.astype(np.float64) - gt) ** 2).reshape(nel, -1).mean(axis=1))
    assert np.allclose(mm.rmse(), want_rmse, rtol=2e-6)
    tv_g, tv_o = mm.tv_gd(4, 0.3), o.tv_gd(4, 0.3)
    assert abs(tv_g - tv_o) <= 1e-5 * abs(tv_o)
    assert _rel(o.X, F.volume4d(mm, nel, n)) <= 1e-5
    assert _rel(x0, o.X) > 1e-3                               # the step really moved the volume
    # rmse follows the update
    want_rmse = np.sqrt(((o.X.astype(np.float64) - gt) ** 2).reshape(nel, -1).mean(axis=1))
    assert np.allclose(mm.rmse(), want_rmse, rtol=1e-5)
