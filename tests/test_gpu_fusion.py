"""GPU parity tests of the multi-modal fusion path: the CUDA engine through
the C ABI (ttv_mm_*) against the CPU oracle (OracleMM) and against the golden vectors the REAL reference
(`multimodal`, oracle/_ref) produced on a B200 (tests/golden/fusion_*.npz, oracle/make_golden_fusion.py).
Tolerance: 1e-4 relative L2 after a fixed iteration count (BASELINE.json north_star)."""
import os

import numpy as np
import pytest

import scenarios_fusion as F
from conftest import GOLDEN

pytestmark = pytest.mark.gpu


def _trig(ttv, n, ang):
    e = ttv.tomoengine(n, n, ang)
    return e.get_trig()


@pytest.mark.parametrize("geom", list(F.FGEOMS))
@pytest.mark.parametrize("scn", list(F.FSCENARIOS))
def test_fusion_cuda_matches_oracle(ttv, orc, geom, scn):
    n, els, _, _ = F.FGEOMS[geom]
    nel = len(els)
    inp = F.make_inputs(geom, orc)
    ha, ca = inp["hang"], inp["cang"]
    o = orc.OracleMM(n, n, nel, ha, ca, 1, _trig(ttv, n, ha), _trig(ttv, n, ca))
    want = F.FSCENARIOS[scn](o, inp, nel, n)
    got = F.FSCENARIOS[scn](ttv.multimodal(n, n, nel, ha, ca), inp, nel, n)
    bad = [r for r in F.compare(scn, got, want) if not r[3]]
    assert not bad, bad


def _fusion_goldens():
    out = []
    if os.path.isdir(GOLDEN):
        for f in sorted(os.listdir(GOLDEN)):
            if f.startswith("fusion_") and f.endswith(".npz") and not f.endswith("_inputs.npz"):
                _, geom, scn = f[:-4].split("_", 2)
                out.append((geom, scn))
    return out


@pytest.mark.parametrize("geom,scn", _fusion_goldens())
def test_fusion_cuda_matches_reference_goldens(ttv, geom, scn):
    inp = dict(np.load(os.path.join(GOLDEN, f"fusion_{geom}_inputs.npz")))
    want = dict(np.load(os.path.join(GOLDEN, f"fusion_{geom}_{scn}.npz")))
    n, els, _, _ = F.FGEOMS[geom]
    got = F.FSCENARIOS[scn](ttv.multimodal(n, n, len(els), inp["hang"], inp["cang"]), inp, len(els), n)
    bad = [r for r in F.compare(scn, got, want) if not r[3]]
    assert not bad, bad


def test_fusion_non_cubic_and_slice_io(ttv, orc):
    """Nslice != Nray (the reference cannot run this: Sigma is sized with Nslice*Ny) and ragged sizes."""
    nx, n, nel = 37, 24, 2
    ha = np.deg2rad(np.arange(-60, 61, 20)).astype(np.float32)
    ca = np.deg2rad(np.arange(-60, 61, 40)).astype(np.float32)
    rng = np.random.default_rng(5)
    mm = ttv.multimodal(nx, n, nel, ha, ca)
    o = orc.OracleMM(nx, n, nel, ha, ca, 1, _trig_nx(ttv, nx, n, ha), _trig_nx(ttv, nx, n, ca))
    bh = rng.uniform(0.1, 1, (nx, ha.size * n)).astype(np.float32)
    bc = rng.uniform(0.1, 1, (nx, nel * ca.size * n)).astype(np.float32)
    sig = F.sigma_triplets(n, ["Co", "O"])
    x0 = rng.uniform(0, 1, (nel, nx, n, n)).astype(np.float32)
    for m in (mm, o):
        m.set_haadf_tilt_series(bh); m.set_chem_tilt_series(bc); m.set_gamma(1.6); m.load_sigma(sig)
        m.initialize_SIRT(); m.estimate_lipschitz(); m.set_measureChem(True); m.set_measureHaadf(True)
    for e in range(nel):
        for s in range(nx):
            mm.set_recon(x0[e, s], e, s)
    o.X[:] = x0
    assert np.array_equal(F.volume4d(mm, nel, nx), x0)
    assert np.array_equal(mm.get_chem_projections(), bc) and np.array_equal(mm.get_haadf_projections(), bh)
    a, b = mm.sirt_data_fusion(2.0, 0.05, 2), o.sirt_data_fusion(2.0, 0.05, 2)
    assert F.rel_l2(F.volume4d(mm, nel, nx), o.X) <= 1e-4
    assert abs(a[0] - b[0]) <= 1e-4 * abs(b[0]) and abs(a[1] - b[1]) <= 1e-4 * abs(b[1])
    assert abs(mm.tv_fgp_4D(3, 0.01) - o.tv_fgp_4D(3, 0.01)) <= 1e-4 * 1e3
    assert F.rel_l2(F.volume4d(mm, nel, nx), o.X) <= 1e-4


def _trig_nx(ttv, nx, n, ang):
    return ttv.tomoengine(nx, n, ang).get_trig()


def test_fusion_errors(ttv):
    ha = np.deg2rad(np.arange(-60, 61, 20)).astype(np.float32)
    mm = ttv.multimodal(8, 16, 2, ha, ha)
    with pytest.raises(ttv.NativeError):
        mm.sirt_data_fusion(1.0, 0.05, 1)            # estimate_lipschitz / load_sigma missing
    with pytest.raises(ttv.NativeError):
        mm.get_recon(2, 0)
    bad = np.array([[0.0], [3.0], [1.0]], np.float32)  # column not of the form p + Npix*e
    with pytest.raises(ttv.NativeError):
        mm.load_sigma(bad)
    with pytest.raises(ttv.NativeError):
        ttv.multimodal(8, 16, 9, ha, ha)
