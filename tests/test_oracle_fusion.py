"""CPU tests of the fusion oracle (OracleMM) against the golden vectors the REAL reference's
`multimodal` extension produced on a B200 (tests/golden/fusion_*.npz), and of the host-side
summation-matrix builder.  No GPU needed."""
import os

import numpy as np
import pytest

import scenarios_fusion as F
from conftest import GOLDEN


def _cases():
    out = []
    if os.path.isdir(GOLDEN):
        for f in sorted(os.listdir(GOLDEN)):
            if f.startswith("fusion_") and f.endswith(".npz") and not f.endswith("_inputs.npz"):
                _, geom, scn = f[:-4].split("_", 2)
                out.append((geom, scn))
    return out


def test_fusion_goldens_present():
    assert len(_cases()) >= 6


@pytest.mark.parametrize("geom,scn", _cases())
def test_oracle_matches_reference_fusion_goldens(orc, geom, scn):
    inp = dict(np.load(os.path.join(GOLDEN, f"fusion_{geom}_inputs.npz")))
    want = dict(np.load(os.path.join(GOLDEN, f"fusion_{geom}_{scn}.npz")))
    n, els, _, _ = F.FGEOMS[geom]
    o = orc.OracleMM(n, n, len(els), inp["hang"], inp["cang"], 1, (inp["cos_h"], inp["sin_h"]), (inp["cos_c"], inp["sin_c"]))
    got = F.FSCENARIOS[scn](o, inp, len(els), n)
    bad = [r for r in F.compare(scn, got, want) if not r[3]]
    assert not bad, bad


def test_inputs_regenerate_bit_exactly(orc):
    """The committed inputs are what make_inputs produces here (the generator is deterministic CPU code)."""
    for geom in F.FGEOMS:
        inp = dict(np.load(os.path.join(GOLDEN, f"fusion_{geom}_inputs.npz")))
        again = F.make_inputs(geom, orc)
        for k in ("bh", "bchem", "sigma"):
            assert np.array_equal(inp[k], again[k]), (geom, k)


def test_summation_matrix_builder_matches_reference_loop():
    """chemistry.create_weighted_summation_matrix == the per-pixel loop of utils/fusion_helper.py:5-32
    followed by the [rows, cols, vals] float32 packing of chemistry/reconstructor.py:150-153."""
    from tomo_tv_b200.chemistry import create_weighted_summation_matrix, get_periodic_table
    nx, nz = 6, 3
    zn = np.array([6, 30, 79])
    for method in range(5):
        vals = np.zeros(nz * nx * nx, np.float16); row = np.zeros(nz * nx * nx, int); col = np.zeros(nz * nx * nx, int)
        vals[:] = 1
        ii = ind = 0
        while ii < nz * nx * nx:
            for jj in range(nz):
                row[ii + jj] = ind; col[ii + jj] = ind + nx * nx * jj
                if method == 1: vals[ii + jj] = zn[jj] / np.mean(zn)
                if method == 2: vals[ii + jj] = zn[jj] ** 1.6 / np.mean(zn ** 1.6)
                if method == 3: vals[ii + jj] = zn[jj] / np.sum(zn)
                if method == 4: vals[ii + jj] = zn[jj] ** 1.6 / np.sum(zn ** 1.6)
            ii += nz; ind += 1
        want = np.array([row, col, vals], dtype=np.float32)
        got = create_weighted_summation_matrix(nx, nx, nz, zn, 1.6, method)
        assert np.array_equal(got, want), method
    pt = get_periodic_table()
    assert pt["c"] == 6 and pt["zn"] == 30 and pt["au"] == 79 and pt["rf"] == 104 and len(pt) == 104
