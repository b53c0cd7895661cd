"""CPU tests that pin the oracle (oracle/oracle.c) against independent known answers:

* ASTRA's analytic rectangle model for the Joseph ("linear"/"cuda") projector
  (thirdparty/astra-toolbox/tests/python/functional/test_line2d.py:476-510; tolerance 2e-2 with
  8-bit interpolation weights, 2e-3 with exact weights) -- restated here in closed form from the
  geometry, independently of oracle.c.
* ASTRA's adjoint tolerance for the CUDA pair (test_line2d.py:591-640: < 1e-1).
* ASTRA's reconstruction-scaling test (test_rec_scaling.py:134-145: a block of ones reconstructs
  to 1 +- 5e-2 with SIRT).
* numpy restatements of the tomofusion TV kernels (tv_gd.cu / tv_fgp.cu), including the TV-GD index
  quirks evaluated with C-style truncating modulo in pure Python.
"""
import math

import numpy as np
import pytest

import scenarios as S
from conftest import golden_cases, load_golden


# ------------------------------------------------------------------------------------------
# analytic Joseph sums of an axis-aligned block of ones
# ------------------------------------------------------------------------------------------
def _tri_sum(p, lo, hi):
    """sum_{c=lo}^{hi-1} max(0, 1-|p-c|): linear-interpolation weight of a run of ones."""
    c = np.arange(lo, hi)
    return np.maximum(0.0, 1.0 - np.abs(p - c)).sum()


def analytic_sino(n, angles, r0, r1, c0, c1, round8):
    """Volume = ones on rows [r0,r1) x cols [c0,c1).  Geometry of SURVEY.md App. A.1."""
    out = np.zeros((len(angles), n))
    for a, al in enumerate(np.asarray(angles, np.float64)):
        ca, sa = math.cos(al), math.sin(al)
        vertical = abs(-sa) <= abs(ca)
        for d in range(n):
            t = d - n / 2 + 0.5
            tot = 0.0
            if vertical:
                for j in range(r0, r1):
                    y = n / 2 - j - 0.5
                    x = t / ca - y * (sa / ca)
                    w = _tri_sum(x + n / 2 - 0.5, c0, c1)
                    if round8:
                        w = round(w * 256.0) / 256.0
                    tot += w
                out[a, d] = tot / abs(ca)
            else:
                for k in range(c0, c1):
                    x = k - n / 2 + 0.5
                    y = t / sa - x * (ca / sa)
                    w = _tri_sum(n / 2 - 0.5 - y, r0, r1)
                    if round8:
                        w = round(w * 256.0) / 256.0
                    tot += w
                out[a, d] = tot / abs(sa)
    return out


@pytest.mark.parametrize("quant,tol", [(1, 2e-2), (0, 2e-3)])
def test_fp_matches_astra_analytic_rectangles(orc, quant, tol):
    rng = np.random.RandomState(123)
    worst = 0.0
    for _ in range(12):
        n = int(rng.randint(24, 72))
        na = int(rng.randint(4, 9))
        angles = (rng.random_sample(na) * 2 * np.pi - np.pi).astype(np.float32)
        r0, c0 = rng.randint(0, n - 4, 2)
        r1 = int(rng.randint(r0 + 1, n + 1)); c1 = int(rng.randint(c0 + 1, n + 1))
        vol = np.zeros((n, n), np.float32)
        vol[r0:r1, c0:c1] = 1.0
        geo = orc.Geometry(angles, n)
        got = orc.fp2d(vol, geo, quant=quant)
        want = analytic_sino(n, angles.astype(np.float64), int(r0), r1, int(c0), c1, round8=bool(quant))
        worst = max(worst, float(np.abs(got - want).max()))
    assert worst <= tol, worst


def test_fp_bp_adjoint_mismatch_is_astra_like(orc):
    """ASTRA's CUDA pair is unmatched (ray-driven FP, voxel-driven BP); its own test accepts
    |<Ax,y> - <x,A^T y>| / <Ax,y> < 1e-1 (test_line2d.py:633)."""
    rng = np.random.RandomState(5)
    for n, na in ((48, 16), (65, 30)):
        geo = orc.Geometry(np.linspace(-np.pi / 2, np.pi / 2, na, endpoint=False).astype(np.float32), n)
        x = rng.random_sample((n, n)).astype(np.float32)
        y = rng.random_sample((na, n)).astype(np.float32)
        ax = orc.fp2d(x, geo)
        aty = orc.bp2d(y, geo, (n, n))
        da, db = float(np.dot(ax.ravel().astype(np.float64), y.ravel())), float(np.dot(x.ravel().astype(np.float64), aty.ravel()))
        assert abs(da - db) / da < 1e-1


def test_sirt_reconstructs_block_of_ones(orc):
    """test_rec_scaling.py:134-145 -- pins the R / C scale conventions."""
    n, na = 64, 48
    ang = np.linspace(0, np.pi, na, endpoint=False).astype(np.float32)
    o = orc.Oracle(1, n, ang)
    vol = np.zeros((1, n, n), np.float32)
    vol[0, 20:44, 22:42] = 1.0
    o.recon[:] = vol
    o.forward_projection()
    o.set_tilt_series(o.g)
    o.restart_recon()
    o.initialize_SIRT()
    o.SIRT(200)
    inside = o.recon[0, 24:40, 26:38]
    assert abs(float(inside.mean()) - 1.0) < 5e-2


# ------------------------------------------------------------------------------------------
# TV kernels
# ------------------------------------------------------------------------------------------
def np_tv(v):
    dx = v - np.roll(v, -1, 0); dy = v - np.roll(v, -1, 1); dz = v - np.roll(v, -1, 2)
    return np.sqrt(np.float32(1e-6) + dx * dx + dy * dy + dz * dz)


def np_fgp(vol, n_iter, lam):
    """tv_fgp.cu:44-115,192-281 with numpy slicing."""
    f = np.float32
    P = [np.zeros_like(vol) for _ in range(3)]
    D = np_tv(vol)
    m = f(1.0) / (f(26.0) * f(lam))
    for _ in range(n_iter):
        div = P[0] + P[1] + P[2]
        div[1:] -= P[0][:-1]; div[:, 1:] -= P[1][:, :-1]; div[:, :, 1:] -= P[2][:, :, :-1]
        D = np.maximum(vol - f(lam) * div, f(0))
        g = [np.zeros_like(vol) for _ in range(3)]
        g[0][:-1] = D[:-1] - D[1:]; g[1][:, :-1] = D[:, :-1] - D[:, 1:]; g[2][:, :, :-1] = D[:, :, :-1] - D[:, :, 1:]
        Q = [P[i] + m * g[i] for i in range(3)]
        den = Q[0] ** 2 + Q[1] ** 2 + Q[2] ** 2
        s = np.where(den > 1, f(1) / np.sqrt(np.maximum(den, f(1e-30))), f(1)).astype(f)
        P = [q * s for q in Q]
    return D


def c_mod(a, b):
    return int(math.fmod(a, b))          # C '%' truncates toward zero


def py_tv_gradient(r):
    """tv_gradient_3D_kernel (tv_gd.cu:52-106) restated literally: linear indices with the
    reference's expressions, C-truncating modulo, out-of-range reads := 0."""
    nx, ny, nz = r.shape
    flat = r.ravel()
    n = flat.size
    eps = np.float32(1e-6)
    g = np.zeros(n, np.float32)

    def rd(i):
        return flat[i] if 0 <= i < n else np.float32(0)

    for i in range(nx):
        for j in range(ny):
            for k in range(nz):
                P = ny * nz
                ijk = P * i + nz * j + k
                ip = P * ((i + 1) % nx) + nz * j + k
                im = P * ((i - 1 + nx) % nx) + nz * j + k
                jp = P * i + nz * ((j + 1) % ny) + k
                jm = P * i + nz * ((j - 1 + ny) % ny) + k
                kp = P * i + nz * j + ((k + 1) % nz)
                km = P * i + nz * j + ((k - 1 + nz) % nz)
                im_jp = P * ((i - 1 + nx) % nx) + nz * ((j + 1) % ny) + k
                ip_jm = P * ((i + nx) % nx) + nz * ((j - 1 + ny) % ny) + k
                jm_kp = P * i + nz * ((j - 1 + ny) % ny) + ((k + 1) % nz)
                jp_km = P * i + nz * ((j + 1) % ny) + c_mod(k - 1 - nz, nz)
                im_kp = P * ((i - 1 + nx) % nx) + nz * j + ((k + 1) % nz)
                ip_km = P * ((i + 1) % nx) + nz * j + c_mod(k - 1 - nz, nz)
                c = flat[ijk]
                f = np.float32
                v1n = f(3.0 * float(c) - float(flat[ip]) - float(flat[jp]) - float(flat[kp]))
                v1d = np.sqrt(eps + (c - flat[ip]) ** 2 + (c - flat[jp]) ** 2 + (c - flat[kp]) ** 2)
                v2n = c - flat[im]
                v2d = np.sqrt(eps + (flat[im] - c) ** 2 + (flat[im] - flat[im_jp]) ** 2 + (flat[im] - flat[im_kp]) ** 2)
                v3n = c - flat[jm]
                v3d = np.sqrt(eps + (flat[jm] - flat[ip_jm]) ** 2 + (flat[jm] - c) ** 2 + (flat[jm] - flat[jm_kp]) ** 2)
                v4n = c - flat[km]
                v4d = np.sqrt(eps + (flat[km] - rd(ip_km)) ** 2 + (flat[km] - rd(jp_km)) ** 2 + (flat[km] - c) ** 2)
                g[ijk] = v1n / v1d + v2n / v2d + v3n / v3d + v4n / v4d
    return g.reshape(r.shape)


def test_tv_value_and_fgp_against_numpy(orc):
    rng = np.random.default_rng(3)
    v = rng.random((7, 9, 8), dtype=np.float32)
    assert abs(orc.lib().orc_tv(v, 7, 9, 8) - float(np_tv(v).astype(np.float64).sum())) < 1e-3
    for it in (0, 1, 2, 7):
        w = v.copy()
        ret = orc.lib().orc_tv_fgp(w, it, 0.1, 7, 9, 8)
        assert abs(ret - np_tv(v).sum()) / np_tv(v).sum() < 1e-5
        want = np_fgp(v, it, 0.1)
        assert S.rel_l2(w, want) < 2e-6, it


def test_tv_gradient_quirks_against_literal_restatement(orc):
    rng = np.random.default_rng(4)
    for shp in ((4, 5, 6), (3, 3, 3), (2, 6, 4)):
        v = rng.random(shp, dtype=np.float32)
        g = np.zeros_like(v)
        orc.lib().orc_tv_gradient(v, g, *shp)
        assert S.rel_l2(g, py_tv_gradient(v)) < 2e-6, shp


def test_elementwise_and_momentum(orc):
    rng = np.random.default_rng(6)
    x = rng.standard_normal(1000).astype(np.float32)
    y = x.copy()
    orc.lib().orc_soft_threshold(y, 0.3, y.size)
    want = np.sign(x) * np.maximum(np.abs(x) - np.float32(0.3), 0)
    assert np.allclose(y, want, atol=1e-7)
    recon = np.zeros(100, np.float32); yk = rng.random(100, dtype=np.float32); old = rng.random(100, dtype=np.float32)
    yk0, old0 = yk.copy(), old.copy()
    orc.lib().orc_momentum(recon, yk, old, 0.4, 100)
    assert np.array_equal(recon, yk0) and np.array_equal(old, yk0)
    assert np.allclose(yk, yk0 + np.float32(0.4) * (yk0 - old0), atol=1e-7)


# ------------------------------------------------------------------------------------------
# oracle vs golden vectors produced by the REAL reference on a B200 (oracle/make_golden.py)
# ------------------------------------------------------------------------------------------
@pytest.mark.parametrize("geom,scn", golden_cases())
def test_oracle_matches_reference_goldens(orc, geom, scn):
    inp, want = load_golden(geom, scn)
    nx, n, _ = S.GEOMS[geom]
    eng = orc.Oracle(nx, n, inp["angles"], quant=1, cosv=inp["cosv"], sinv=inp["sinv"])
    got = S.SCENARIOS[scn](eng, inp, nx)
    bad = [r for r in S.compare(scn, got, want, geom) if not r[3]]
    assert not bad, bad
