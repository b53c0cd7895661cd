"""Host-side logic that needs no GPU: algorithm dispatch, tilt-series packing, slab partitioning."""
import numpy as np
import pytest


class Spy:
    def __init__(self):
        self.calls = []

    def __getattr__(self, name):
        def f(*a):
            self.calls.append((name,) + a)
            return 1.25
        return f


def test_dispatch_matches_reference_table(ttv):
    """tomofusion/pytvlib.py:5-31 -- note 'fista' runs one SIRT iteration."""
    P = ttv.pytvlib
    cases = {
        "SIRT": ["initialize_SIRT"], "fista": ["initialize_fista"], "kl-divergence": ["initialize_poisson_ML"],
        "CGLS": ["initialize_CGLS"],
    }
    for alg, first in cases.items():
        s = Spy(); P.initialize_algorithm(s, alg)
        assert [c[0] for c in s.calls] == first + ["initialize_FP"]
    s = Spy(); P.initialize_algorithm(s, "SART", "random")
    assert s.calls == [("initialize_SART", "random"), ("initialize_FP",)]
    s = Spy(); P.initialize_algorithm(s, "asd-pocs", "sequential")
    assert s.calls[0] == ("initialize_SART", "sequential")
    s = Spy(); P.initialize_algorithm(s, "FBP", "hamming")
    assert s.calls[0] == ("initialize_FBP", "hamming")
    s = Spy(); P.run(s, "fista"); assert s.calls == [("SIRT", 1)]
    s = Spy(); P.run(s, "SIRT"); assert s.calls == [("SIRT", 1)]
    s = Spy(); P.run(s, "sart", 0.3); assert s.calls == [("SART", 0.3, 1)]
    s = Spy(); assert P.run(s, "kl-divergence", 0.2) == 1.25 and s.calls == [("poisson_ML", 0.2)]
    s = Spy(); P.run(s, "FBP"); assert s.calls == [("FBP", True)]
    s = Spy(); P.run(s, "CGLS"); assert s.calls == [("CGLS", 1)]
    assert len(P.wbp_filters()) == 12 and P.sart_orders() == ["sequential", "random"]


def test_pack_tilt_series_matches_reference_loop(ttv):
    from tomo_tv_b200.reconstructor import pack_tilt_series
    rng = np.random.default_rng(0)
    ts = rng.random((5, 7, 3))
    b = np.zeros((5, 7 * 3))
    for s in range(5):                       # the reference's per-slice loop, reconstructor.py:54-56
        b[s, :] = ts[s, :, :].transpose().ravel()
    got = pack_tilt_series(ts)
    assert got.dtype == np.float32 and np.allclose(got, b.astype(np.float32))
    with pytest.raises(ValueError):
        pack_tilt_series(np.zeros((3, 3)))


def test_slab_bounds_cover_the_volume(ttv):
    from tomo_tv_b200.phantom import slab_bounds
    for nslice in (1, 7, 8, 100, 512, 2048):
        for nr in (1, 2, 3, 4, 8):
            spans = [slab_bounds(nslice, nr, r) for r in range(nr)]
            assert spans[0][0] == 0 and sum(c for _, c in spans) == nslice
            for (f0, c0), (f1, _) in zip(spans, spans[1:]):
                assert f0 + c0 == f1
            assert max(c for _, c in spans) - min(c for _, c in spans) <= 1


def test_phantom_is_deterministic(ttv):
    from tomo_tv_b200.phantom import phantom3d
    a = phantom3d(6, 32); b = phantom3d(6, 32)
    assert a.dtype == np.float32 and a.shape == (6, 32, 32) and np.array_equal(a, b)
    assert a.min() >= 0 and 0.2 < float(a.max()) <= 1.0
    assert np.array_equal(phantom3d(6, 32, x0=2, count=3), a[2:5])


def test_tomogpu_get_projections_inverts_the_packing(ttv):
    """TomoGPU.get_projections (a `pass` stub in the reference, reconstructor.py:217-219) returns the tilt
    series in the constructor's [Nslice, Nray, Nangles] layout."""
    from tomo_tv_b200.reconstructor import pack_tilt_series, TomoGPU
    ts = np.random.default_rng(0).random((4, 6, 3)).astype(np.float32)

    class FakeEngine:
        def get_projections(self):
            return pack_tilt_series(ts)

    t = TomoGPU.__new__(TomoGPU)
    t.tomo, t.Nangles, t.Nray = FakeEngine(), 3, 6
    assert np.array_equal(t.get_projections(), ts)
    with pytest.raises(RuntimeError):
        t.show_recon()


def test_parity_register_bookkeeping_of_the_table_driven_projectors():
    """Integer bookkeeping of ttv_bpt.cuh / ttv_fpt.cuh restated in numpy: of the cell pair (c, c+1) the even cell
    is (c+1) & ~1 and the odd one c | 1, the weight of the lower cell goes to whichever register holds it, and a run
    that keeps the cells it already holds loads exactly the missing ones."""
    c = np.arange(0, 300)
    cE, cO = (c + 1) & ~1, c | 1
    assert np.array_equal(np.sort(np.stack([cE, cO]), axis=0), np.stack([c, c + 1]))
    assert np.all(cE % 2 == 0) and np.all(cO % 2 == 1)
    odd = (c & 1) != 0
    w_lo, w_hi = 0.25, 0.75
    wE, wO = np.where(odd, w_hi, w_lo), np.where(odd, w_lo, w_hi)
    assert np.all(np.where(cE == c, wE, wO) == w_lo) and np.all(np.where(cE == c + 1, wE, wO) == w_hi)
    # fpt entry: [even index | odd index << 8 | weight of the even cell << 16], 0xFF = held; weights 0..256 in 1/256
    a8 = np.arange(256)
    for cc in (6, 7):
        wE8 = a8 if cc & 1 else 256 - a8
        e = ((cc + 1) & ~1) | ((cc | 1) << 8) | (wE8 << 16)
        assert np.array_equal(e & 0xFF, np.full(256, (cc + 1) & ~1)) and np.array_equal((e >> 8) & 0xFF, np.full(256, cc | 1))
        dec = (((e >> 16) & 0xFFFF) | 0x4B000000).astype(np.uint32).view(np.float32) - np.float32(8388608.0)
        assert np.array_equal(dec, wE8.astype(np.float32)) and np.all(dec + (256 - dec) == 256)
    # walking a run of cells with steps of 0, +1, -1 or +2 cells reloads exactly the cells not already held
    rng = np.random.default_rng(0)
    cells = np.cumsum(rng.choice([0, 1, 1, 2, -1], size=200)) + 50
    hE = hO = -1
    loads = 0
    for cc in cells:
        e, o = (cc + 1) & ~1, cc | 1
        loads += (e != hE) + (o != hO)
        hE, hO = e, o
        assert {hE, hO} == {cc, cc + 1}
    assert loads < 2 * len(cells)


def test_result_files_follow_the_reference_tree(ttv, tmp_path, monkeypatch):
    """pytvlib.save_results / save_recon (tomofusion/pytvlib.py:120-162): /parameters attributes, /results datasets,
    /Reconstruction/recon + size attributes -- HDF5 with h5py, the same tree as .npz keys without it."""
    monkeypatch.chdir(tmp_path)
    P = ttv.pytvlib

    class T:
        def get_recon(self, s):
            return np.full((4, 4), float(s), np.float32)

    path = P.save_results(("run1", "fista"), {"lambda": 0.1, "Niter": 5}, {"cost": np.arange(5.0), "tv": [1, 2, 3]})
    assert P.save_recon(("run1", "fista"), (3, 4, 7), T()) == path
    if path.endswith(".npz"):
        d = np.load(path)
        assert float(d["parameters@lambda"]) == 0.1 and int(d["parameters@Niter"]) == 5
        assert d["results/cost"].dtype == np.float32 and np.array_equal(d["results/tv"], [1, 2, 3])
        assert d["Reconstruction/recon"].shape == (3, 4, 4) and d["Reconstruction/recon"][2, 0, 0] == 2.0
        assert int(d["Reconstruction@Nproj"]) == 7
    else:
        import h5py
        with h5py.File(path, "r") as h5:
            assert h5["parameters"].attrs["lambda"] == 0.1 and h5["results/cost"].dtype == np.float32
            assert h5["Reconstruction/recon"].shape == (3, 4, 4) and h5["Reconstruction"].attrs["Nproj"] == 7


def test_lane_mappings_of_the_table_driven_projectors_cover_their_tiles_once():
    """Index bookkeeping of ttv_bpt.cuh / ttv_fpt.cuh restated in numpy: every (row, column, slice) of a BP tile and
    every (angle slot, detector, slice) of an FP CTA is owned by exactly one (warp, lane, t, half), every quarter warp
    reads one full 128-byte line per LDS.128, and helper threads and consumers agree on the table addresses."""
    # --- bpt_kernel: 16 warps = 16 runs of 8 voxels in a 16 (X) x 8 (Y) tile; 8 x-blocks of 32 slices per CTA
    TX, TY, NW, VT, XB = 16, 8, 16, 8, 8
    seen = np.zeros((TY, TX, XB * 32), int)
    for warp in range(NW):
        Y, Xb = warp >> 1, (warp & 1) * VT
        for lane in range(32):
            xa, q = (lane >> 3) * 2, lane & 7
            for t in range(VT):
                for xb in (xa, xa + 1):
                    seen[Y, Xb + t, xb * 32 + q * 4: xb * 32 + q * 4 + 4] += 1
    assert np.all(seen == 1)
    # helper warp: entry v = i*32 + lane belongs to run v >> 3, voxel v & 7; consumers read run*8 + t
    assert sorted(i * 32 + lane for i in range(4) for lane in range(32)) == [r * 8 + t for r in range(16) for t in range(8)]
    for quarter in range(4):                      # LDS.128 of a quarter warp: 8 lanes x 16 B of one x-block = 32 distinct banks
        banks = [(((lane >> 3) * 256 + (lane & 7) * 16) // 4 + i) % 32 for lane in range(quarter * 8, quarter * 8 + 8) for i in range(4)]
        assert len(set(banks)) == 32
    # --- fpt_kernel: 16 warps x 4 quarter warps = 64 runs = 8 angles x 8 detector runs of 8; 2 x-blocks per CTA
    seen = np.zeros((8, 64, 64), int)
    for warp in range(16):
        for lane in range(32):
            run, lx = warp * 4 + (lane >> 3), lane & 7
            ai, ds = run >> 3, run & 7
            assert ai == warp // 2                                # warp-uniform angle
            for t in range(8):
                for xb in (0, 1):
                    seen[ai, ds * 8 + t, xb * 32 + lx * 4: xb * 32 + lx * 4 + 4] += 1
    assert np.all(seen == 1)
    # helper thread h = run*2 + half writes 16 bytes at h*16; the consumer of `run` reads 32 bytes at run*32
    assert [h * 16 for h in range(128)] == [run * 32 + half * 16 for run in range(64) for half in range(2)]


@pytest.mark.parametrize("tx,ty,cw", [(16, 16, 27), (16, 8, 22), (32, 16, 40)])
def test_bp_detector_windows_hold_every_sample_of_the_tile(tx, ty, cw):
    """Window sizing of bp_kernel (16 x 16 tile, BP_CW = 27) and bpt_kernel (16 x 8: CW = 22; 32 x 16: CW = 40),
    restated from the producer code: w0 = floor(fmin - 0.5) - 1 with fmin the smallest detector coordinate of the
    tile's corners; every voxel's cell pair (c, c+1) relative to w0 must lie inside [0, cw)."""
    rng = np.random.default_rng(tx)
    lo, hi = 10 ** 9, -10 ** 9
    for n in (512, 2048, 200):
        for _ in range(1500):
            th = rng.uniform(0, 2 * np.pi)
            c, s, off = np.float32(np.cos(th)), np.float32(np.sin(th)), np.float32(n / 2)
            x0 = int(rng.integers(0, (n + tx - 1) // tx)) * tx
            y0 = int(rng.integers(0, (n + ty - 1) // ty)) * ty
            fx0, fy0 = np.float32(x0 - 0.5 * n + 0.5), np.float32(y0 - 0.5 * n + 0.5)
            ax0, ax1 = fx0 * c, np.float32(fx0 + np.float32(tx - 1)) * c
            ay0, ay1 = -fy0 * s, -np.float32(fy0 + np.float32(ty - 1)) * s
            fmin = np.float32(np.float32(min(ax0, ax1) + min(ay0, ay1)) + off)
            w0 = (int(np.floor(np.float32(fmin - np.float32(0.5)))) - 1) & ~1      # even start
            fx = (np.arange(x0, x0 + tx) - 0.5 * n + 0.5)[None, :]
            fy = (np.arange(y0, y0 + ty) - 0.5 * n + 0.5)[:, None]
            cell = np.floor(fx * float(c) - fy * float(s) + float(off) - 0.5).astype(int) - w0
            lo, hi = min(lo, int(cell.min())), max(hi, int(cell.max()))
    assert lo >= 0 and hi + 1 <= cw - 2          # at least one spare cell for float32 rounding at either end


def test_tomofusion_import_paths_resolve_to_this_package(ttv):
    """The reference's import spellings (tomofusion/gpu/reconstructor.py:1-3, chemistry/reconstructor.py:1-2,
    gpu/utils/__init__.py:2, gpu/simulator.py:1-2) work against the `tomofusion/` shim and name this package's classes."""
    import tomofusion
    from tomofusion.gpu.utils import tomoengine, multigpuengine
    from tomofusion.gpu.reconstructor import TomoGPU, reconstructor
    from tomofusion import pytvlib
    from tomofusion.chemistry.utils import multimodal, multigpufusion
    from tomofusion.chemistry.utils import fusion_helper as utils
    from tomofusion.chemistry.reconstructor import ChemicalTomo
    assert tomoengine.tomoengine is ttv.tomoengine and multigpuengine.multigpuengine is ttv.multigpuengine
    assert multimodal is ttv.multimodal and multigpufusion.multigpufusion is ttv.multigpufusion
    assert TomoGPU is ttv.TomoGPU and reconstructor is ttv.TomoGPU and ChemicalTomo is ttv.ChemicalTomo
    assert pytvlib.wbp_filters() == ttv.pytvlib.wbp_filters()
    assert utils.get_periodic_table()["zn"] == 30
    assert isinstance(tomofusion.device_count(), list)
    assert tomofusion.determine_gpu_config is ttv.determine_gpu_config
    # the modules are callable like the classes they hold (both spellings are seen in user scripts)
    assert callable(tomoengine) and callable(multigpuengine)


def test_digest_goldens_compare_slices_and_norms():
    """scenarios.digest / compare (fixtures of the 70 x 136^2 geometry hold six slices + whole-array norms): a full result
    matches its own digest, a change inside a stored slice or anywhere else in the volume is caught, and the ASD-POCS volume
    tolerance comes from the reference's measured spread."""
    import scenarios as S
    rng = np.random.default_rng(0)
    nx = S.GEOMS["C"][0]
    full = {"recon": rng.random((nx, 8, 8), dtype=np.float32), "cost": np.float32(3.5)}
    dg = S.digest("C", full)
    assert dg["recon"].shape[0] == len(S.DIGEST_SLICES["C"]) and "recon__norm" in dg and dg["cost"] == np.float32(3.5)
    assert all(r[3] for r in S.compare("fista", full, dg, "C"))
    bad = {"recon": full["recon"].copy(), "cost": full["cost"]}
    bad["recon"][S.DIGEST_SLICES["C"][1], 2, 3] += 0.5                  # inside a stored slice
    assert not all(r[3] for r in S.compare("fista", bad, dg, "C"))
    bad2 = {"recon": full["recon"].copy(), "cost": full["cost"]}
    bad2["recon"][5] += 0.5                                              # a slice that is NOT stored: the norm catches it
    assert not all(r[3] for r in S.compare("fista", bad2, dg, "C"))
    assert S.digest("A", full) is full                                   # small geometries keep whole arrays
    assert 1e-3 < S.tol_for("C", "asd", "recon") < 1e-2 and S.tol_for("C", "asd", "dd") == 1e-4 and S.tol_for("C", "fista", "recon") == 1e-4
