"""ttv_tv_value_noalloc (include/tomo_tv_b200.h): the allocation-free TV value against the engine's ttv_tv.

Collected last on purpose.  The entry point keeps its reduction scratch in a static
__device__ array that all calls share and that keeps the last call's contents, so the test calls it twice in a row -- the
second time on a SMALLER, different volume, with the first call's partial sums still in the array.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

RTOL = 1e-5     # both sides sum float32 terms in float64 and round once to float32 (the suite's default tolerance is 1e-4)


def _engine_tv(vol):
    from tomo_tv_b200.engine import tomoengine
    nx, n = vol.shape[0], vol.shape[1]
    e = tomoengine(nx, n, np.linspace(0, np.pi, 4, endpoint=False).astype(np.float32))
    e.initialize_initial_volume()
    e.set_recon_all(vol)
    return float(e.tv())


def _volume(nx, n, seed):
    rng = np.random.default_rng(seed)
    v = rng.random((nx, n, n), dtype=np.float32)
    v[:, n // 4: n // 2, n // 3:] += 2.0          # a step, so the value is not only noise
    return v


def test_tv_value_noalloc_matches_engine_on_two_calls_in_a_row():
    import torch
    from tomo_tv_b200.noalloc import tv_value_noalloc
    torch.cuda.set_device(0)
    out = torch.full((1,), -1.0, dtype=torch.float32, device="cuda:0")
    got, want = [], []
    for (nx, n, seed) in ((70, 136, 1), (37, 45, 2)):       # second call: fewer CTAs than the first left partials for
        vol = _volume(nx, n, seed)
        d = torch.from_numpy(vol).to("cuda:0")
        torch.cuda.synchronize()
        before = torch.cuda.memory_allocated()
        tv_value_noalloc(d.data_ptr(), nx, n, n, out.data_ptr())
        torch.cuda.synchronize()
        assert torch.cuda.memory_allocated() == before
        got.append(float(out.item()))
        want.append(_engine_tv(vol))
        del d
    for g, w in zip(got, want):
        assert w > 0 and abs(g - w) <= RTOL * abs(w), (got, want)
    # same input again: fixed summation order -> the same bits
    vol = _volume(37, 45, 2)
    d = torch.from_numpy(vol).to("cuda:0")
    tv_value_noalloc(d.data_ptr(), 37, 45, 45, out.data_ptr())
    torch.cuda.synchronize()
    assert float(out.item()) == got[1]


def test_tv_value_noalloc_non_cubic_against_numpy():
    # ny != nz (the engine needs ny == nz): float64 restatement of tv_gd.cu:27-47 on the spot
    import torch
    from tomo_tv_b200.noalloc import tv_value_noalloc
    rng = np.random.default_rng(3)
    v = rng.random((5, 33, 70), dtype=np.float32)
    d = torch.from_numpy(v).to("cuda:0")
    out = torch.zeros(1, dtype=torch.float32, device="cuda:0")
    tv_value_noalloc(d.data_ptr(), 5, 33, 70, out.data_ptr())
    torch.cuda.synchronize()
    w = v.astype(np.float64)
    ref = np.sqrt(1e-6 + (w - np.roll(w, -1, 0)) ** 2 + (w - np.roll(w, -1, 1)) ** 2 + (w - np.roll(w, -1, 2)) ** 2).sum()
    assert abs(float(out.item()) - ref) <= RTOL * ref


def test_tv_value_noalloc_rejects_oversize_and_host_pointers():
    import torch
    from tomo_tv_b200.noalloc import tv_value_noalloc, MAX_ROWS
    from tomo_tv_b200._native import NativeError
    d = torch.zeros(16, dtype=torch.float32, device="cuda:0")
    out = torch.zeros(1, dtype=torch.float32, device="cuda:0")
    with pytest.raises(NativeError, match="maximum"):
        tv_value_noalloc(d.data_ptr(), 1025, 1024, 4, out.data_ptr())       # 1025 * 1024 rows > MAX_ROWS; rejected before any launch
    assert 1025 * 1024 > MAX_ROWS
    h = np.zeros(16, np.float32)
    with pytest.raises(NativeError, match="device pointers"):
        tv_value_noalloc(h.ctypes.data, 1, 4, 4, out.data_ptr())
    # the entry point still works after the rejected calls
    tv_value_noalloc(d.data_ptr(), 1, 4, 4, out.data_ptr())
    torch.cuda.synchronize()
    assert abs(float(out.item()) - 16 * 1e-3) < 1e-6
