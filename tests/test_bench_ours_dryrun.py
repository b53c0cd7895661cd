"""bench.py's own arm, dry-run on CPU: the JSON assembly, roofline arithmetic and timing plumbing of run_ours()
executed against a stand-in engine (the CPU oracle behind the tomoengine method names, with fabricated kernel
timings) and a stubbed torch.cuda.  Nothing here measures anything -- it only proves that the code the driver
runs at round end cannot die on a typo, and that the line carries every key of the contract.  The product
itself has no CPU path (bench.py exits without a CUDA device; see test_no_gpu_no_fallback)."""
import json
import os
import subprocess
import sys
import time
import types

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


class _StandInEngine:
    PROFILE = {"fp": 6.0, "bp": 5.0, "fgp_first": 0.4, "fgp_sweep": 0.6, "fgp_final": 0.4, "tv": 0.2, "momentum": 0.3}

    def __init__(self, nslice, nray, angles, device=-1, x_offset=0, nslice_global=None):
        import oracle as orc
        self.o = orc.Oracle(nslice, nray, np.asarray(angles, np.float32))
        self.n, self.nx, self.na = nray, nslice, len(angles)
        self.launches, self.calls, self.t0 = 0, {}, 0.0

    def __getattr__(self, name):                      # algorithm calls go to the oracle, counted like kernel launches
        f = getattr(self.o, name)

        def call(*a):
            self.launches += 1
            self.calls[name] = self.calls.get(name, 0) + 1
            return f(*a)
        return call

    def set_original_volume_all(self, v): self.o.original = np.array(v, np.float32)
    def create_projections(self): self.o.create_projections()
    def get_projections(self): return self.o.get_projections()
    def get_recon_all_into(self, out): out[:] = self.o.recon
    def sizes(self): return self.nx * self.n * self.n * 4, self.nx * self.na * self.n * 4, self.nx * self.na * self.n * self.n
    def profile(self, on=True): self.calls = {}
    def launch_count(self): return self.launches
    def timer_start(self): self.t0 = time.perf_counter()
    def timer_stop(self): return (time.perf_counter() - self.t0) * 1e3
    def sync(self): pass

    def profile_report(self):
        it = max(self.calls.get("SIRT", 1), 1)
        cnt = {"fp": it, "bp": it, "fgp_first": it, "fgp_sweep": 8 * it, "fgp_final": it, "tv": it, "momentum": it}
        return {g: (ms * cnt[g], cnt[g]) for g, ms in self.PROFILE.items()}


def test_run_ours_dry_run_emits_the_full_contract(monkeypatch, capsys):
    import torch
    sys.path.insert(0, ROOT)
    import bench
    import tomo_tv_b200 as ttv
    monkeypatch.setattr(ttv, "tomoengine", _StandInEngine)
    monkeypatch.setattr(torch.cuda, "is_available", lambda: True)
    monkeypatch.setattr(torch.cuda, "set_device", lambda d: None)
    monkeypatch.setattr(torch.cuda, "synchronize", lambda *a: None)
    monkeypatch.setattr(torch.cuda, "get_device_properties", lambda d: types.SimpleNamespace(multi_processor_count=148))
    real_empty = torch.empty
    monkeypatch.setattr(torch, "empty", lambda *a, **k: real_empty(*a, **{x: y for x, y in k.items() if x != "pin_memory"}))
    for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK"):
        monkeypatch.delenv(k, raising=False)
    args = types.SimpleNamespace(gpus=1, steps=2, warmup=1, impl="ours", size=24, angles=6, no_cpu=True, phantom_period=0, max_seconds=60.0)
    bench.run_ours(args)
    lines = [l for l in capsys.readouterr().out.splitlines() if l.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    for key in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling", "vs_baseline",
                "dtype", "data", "config", "e2e", "gpu_launches", "clocks", "roofline", "kernels"):
        assert key in d, key
    assert d["metric"] == "fista_tv_iters_per_sec" and d["n_gpus"] == 1 and d["steps"] == 2 and d["warmup"] >= 3
    assert d["dtype"] == "f32" and d["vs_baseline"] is None and "workload" in d["config"] and "model" not in d["config"]
    assert set(d["e2e"]) == {"value", "unit", "h2d_bytes_per_step", "d2h_bytes_per_step"} and d["e2e"]["h2d_bytes_per_step"] > 0
    r = d["roofline"]
    assert r["bound"] == "hbm" and r["unit"] == "GB/s" and abs(r["frac"] - r["achieved"] / r["peak"]) < 1e-12
    V = 24 * 24 * 24 * 4
    assert r["kernel"] == "fgp_tile_kernel<SWEEP>" and r["algorithmic_bytes_per_launch"] == 7 * V
    assert abs(r["achieved"] - 7 * V / 0.6e-3 / 1e9) < 1e-6          # algorithmic bytes / fabricated 0.6 ms
    assert d["gpu_launches"] > 0
    # the phantom-period path used for 2048^3 runs builds the same shapes
    args.phantom_period = 8
    bench.run_ours(args)
    assert json.loads([l for l in capsys.readouterr().out.splitlines() if l.startswith("{")][0])["value"] > 0


def test_no_gpu_no_fallback():
    p = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--steps", "1", "--warmup", "0", "--size", "24", "--angles", "6"],
                       capture_output=True, text=True, timeout=300, cwd=ROOT)
    assert p.returncode != 0 and "no CUDA device" in (p.stderr + p.stdout)
