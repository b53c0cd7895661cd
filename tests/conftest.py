import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.dirname(os.path.abspath(__file__))):
    if p not in sys.path:
        sys.path.insert(0, p)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def pytest_collection_modifyitems(config, items):
    # GPU tests never run silently on a CPU fallback: without a device they are skipped here,
    # and on a GPU box a missing CUDA library is a hard failure (see cuda_lib fixture).
    try:
        import torch
        has_gpu = torch.cuda.is_available()
    except Exception:
        has_gpu = False
    if has_gpu:
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for it in items:
        if "gpu" in it.keywords:
            it.add_marker(skip)


@pytest.fixture(scope="session")
def orc():
    import oracle
    oracle.build()
    return oracle


@pytest.fixture(scope="session")
def ttv():
    """The product package with its CUDA library loaded; fails loudly if it is missing."""
    import tomo_tv_b200
    from tomo_tv_b200 import _native
    _native.lib()
    return tomo_tv_b200


def golden_cases():
    if not os.path.isdir(GOLDEN):
        return []
    out = []
    for f in sorted(os.listdir(GOLDEN)):
        if f.endswith(".npz") and not f.endswith("_inputs.npz") and not f.startswith(("fusion_", "cpuref_", "spread_")):
            geom, scn = f[:-4].split("_", 1)
            out.append((geom, scn))
    return out


def load_golden(geom, scn):
    inp = dict(np.load(os.path.join(GOLDEN, f"{geom}_inputs.npz")))
    want = dict(np.load(os.path.join(GOLDEN, f"{geom}_{scn}.npz")))
    if "vol__idx" in inp:                      # digest geometry: inputs are rebuilt from their seeds
        inp = _regenerated_inputs(geom, inp)
    return inp, want


_REGEN = {}


def _regenerated_inputs(geom, stored):
    """Digest geometries store only the angles and digests of their inputs: the seeded inputs are rebuilt
    (scenarios.make_inputs + the oracle's forward projection for `b`, as oracle/make_golden.py did) and must
    reproduce the stored digests bit for bit."""
    if geom in _REGEN:
        return dict(_REGEN[geom])
    import oracle
    import scenarios as S
    oracle.build()
    nx, n, _ = S.GEOMS[geom]
    inp = S.make_inputs(geom)
    o0 = oracle.Oracle(nx, n, stored["angles"])
    o0.recon[:] = inp["vol"]
    o0.forward_projection()
    inp["b"] = o0.g.copy()
    for k, v in S.digest(geom, inp).items():
        # slices bit for bit; the float64 norm / sum only to summation-order accuracy (numpy's reduction order is per machine)
        same = k in stored and (np.allclose(v, stored[k], rtol=1e-12, atol=0) if k.endswith(("__norm", "__sum")) else np.array_equal(np.asarray(v), stored[k]))
        assert same, f"regenerated input {geom}/{k} differs from the golden generation"
    for k in ("angles", "cosv", "sinv"):
        inp[k] = stored[k]
    _REGEN[geom] = inp
    return dict(inp)
