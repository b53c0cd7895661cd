"""CPU-side checks of the drop-in boundary: the C-ABI library loads without a GPU, exports every
symbol include/tomo_tv_b200.h declares, and refuses to compute without a device (no CPU fallback)."""
import ctypes as C
import os

import numpy as np
import pytest

from conftest import ROOT


def test_library_exports_every_declared_symbol(ttv):
    from tomo_tv_b200 import _native
    names = _native.exported_symbols()
    assert len(names) >= 55
    L = _native.lib()
    missing = [n for n in names if not hasattr(L, n)]
    assert not missing, missing


def test_header_cites_reference_for_every_entry_point():
    txt = open(os.path.join(ROOT, "include", "tomo_tv_b200.h")).read()
    assert "tomoengine.cpp:487-534" in txt
    assert txt.count("/* :") + txt.count("tomoengine.cpp") > 30


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "tomo_tv_b200")
    for dp, _, fs in os.walk(pkg):
        for f in fs:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dp, f)).read()
                assert "import oracle" not in src and "from oracle" not in src and "liboracle" not in src, f


def _no_gpu():
    try:
        import torch
        return not torch.cuda.is_available()
    except Exception:
        return True


@pytest.mark.skipif(not _no_gpu(), reason="only meaningful on a box without a GPU")
def test_no_cpu_fallback(ttv):
    assert ttv.device_count() == 0
    with pytest.raises(ValueError):
        ttv.determine_gpu_config()
    with pytest.raises(ttv.NativeError):
        ttv.tomoengine(4, 32, np.zeros(3, np.float32))
    with pytest.raises(SystemExit):
        ttv.pytvlib.check_cuda()


def test_noalloc_limit_in_header_matches_python_wrapper():
    """TTV_NOALLOC_MAX_ROWS (include/tomo_tv_b200.h) is the limit tomo_tv_b200.noalloc documents; the entry point is declared."""
    import os
    import re
    from tomo_tv_b200 import noalloc
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    txt = open(os.path.join(root, "include", "tomo_tv_b200.h")).read()
    m = re.search(r"#define\s+TTV_NOALLOC_MAX_ROWS\s+(\d+)", txt)
    assert m and int(m.group(1)) == noalloc.MAX_ROWS
    assert "int ttv_tv_value_noalloc(const float* d_vol, int nx, int ny, int nz, float* d_tv, void* stream);" in txt
