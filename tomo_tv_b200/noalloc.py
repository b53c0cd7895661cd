"""Allocation-free entry points of the C ABI (include/tomo_tv_b200.h: ttv_tv_value_noalloc).

No counterpart in the reference; for callers that may not call cudaMalloc.  Arguments are raw DEVICE addresses (e.g.
`tensor.data_ptr()`), nothing is copied or allocated on either side; see the header for the limit and the sharing rules.
"""
import ctypes as C

from ._native import lib, check

MAX_ROWS = 1048576   # TTV_NOALLOC_MAX_ROWS: nx * ny of one call


def tv_value_noalloc(d_vol, nx, ny, nz, d_tv, stream=0):
    """TV value of the float32 device array [nx][ny][nz] at address `d_vol`, written (stream-ordered, no synchronisation)
    to the float32 at device address `d_tv`.  `stream`: a cudaStream_t as an integer, 0 = default stream."""
    f = lib().ttv_tv_value_noalloc
    f.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
    f.restype = C.c_int
    check(f(C.c_void_p(int(d_vol)), int(nx), int(ny), int(nz), C.c_void_p(int(d_tv)), C.c_void_p(int(stream))))
