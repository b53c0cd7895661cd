"""Host-side mirror of the reference's pybind11 class `tomoengine`
(/root/reference/tomofusion/gpu/utils/tomoengine.cpp:487-534): same method names, argument meaning
and array layouts, implemented as thin ctypes calls into the C ABI (include/tomo_tv_b200.h).

All compute runs in the CUDA library; this file only marshals numpy arrays.  Arrays cross the
boundary by copy, exactly like the reference's by-value Eigen arguments (tomoengine.hpp:38).
"""
import ctypes as C
import numpy as np

from ._native import lib, check, NativeError  # noqa: F401

_f32p = C.POINTER(C.c_float)


def _as_f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _p(a):
    return a.ctypes.data_as(_f32p)


def device_count():
    """Number of usable CUDA devices (replaces the pycuda probe in tomofusion/__init__.py:10-18)."""
    n = C.c_int(0)
    check(lib().ttv_device_count(C.byref(n)))
    return n.value


class tomoengine:
    """tomoengine(Nslice, Nray) / tomoengine(Nslice, Nray, angles[rad])  -- tomoengine.cpp:491-492.

    Extra keyword arguments (not in the reference): `device` (CUDA ordinal, default current),
    `x_offset` / `nslice_global` for a tilt-axis slab of a sharded volume.
    """

    def __init__(self, Nslice, Nray, angles=None, device=-1, x_offset=0, nslice_global=None):
        self._h = C.c_void_p()
        self.Nslice, self.Nray = int(Nslice), int(Nray)
        ang = None if angles is None else _as_f32(np.ravel(angles))
        self.Nproj = 0 if ang is None else int(ang.size)
        nsg = self.Nslice if nslice_global is None else int(nslice_global)
        check(lib().ttv_create_slab(C.byref(self._h), self.Nslice, self.Nray,
                                    None if ang is None else _p(ang), self.Nproj, int(device),
                                    int(x_offset), nsg))

    def __del__(self):
        h = getattr(self, "_h", None)
        if h is not None and h.value:
            try:
                lib().ttv_destroy(h)
            except Exception:
                pass
            self._h = C.c_void_p()

    # ---- data -------------------------------------------------------------------------------
    def set_tilt_series(self, b):
        b = _as_f32(b)
        if b.ndim != 2:
            raise ValueError("set_tilt_series expects a [Nslice, Nproj*Nray] matrix")
        check(lib().ttv_set_tilt_series(self._h, _p(b), b.shape[0], b.shape[1]))

    def get_projections(self):
        out = np.empty((self.Nslice, self.Nproj * self.Nray), np.float32)
        check(lib().ttv_get_projections(self._h, _p(out)))
        return out

    def get_model_projections(self):
        out = np.empty((self.Nslice, self.Nproj * self.Nray), np.float32)
        check(lib().ttv_get_model_projections(self._h, _p(out)))
        return out

    def initialize_initial_volume(self):
        check(lib().ttv_initialize_initial_volume(self._h))

    def initialize_recon_copy(self):
        check(lib().ttv_initialize_recon_copy(self._h))

    def set_original_volume(self, vol, s):
        v = _as_f32(vol)
        if v.shape != (self.Nray, self.Nray):
            raise ValueError("set_original_volume expects an [Nray, Nray] slice")
        check(lib().ttv_set_original_volume(self._h, _p(v), int(s)))

    def set_recon(self, vol, s):
        v = _as_f32(vol)
        if v.shape != (self.Nray, self.Nray):
            raise ValueError("set_recon expects an [Nray, Nray] slice")
        check(lib().ttv_set_recon(self._h, _p(v), int(s)))

    def get_recon(self, s):
        out = np.empty((self.Nray, self.Nray), np.float32)
        check(lib().ttv_get_recon(self._h, _p(out), int(s)))
        return out

    # bulk additions -- not in the reference, which loops slices in Python
    def get_recon_all(self):
        out = np.empty((self.Nslice, self.Nray, self.Nray), np.float32)
        check(lib().ttv_get_recon_all(self._h, _p(out)))
        return out

    def get_recon_all_into(self, out):
        """Bulk copy into a caller-provided C-contiguous float32 [Nslice, Nray, Nray] buffer
        (e.g. pinned memory)."""
        if out.dtype != np.float32 or not out.flags.c_contiguous or out.shape != (self.Nslice, self.Nray, self.Nray):
            raise ValueError("get_recon_all_into expects a C-contiguous float32 [Nslice, Nray, Nray] array")
        check(lib().ttv_get_recon_all(self._h, _p(out)))
        return out

    def set_recon_all(self, vol):
        v = _as_f32(vol)
        if v.shape != (self.Nslice, self.Nray, self.Nray):
            raise ValueError("set_recon_all expects [Nslice, Nray, Nray]")
        check(lib().ttv_set_recon_all(self._h, _p(v)))

    def set_original_volume_all(self, vol):
        v = _as_f32(vol)
        if v.shape != (self.Nslice, self.Nray, self.Nray):
            raise ValueError("set_original_volume_all expects [Nslice, Nray, Nray]")
        check(lib().ttv_set_original_volume_all(self._h, _p(v)))

    def update_projection_angles(self, angles):
        a = _as_f32(np.ravel(angles))
        check(lib().ttv_update_projection_angles(self._h, _p(a), int(a.size)))
        self.Nproj = int(a.size)

    def create_projections(self):
        check(lib().ttv_create_projections(self._h))

    def restart_recon(self):
        check(lib().ttv_restart_recon(self._h))

    def get_gpu_id(self):
        v = C.c_int(0)
        check(lib().ttv_get_gpu_id(self._h, C.byref(v)))
        return v.value

    def set_gpu(self, gpu_id):
        check(lib().ttv_set_gpu(self._h, int(gpu_id)))

    def poisson_noise(self, Nc):
        check(lib().ttv_poisson_noise(self._h, int(Nc)))

    # ---- algorithms ---------------------------------------------------------------------------
    def initialize_SART(self, order):
        check(lib().ttv_initialize_SART(self._h, str(order).encode()))

    def SART(self, beta, nIter):
        check(lib().ttv_SART(self._h, C.c_float(beta), int(nIter)))

    def initialize_SIRT(self):
        check(lib().ttv_initialize_SIRT(self._h))

    def SIRT(self, nIter):
        check(lib().ttv_SIRT(self._h, int(nIter)))

    def initialize_FBP(self, filter_name):
        check(lib().ttv_initialize_FBP(self._h, str(filter_name).encode()))

    def FBP(self, apply_positivity):
        check(lib().ttv_FBP(self._h, int(bool(apply_positivity))))

    def initialize_FP(self):
        check(lib().ttv_initialize_FP(self._h))

    def initialize_BP(self):
        check(lib().ttv_initialize_BP(self._h))

    def initialize_CGLS(self):
        check(lib().ttv_initialize_CGLS(self._h))

    def CGLS(self, nIter):
        check(lib().ttv_CGLS(self._h, int(nIter)))

    def initialize_poisson_ML(self):
        check(lib().ttv_initialize_poisson_ML(self._h))

    def poisson_ML(self, lam):
        v = C.c_float(0)
        check(lib().ttv_poisson_ML(self._h, C.c_float(lam), C.byref(v)))
        return v.value

    def forward_projection(self):
        check(lib().ttv_forward_projection(self._h))

    def soft_threshold(self, lam):
        check(lib().ttv_soft_threshold(self._h, C.c_float(lam)))

    def initialize_fista(self):
        check(lib().ttv_initialize_fista(self._h))

    def fista_momentum(self, beta):
        check(lib().ttv_fista_momentum(self._h, C.c_float(beta)))

    def remove_momentum(self):
        check(lib().ttv_remove_momentum(self._h))

    def copy_recon(self):
        check(lib().ttv_copy_recon(self._h))

    def _scalar(self, fn, *args):
        v = C.c_float(0)
        check(fn(self._h, *args, C.byref(v)))
        return v.value

    def get_lipschitz(self):
        return self._scalar(lib().ttv_get_lipschitz)

    def l1_norm(self):
        return self._scalar(lib().ttv_l1_norm)

    def matrix_2norm(self):
        return self._scalar(lib().ttv_matrix_2norm)

    def data_distance(self):
        return self._scalar(lib().ttv_data_distance)

    def rmse(self):
        return self._scalar(lib().ttv_rmse)

    def tv(self):
        return self._scalar(lib().ttv_tv)

    def original_tv(self):
        return self._scalar(lib().ttv_original_tv)

    def tv_gd(self, ng, dPOCS):
        return self._scalar(lib().ttv_tv_gd, int(ng), C.c_float(dPOCS))

    def tv_fgp(self, ng, lam):
        return self._scalar(lib().ttv_tv_fgp, int(ng), C.c_float(lam))

    # ---- multi-GPU plumbing ---------------------------------------------------------------------
    @staticmethod
    def nccl_unique_id():
        buf = (C.c_char * 128)()
        check(lib().ttv_nccl_unique_id(buf))
        return bytes(buf)

    def comm_init(self, uid, rank, nranks):
        buf = (C.c_char * 128).from_buffer_copy(uid)
        check(lib().ttv_comm_init(self._h, buf, int(rank), int(nranks)))

    def p2p_export(self):
        buf = (C.c_char * 64)()
        check(lib().ttv_p2p_export(self._h, buf))
        return bytes(buf)

    def p2p_attach(self, left, right):
        l = None if left is None else (C.c_char * 64).from_buffer_copy(left)
        r = None if right is None else (C.c_char * 64).from_buffer_copy(right)
        check(lib().ttv_p2p_attach(self._h, l, r))

    def p2p_attach_local(self, left, right):
        check(lib().ttv_p2p_attach_local(self._h, None if left is None else left._h, None if right is None else right._h))

    def set_rank(self, rank, nranks):
        check(lib().ttv_set_rank(self._h, int(rank), int(nranks)))

    def setup_sharding(self, dist, rank, world, p2p=True):
        """NCCL communicator + (optionally) CUDA-IPC peer buffers, using an initialised
        torch.distributed process group for the rendezvous."""
        from .sharded import exchange_handles
        uid, left, right = exchange_handles(dist, rank, world, tomoengine.nccl_unique_id,
                                            self.p2p_export() if (p2p and world > 1) else None)
        self.comm_init(uid, rank, world)
        if p2p and world > 1:
            # ring neighbours (the periodic TV value needs rank 0 <-> last); the FGP chain ignores the wrap-around
            self.p2p_attach(left, right)

    # ---- measurement / test hooks (not part of the reference API) -------------------------------
    def sync(self):
        check(lib().ttv_sync(self._h))

    def set_option(self, name, value):
        check(lib().ttv_set_option(self._h, str(name).encode(), int(value)))

    def get_trig(self):
        c = np.empty(self.Nproj, np.float32)
        s = np.empty(self.Nproj, np.float32)
        check(lib().ttv_get_trig(self._h, _p(c), _p(s)))
        return c, s

    def get_sirt_weights(self):
        R = np.empty((self.Nproj, self.Nray), np.float32)
        Cw = np.empty((self.Nray, self.Nray), np.float32)
        check(lib().ttv_get_sirt_weights(self._h, _p(R), _p(Cw)))
        return R, Cw

    def set_model_projections(self, g):
        g = _as_f32(g)
        check(lib().ttv_set_model_projections(self._h, _p(g), g.shape[0], g.shape[1]))

    def back_projection(self):
        check(lib().ttv_back_projection(self._h))

    def time_kernel(self, name, reps=10, arg_i=0, arg_f=0.0):
        v = C.c_float(0)
        check(lib().ttv_time_kernel(self._h, name.encode(), int(reps), int(arg_i), C.c_float(arg_f), C.byref(v)))
        return v.value

    PROFILE_GROUPS = ("fp", "bp", "fgp_first", "fgp_sweep", "fgp_final", "tv", "momentum", "tvgd_grad",
                      "tvgd_update", "reduce", "elementwise", "sart_fp", "sart_bp", "io", "halo", "fgp2_first", "fgp2_sweep", "fgp2_final")

    def profile(self, enable=True):
        check(lib().ttv_profile(self._h, int(bool(enable))))

    def profile_report(self):
        """{group: (total_ms, launches)} measured with CUDA events on the engine's stream."""
        out = {}
        for g in self.PROFILE_GROUPS:
            ms, n = C.c_double(0), C.c_longlong(0)
            check(lib().ttv_profile_get(self._h, g.encode(), C.byref(ms), C.byref(n)))
            if n.value:
                out[g] = (ms.value, n.value)
        return out

    def timer_start(self):
        check(lib().ttv_timer_start(self._h))

    def timer_stop(self):
        v = C.c_float(0)
        check(lib().ttv_timer_stop(self._h, C.byref(v)))
        return v.value

    def selftest_invsqrt(self, n=1 << 24, seed=1):
        v = C.c_ulonglong(0)
        check(lib().ttv_selftest_invsqrt(self._h, C.c_ulonglong(int(n)), C.c_uint(int(seed)), C.byref(v)))
        return v.value

    def launch_count(self):
        v = C.c_longlong(0)
        check(lib().ttv_launch_count(self._h, C.byref(v)))
        return v.value

    def sizes(self):
        V, S, n = C.c_longlong(0), C.c_longlong(0), C.c_longlong(0)
        check(lib().ttv_sizes(self._h, C.byref(V), C.byref(S), C.byref(n)))
        return V.value, S.value, n.value
