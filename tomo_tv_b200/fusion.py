"""Host-side mirror of the reference's pybind11 class `multimodal`
(/root/reference/tomofusion/chemistry/utils/multimodal.cpp:520-564): same method names, argument
meaning and array layouts, implemented as thin ctypes calls into the C ABI (ttv_mm_*,
include/tomo_tv_b200.h).  All compute runs in the CUDA library.
"""
import ctypes as C
import numpy as np

from ._native import lib, check
from .engine import _as_f32, _p


class multimodal:
    """multimodal(Ns, Nray, Nel, haadfAngles[rad], chemAngles[rad])  -- multimodal.cpp:49-109.

    Extra keyword arguments (not in the reference): `device`, `x_offset` / `nslice_global` for a
    tilt-axis slab.  The reference's 3-argument constructor (no geometry) cannot project and is not
    mirrored.
    """

    def __init__(self, Ns, Nray, Nelements, haadfAngles, chemAngles, device=-1, x_offset=0, nslice_global=None):
        self._h = C.c_void_p()
        self.Nslice, self.Nray, self.Nel = int(Ns), int(Nray), int(Nelements)
        ha, ca = _as_f32(np.ravel(haadfAngles)), _as_f32(np.ravel(chemAngles))
        self.NprojHaadf, self.NprojChem = int(ha.size), int(ca.size)
        nsg = self.Nslice if nslice_global is None else int(nslice_global)
        check(lib().ttv_mm_create_slab(C.byref(self._h), self.Nslice, self.Nray, self.Nel, _p(ha), self.NprojHaadf,
                                       _p(ca), self.NprojChem, int(device), int(x_offset), nsg))

    def __del__(self):
        h = getattr(self, "_h", None)
        if h is not None and h.value:
            try:
                lib().ttv_mm_destroy(h)
            except Exception:
                pass
            self._h = C.c_void_p()

    # ---- data ---------------------------------------------------------------------------------
    def initialize_initial_volume(self):
        check(lib().ttv_mm_initialize_initial_volume(self._h))

    def set_haadf_tilt_series(self, bh):
        b = _as_f32(bh)
        if b.ndim != 2:
            raise ValueError("set_haadf_tilt_series expects a [Nslice, NprojHaadf*Nray] matrix")
        check(lib().ttv_mm_set_haadf_tilt_series(self._h, _p(b), b.shape[0], b.shape[1]))

    def set_chem_tilt_series(self, bchem):
        b = _as_f32(bchem)
        if b.ndim != 2:
            raise ValueError("set_chem_tilt_series expects a [Nslice, Nel*NprojChem*Nray] matrix")
        check(lib().ttv_mm_set_chem_tilt_series(self._h, _p(b), b.shape[0], b.shape[1]))

    def set_measureHaadf(self, flag):
        check(lib().ttv_mm_set_measureHaadf(self._h, int(bool(flag))))

    def set_measureChem(self, flag):
        check(lib().ttv_mm_set_measureChem(self._h, int(bool(flag))))

    def measureHaadf(self):
        v = C.c_int(0); check(lib().ttv_mm_measureHaadf(self._h, C.byref(v))); return bool(v.value)

    def measureChem(self):
        v = C.c_int(0); check(lib().ttv_mm_measureChem(self._h, C.byref(v))); return bool(v.value)

    def set_gamma(self, gamma):
        check(lib().ttv_mm_set_gamma(self._h, C.c_float(gamma)))

    def gamma(self):
        v = C.c_float(0); check(lib().ttv_mm_gamma(self._h, C.byref(v))); return v.value

    def _slice(self, m):
        v = _as_f32(m)
        if v.shape != (self.Nray, self.Nray):
            raise ValueError("expected an [Nray, Nray] slice")
        return v

    def set_recon(self, m, element, s):
        check(lib().ttv_mm_set_recon(self._h, _p(self._slice(m)), int(element), int(s)))

    def set_original_volume(self, m, element, s):
        check(lib().ttv_mm_set_original_volume(self._h, _p(self._slice(m)), int(element), int(s)))

    def get_recon(self, element, s):
        out = np.empty((self.Nray, self.Nray), np.float32)
        check(lib().ttv_mm_get_recon(self._h, _p(out), int(element), int(s)))
        return out

    # bulk additions (not in the reference, which loops elements and slices in Python)
    def get_recon_all(self, element):
        out = np.empty((self.Nslice, self.Nray, self.Nray), np.float32)
        check(lib().ttv_mm_get_recon_all(self._h, _p(out), int(element)))
        return out

    def set_recon_all(self, vol, element):
        v = _as_f32(vol)
        if v.shape != (self.Nslice, self.Nray, self.Nray):
            raise ValueError("set_recon_all expects [Nslice, Nray, Nray]")
        check(lib().ttv_mm_set_recon_all(self._h, _p(v), int(element)))

    def get_gpu_id(self):
        v = C.c_int(0); check(lib().ttv_mm_get_gpu_id(self._h, C.byref(v))); return v.value

    get_gpu = get_gpu_id

    def set_gpu(self, gpu_id):
        check(lib().ttv_mm_set_gpu(self._h, int(gpu_id)))

    def load_sigma(self, sigma):
        """[3, nnz] float32 array of rows / cols / vals (chemistry/reconstructor.py:150-153)."""
        s = _as_f32(sigma)
        if s.ndim != 2 or s.shape[0] != 3:
            raise ValueError("load_sigma expects a [3, nnz] array (rows, cols, vals)")
        check(lib().ttv_mm_load_sigma(self._h, _p(s), s.shape[1]))

    # ---- operators ----------------------------------------------------------------------------
    def initialize_FP(self):
        check(lib().ttv_mm_initialize_FP(self._h))

    def initialize_BP(self):
        check(lib().ttv_mm_initialize_BP(self._h))

    def initialize_SIRT(self):
        check(lib().ttv_mm_initialize_SIRT(self._h))

    def initialize_SART(self, order):
        check(lib().ttv_mm_initialize_SART(self._h, str(order).encode()))

    def forward_projection(self, vol):
        v = _as_f32(np.ravel(vol))
        if v.size != self.Nray * self.Nray:
            raise ValueError("forward_projection expects Nray*Nray values")
        out = np.empty(self.NprojHaadf * self.Nray, np.float32)
        check(lib().ttv_mm_forward_projection(self._h, _p(v), _p(out)))
        return out

    def estimate_lipschitz(self):
        check(lib().ttv_mm_estimate_lipschitz(self._h))

    def rescale_projections(self):
        check(lib().ttv_mm_rescale_projections(self._h))

    def rescale_tomograms(self, scale):
        check(lib().ttv_mm_rescale_tomograms(self._h, C.c_float(scale)))

    def poisson_ml(self, lambdaCHEM):
        v = C.c_float(0)
        check(lib().ttv_mm_poisson_ml(self._h, C.c_float(lambdaCHEM), C.byref(v)))
        return v.value

    def data_fusion(self, lambdaHAADF, lambdaCHEM, nIter, method):
        a, b = C.c_float(0), C.c_float(0)
        check(lib().ttv_mm_data_fusion(self._h, C.c_float(lambdaHAADF), C.c_float(lambdaCHEM), int(nIter), str(method).encode(),
                                       C.byref(a), C.byref(b)))
        return a.value, b.value

    def sirt_data_fusion(self, lambdaHAADF, lambdaCHEM, nIter):
        return self.data_fusion(lambdaHAADF, lambdaCHEM, nIter, "SIRT")

    def sart_data_fusion(self, lambdaHAADF, lambdaCHEM):
        return self.data_fusion(lambdaHAADF, lambdaCHEM, 1, "SART")

    def chemical_SIRT(self, nIter):
        check(lib().ttv_mm_chemical_SIRT(self._h, int(nIter)))

    def chemical_SART(self, nIter):
        check(lib().ttv_mm_chemical_SART(self._h, int(nIter)))

    def data_distance(self):
        v = C.c_float(0); check(lib().ttv_mm_data_distance(self._h, C.byref(v))); return v.value

    def tv_gd(self, ng, dPOCS):
        v = C.c_float(0); check(lib().ttv_mm_tv_gd(self._h, int(ng), C.c_float(dPOCS), C.byref(v))); return v.value

    def tv_fgp_4D(self, ng, lambdaTV):
        v = C.c_float(0); check(lib().ttv_mm_tv_fgp_4D(self._h, int(ng), C.c_float(lambdaTV), C.byref(v))); return v.value

    def rmse(self):
        out = np.empty(self.Nel, np.float32)
        check(lib().ttv_mm_rmse(self._h, _p(out)))
        return out

    def get_haadf_projections(self):
        out = np.empty((self.Nslice, self.NprojHaadf * self.Nray), np.float32)
        check(lib().ttv_mm_get_haadf_projections(self._h, _p(out)))
        return out

    def get_chem_projections(self):
        out = np.empty((self.Nslice, self.Nel * self.NprojChem * self.Nray), np.float32)
        check(lib().ttv_mm_get_chem_projections(self._h, _p(out)))
        return out

    def get_model_projections(self):
        out = np.empty((self.Nslice, self.NprojHaadf * self.Nray), np.float32)
        check(lib().ttv_mm_get_model_projections(self._h, _p(out)))
        return out

    def restart_recon(self):
        check(lib().ttv_mm_restart_recon(self._h))

    # ---- measurement hooks ----------------------------------------------------------------------
    def launch_count(self):
        v = C.c_longlong(0); check(lib().ttv_mm_launch_count(self._h, C.byref(v))); return v.value

    def timer_start(self):
        check(lib().ttv_mm_timer_start(self._h))

    def timer_stop(self):
        v = C.c_float(0); check(lib().ttv_mm_timer_stop(self._h, C.byref(v))); return v.value


class multigpufusion(multimodal):
    """multigpufusion(Ns, Nray, Nel, haadfAngles, chemAngles) -- chemistry/utils/multigpufusion.hpp:21-54, binding
    multigpufusion.cpp:459-475.

    The reference fans the slices of ONE process out over the visible GPUs with OpenMP.  Here it is one process per
    GPU: under `torchrun --nproc-per-node G` (an initialised torch.distributed group) every rank constructs this with
    the GLOBAL Ns and gets the fusion engine of its own contiguous x-slab on GPU LOCAL_RANK (data methods then take /
    return the rank's slab, see tomo_tv_b200.sharded.slab_of / gather_slabs); cost values, data distances and the 4-D
    TV norms are all-reduced inside the engine (NCCL), so every rank sees the whole-volume numbers.  Without a process
    group it is a single-GPU `multimodal`.
    """

    def __init__(self, Ns, Nray, Nelements, haadfAngles, chemAngles, dist=None):
        import os
        from .sharded import world
        from .phantom import slab_bounds
        if dist is None:
            try:
                import torch.distributed as td
                dist = td if td.is_available() and td.is_initialized() else None
            except Exception:
                dist = None
        rank, ws = world(dist)
        self.global_nslice, self.rank, self.world_size, self._dist = int(Ns), rank, ws, dist
        x0, cnt = slab_bounds(int(Ns), ws, rank)
        self.x_offset = x0
        local = int(os.environ.get("LOCAL_RANK", "0")) if ws > 1 else -1
        super().__init__(cnt, Nray, Nelements, haadfAngles, chemAngles, device=local, x_offset=x0, nslice_global=int(Ns))
        if ws > 1:
            from .engine import tomoengine
            uid = [tomoengine.nccl_unique_id() if rank == 0 else None]
            dist.broadcast_object_list(uid, src=0)
            self.comm_init(uid[0], rank, ws)

    def comm_init(self, uid128, rank, nranks):
        buf = (C.c_char * 128).from_buffer_copy(bytes(uid128))
        check(lib().ttv_mm_comm_init(self._h, buf, int(rank), int(nranks)))

    def get_gpu_ids(self):
        import os
        if self.world_size == 1:
            return [max(self.get_gpu_id(), 0)]
        ids = [None] * self.world_size
        self._dist.all_gather_object(ids, int(os.environ.get("LOCAL_RANK", "0")))
        return ids

    def is_multi_gpu_enabled(self):
        return self.world_size > 1

    def print_gpu_usage(self):
        print(f"rank {self.rank}/{self.world_size}: slices [{self.x_offset}, {self.x_offset + self.Nslice}) of {self.global_nslice}")
