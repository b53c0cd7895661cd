"""Tilt-axis sharding: one process per GPU, each owning a contiguous x-slab.

This replaces the reference's `multigpuengine` (OpenMP slice fan-out inside one process,
/root/reference/tomofusion/gpu/utils/multigpuengine.cpp:113-361) and its unbuilt MPI variant
(mpi_astra_ctvlib.cpp): FP / BP / SIRT / SART / KL / fusion terms are slice-independent and need no
communication; the TV stencils exchange one-slice halos (peer memory or NCCL) and scalar results are
all-reduced inside the engine.  `torch.distributed` is used for the rendezvous only.

Host-side helpers (pure numpy + torch.distributed, usable with the gloo backend on CPU):
    slab_of(x, dist)             rows of a [Nslice, ...] host array owned by this rank
    gather_slabs(local, dist)    the full [Nslice, ...] array on every rank
    multigpuengine               drop-in for the reference class: a slab `tomoengine` when launched
                                 under torchrun, a plain single-GPU engine otherwise
"""
import os

import numpy as np

from .engine import tomoengine
from .phantom import slab_bounds


def world(dist=None):
    """(rank, world_size) of the initialised process group, (0, 1) without one."""
    if dist is not None and dist.is_available() and dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    return 0, 1


def slab_of(x, dist=None, nslice=None):
    """The contiguous block of slices of `x` ([Nslice, ...]) this rank owns (same remainder rule as the
    reference's MPI variant, mpi_astra_ctvlib.cpp:53-58)."""
    rank, ws = world(dist)
    n = int(np.shape(x)[0] if nslice is None else nslice)
    x0, cnt = slab_bounds(n, ws, rank)
    return np.ascontiguousarray(x[x0:x0 + cnt])


def gather_slabs(local, dist=None):
    """Concatenate the ranks' slabs along axis 0 on every rank (host side; works on gloo and nccl)."""
    rank, ws = world(dist)
    if ws == 1:
        return np.ascontiguousarray(local)
    parts = [None] * ws
    dist.all_gather_object(parts, np.ascontiguousarray(local))
    return np.concatenate(parts, axis=0)


def exchange_handles(dist, rank, ws, make_uid, my_handle):
    """Rendezvous used by tomoengine.setup_sharding: rank 0's NCCL unique id reaches every rank, every rank's
    CUDA-IPC handle reaches its ring neighbours.  Returns (uid, left_handle, right_handle)."""
    uid = [make_uid() if rank == 0 else None]
    dist.broadcast_object_list(uid, src=0)
    handles = [None] * ws
    dist.all_gather_object(handles, my_handle)
    return uid[0], handles[(rank - 1) % ws], handles[(rank + 1) % ws]


class multigpuengine(tomoengine):
    """multigpuengine(Nslice, Nray[, angles]) -- multigpuengine.cpp:385-421.

    Under `torchrun --nproc-per-node G` (an initialised torch.distributed group) every rank constructs this
    with the GLOBAL Nslice and gets the engine of its own slab on GPU LOCAL_RANK; data methods then take /
    return the rank's slab (use slab_of / gather_slabs).  Without a process group it is a single-GPU engine.
    """

    def __init__(self, Nslice, Nray, angles=None, dist=None, p2p=True):
        if dist is None:
            try:
                import torch.distributed as td
                dist = td if td.is_available() and td.is_initialized() else None
            except Exception:
                dist = None
        rank, ws = world(dist)
        self.global_nslice = int(Nslice)
        self.rank, self.world_size = rank, ws
        x0, cnt = slab_bounds(Nslice, ws, rank)
        self.x_offset = x0
        local = int(os.environ.get("LOCAL_RANK", "0")) if ws > 1 else -1
        super().__init__(cnt, Nray, angles, device=local, x_offset=x0, nslice_global=int(Nslice))
        if ws > 1:
            self.setup_sharding(dist, rank, ws, p2p=p2p)
        self._dist = dist

    def get_gpu_ids(self):
        """One GPU per rank: the ordinals of all ranks' devices on this node."""
        if self.world_size == 1:
            return [max(self.get_gpu_id(), 0)]
        ids = [None] * self.world_size
        self._dist.all_gather_object(ids, int(os.environ.get("LOCAL_RANK", "0")))
        return ids

    def is_multi_gpu_enabled(self):
        return self.world_size > 1

    def print_gpu_usage(self):
        print(f"rank {self.rank}/{self.world_size}: slices [{self.x_offset}, {self.x_offset + self.Nslice}) of {self.global_nslice}")
