"""N > 1 host logic on CPU: two processes over the gloo backend (no GPU, no compute calls).

Covers what `bench.py --gpus N` / `multigpuengine` do around the CUDA engine: the x-slab partition
(remainder rule of mpi_astra_ctvlib.cpp:53-58), the rendezvous that hands rank 0's NCCL id to every
rank and each rank's CUDA-IPC handle to its ring neighbours, the slab scatter / gather of host
arrays, and the max-over-ranks timing reduction of bench.py.
"""
import os
import socket

import numpy as np
import pytest


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close()
    return p


def _worker(rank, ws, port, nslice, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(ws))
    import torch
    import torch.distributed as dist
    dist.init_process_group("gloo", rank=rank, world_size=ws)
    try:
        from tomo_tv_b200 import sharded
        from tomo_tv_b200.phantom import slab_bounds
        out = {}
        assert sharded.world(dist) == (rank, ws)
        full = np.arange(nslice * 6, dtype=np.float32).reshape(nslice, 2, 3)
        mine = sharded.slab_of(full, dist)
        x0, cnt = slab_bounds(nslice, ws, rank)
        assert mine.shape == (cnt, 2, 3) and np.array_equal(mine, full[x0:x0 + cnt])
        back = sharded.gather_slabs(mine * 2, dist)
        assert np.array_equal(back, full * 2)
        uid, left, right = sharded.exchange_handles(dist, rank, ws, lambda: b"U" * 128, bytes([rank]) * 64)
        out["uid"], out["left"], out["right"] = uid, left, right
        # bench.py's timing rule: the step time is the max over ranks
        t = torch.tensor([10.0 + rank], dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        out["tmax"] = float(t)
        # scalar results are sums of per-slab partial sums (the engine all-reduces them in double)
        part = torch.tensor([float(np.sum(mine.astype(np.float64) ** 2))], dtype=torch.float64)
        dist.all_reduce(part)
        out["norm2"] = float(part)
        q.put((rank, out))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("nslice", [8, 7])
def test_two_rank_slab_plumbing_over_gloo(nslice):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, nslice, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = dict(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    for r in (0, 1):
        assert res[r]["uid"] == b"U" * 128
        assert res[r]["left"] == bytes([(r - 1) % 2]) * 64 and res[r]["right"] == bytes([(r + 1) % 2]) * 64
        assert res[r]["tmax"] == 11.0
    full = np.arange(nslice * 6, dtype=np.float64)
    assert res[0]["norm2"] == res[1]["norm2"] == float(np.sum(full ** 2))


def test_multigpuengine_is_exported_and_subclasses_tomoengine(ttv):
    from tomo_tv_b200.sharded import multigpuengine
    assert issubclass(multigpuengine, ttv.tomoengine)
    for m in ("get_gpu_ids", "is_multi_gpu_enabled", "print_gpu_usage", "SIRT", "SART", "CGLS", "FBP",
              "forward_projection", "poisson_ML"):
        assert callable(getattr(multigpuengine, m))      # multigpuengine.cpp:385-421
