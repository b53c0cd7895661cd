"""Sharded multi-modal fusion == whole-volume fusion (run under torchrun by tests/test_gpu_fullsize.py, one rank per GPU).

Every rank builds `multigpufusion` (tomo_tv_b200.fusion: the rank's x-slab engine + NCCL all-reduces for the cost values, data
distances and 4-D TV norms) on its rows of the ChemicalTomo inputs and runs the data-fusion scenario of tests/scenarios_fusion.py;
rank 0 runs the same scenario on a whole-volume `multimodal` and compares.  Test infrastructure: inputs come from the CPU oracle.
Usage: torchrun --nproc-per-node 2 tests/mgpu_fusion_check.py [geom]"""
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE)); sys.path.insert(0, HERE)
import numpy as np
import torch
import torch.distributed as dist

import oracle as orc
import scenarios_fusion as SF
import tomo_tv_b200 as ttv
from tomo_tv_b200.phantom import slab_bounds

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
geom = sys.argv[1] if len(sys.argv) > 1 else "G"
n, els, hang, cang = SF.FGEOMS[geom]
nel = len(els)
orc.build()
inp = SF.make_inputs(geom, orc)
x0, cnt = slab_bounds(n, world, rank)
mine_in = dict(inp, bh=inp["bh"][x0:x0 + cnt], bchem=inp["bchem"][x0:x0 + cnt])
mm = ttv.multigpufusion(n, n, nel, inp["hang"], inp["cang"])
assert mm.is_multi_gpu_enabled() and mm.Nslice == cnt and len(mm.get_gpu_ids()) == world
mine = SF.sc_fusion(mm, mine_in, nel, cnt)
gathered = [None] * world
dist.all_gather_object(gathered, mine)
ok = True
if rank == 0:
    whole = SF.sc_fusion(ttv.multimodal(n, n, nel, inp["hang"], inp["cang"], device=local), inp, nel, n)
    for k, w in whole.items():
        w = np.asarray(w)
        if w.ndim == 4:
            g = np.concatenate([d[k] for d in gathered], axis=1)
        elif w.ndim >= 2:
            g = np.concatenate([d[k] for d in gathered], axis=0)
        else:
            g = np.asarray(gathered[0][k])
            for d in gathered[1:]:
                assert np.allclose(np.asarray(d[k]), g), f"{k}: ranks disagree on a global scalar"
        err = SF.rel_l2(g, w)
        flag = err <= 2e-6
        ok &= bool(flag)
        print(f"mgpu_fusion[{world}] {k:12s} rel err {err:.3e} {'ok' if flag else 'FAIL'}", flush=True)
    print("MGPU_FUSION_CHECK", "PASS" if ok else "FAIL", flush=True)
dist.barrier()
del mm
dist.destroy_process_group()
sys.exit(0 if ok else 1)
