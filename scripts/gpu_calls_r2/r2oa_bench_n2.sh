#!/bin/bash
# bench at 2 GPUs with the planned mixed forward-projector launch (768 pair items -> 740 + 56 short CTAs per rank); one run, hard limits
mkdir -p gpurun_out
timeout 200 python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --nproc-per-node 2 --master-port 29612 bench.py --gpus 2 --steps 40 --warmup 3 > gpurun_out/r2oa_bench512_n2.json 2> gpurun_out/r2oa_n2.err
rc=$?; echo "bench n2 rc=$rc"
python - <<'PY'
import json
try:
    d=json.loads([l for l in open('gpurun_out/r2oa_bench512_n2.json') if l.startswith('{')][-1])
    print(d['value'], d['ms_per_step'], d['e2e']['value'], d['recon_sha256'], d.get('tv_fgp_sha256'))
    print({k:round(v['ms_per_launch'],3) for k,v in d['kernels'].items()})
except Exception as e: print('no json', e)
PY
exit $rc
