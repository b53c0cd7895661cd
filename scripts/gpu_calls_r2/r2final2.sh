#!/bin/bash
# final tree of round 2: GPU suite, smoke, bench at N = 1 (hard limits; stops at the first failure)
mkdir -p gpurun_out
timeout 110 python -m pytest tests -x -q -m gpu > gpurun_out/r2final2_pytest_gpu.log 2>&1
rc=$?; tail -n 3 gpurun_out/r2final2_pytest_gpu.log; [ $rc -ne 0 ] && exit $rc
timeout 40 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2final2_smoke.log 2>&1
rc=$?; tail -n 2 gpurun_out/r2final2_smoke.log; [ $rc -ne 0 ] && exit $rc
timeout 60 python bench.py --no-cpu > gpurun_out/r2final2_bench512.json 2> gpurun_out/r2final2_bench512.err
rc=$?; python - <<'PY'
import json
try:
    d=json.loads([l for l in open('gpurun_out/r2final2_bench512.json') if l.startswith('{')][-1])
    print(d['value'], d['ms_per_step'], d['e2e']['value'], d['recon_sha256'][:16], d['tv_fgp_sha256'][:16])
except Exception as e: print('no json', e)
PY
exit $rc
