#!/bin/bash
# mixed forward-projector launch chosen by the plan (fp_xb = 0): GPU suite, slab times, bench at N = 1
mkdir -p gpurun_out
timeout 900 python -m pytest tests -x -q -m gpu > gpurun_out/r2mz_pytest.log 2>&1; echo "pytest rc=$?"
tail -n 3 gpurun_out/r2mz_pytest.log
L=gpurun_out/r2mz_times.log; : > $L
for nx in 512 256 128 96 64; do
  echo "== $nx slices" >> $L
  timeout 300 python scripts/kernel_times.py 512 $nx 180 10 fp_xb=2:fp_sirt fp_xb=1:fp_sirt fp_xb=0:fp_sirt >> $L 2>&1
done
cat $L
timeout 600 python bench.py > gpurun_out/r2mz_bench512.json 2> gpurun_out/r2mz_bench512.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/r2mz_bench512.json') if l.startswith('{')][-1])
print(d['value'], d['ms_per_step'], d['e2e']['value'], d['recon_sha256'], d['tv_fgp_sha256'])
print({k:round(v['ms_per_launch'],3) for k,v in d['kernels'].items()})
PY
