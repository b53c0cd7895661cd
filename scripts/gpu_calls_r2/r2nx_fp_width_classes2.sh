#!/bin/bash
# per-line TMA box widths, packed table + one-line-ahead fetch: A/B against the widest box on every line (fp_wcls = 0)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_table_kernels.py -x -q -m gpu > gpurun_out/r2nx_pytest.log 2>&1; echo "pytest rc=$?"
tail -n 3 gpurun_out/r2nx_pytest.log
L=gpurun_out/r2nx_times.log; : > $L
for nx in 512 256 64; do
  echo "== $nx slices" >> $L
  timeout 300 python scripts/kernel_times.py 512 $nx 180 10 fp_xb=2:fp_sirt fp_xb=1:fp_sirt fp_xb=0:fp_sirt fp_wcls=0:fp_sirt fp_xb=1:fp_sirt fp_xb=2:fp_sirt >> $L 2>&1
done
cat $L
