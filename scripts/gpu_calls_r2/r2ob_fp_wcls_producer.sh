#!/bin/bash
# forward projector: producer warp of its own (producer warpgroup at 24 registers, 768 threads) + per-line box widths.  Stops at the first failure.
mkdir -p gpurun_out
timeout 70 python -m pytest tests/test_gpu_table_kernels.py -x -q -m gpu > gpurun_out/r2ob_pytest.log 2>&1
rc=$?; tail -n 3 gpurun_out/r2ob_pytest.log; [ $rc -ne 0 ] && exit $rc
L=gpurun_out/r2ob_times.log; : > $L
for nx in 512 64; do
  echo "== $nx slices" >> $L
  timeout 60 python scripts/kernel_times.py 512 $nx 180 10 fp_xb=2:fp_sirt fp_xb=1:fp_sirt fp_xb=0:fp_sirt fp_wcls=0:fp_sirt fp_xb=1:fp_sirt fp_xb=2:fp_sirt >> $L 2>&1 || { cat $L; exit 1; }
done
cat $L
