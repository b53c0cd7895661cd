#!/bin/bash
# forward projector with per-line TMA box widths (FP_NCLS width classes): parity of the table kernels, then times
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_table_kernels.py tests/test_gpu_parity.py -x -q -m gpu > gpurun_out/r2nw_pytest.log 2>&1; echo "pytest rc=$?"
tail -n 3 gpurun_out/r2nw_pytest.log
L=gpurun_out/r2nw_times.log; : > $L
for nx in 512 256 64; do
  echo "== $nx slices" >> $L
  timeout 300 python scripts/kernel_times.py 512 $nx 180 10 fp_variant=1:fp_sirt fp_variant=0:fp_sirt fp_xb=2:fp_sirt fp_xb=1:fp_sirt fp_xb=0:fp_sirt >> $L 2>&1
done
cat $L
