#!/bin/bash
# forward projector with both CTA sizes in one launch (fptm_kernel): parity of the variants, then times at 512, 256 and 64 slices
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_table_kernels.py -x -q -m gpu > gpurun_out/r2mx_pytest.log 2>&1; echo "pytest rc=$?"
tail -n 3 gpurun_out/r2mx_pytest.log
for nx in 512 256 128 64; do
  echo "== $nx slices" >> gpurun_out/r2mx_times.log
  timeout 300 python scripts/kernel_times.py 512 $nx 180 10 fp_xb=2:fp_sirt fp_xb=1:fp_sirt fp_xb=0:fp_sirt fp_xb=3:fp_sirt fp_xb=0:fp_sirt >> gpurun_out/r2mx_times.log 2>&1
done
cat gpurun_out/r2mx_times.log
