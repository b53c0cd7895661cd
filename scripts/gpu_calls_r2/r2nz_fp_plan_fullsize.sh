#!/bin/bash
# the planned mixed forward-projector launch at slab sizes of the 512^3 x 180 volume; stops at the first failure
mkdir -p gpurun_out
timeout 150 python -m pytest tests/test_gpu_table_kernels.py -x -q -m gpu -k "planned_mixed or default_choice" > gpurun_out/r2nz_pytest.log 2>&1
rc=$?; tail -n 5 gpurun_out/r2nz_pytest.log; exit $rc
