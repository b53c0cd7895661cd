#!/bin/bash
# the allocation-free entry point: alone in a fresh process, then at the end of the whole GPU suite
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu > gpurun_out/r2na_alone.log 2>&1; echo "alone rc=$?" | tee -a gpurun_out/r2na_alone.log
timeout 900 python -m pytest tests -x -q -m gpu > gpurun_out/r2na_suite.log 2>&1; echo "suite rc=$?" | tee -a gpurun_out/r2na_suite.log
tail -n 3 gpurun_out/r2na_alone.log gpurun_out/r2na_suite.log
