#!/bin/bash
# sharded == whole volume on 2 GPUs, final tree (what tests/test_gpu_fullsize.py runs when it sees two GPUs); hard limit
mkdir -p gpurun_out
timeout 34 python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --master-port 29541 --nproc-per-node 2 scripts/mgpu_check.py 150 96 20 > gpurun_out/r2oc_mgpu2.log 2>&1
rc=$?; grep -E "PASS|FAIL|Error" gpurun_out/r2oc_mgpu2.log | tail -n 5; exit $rc
