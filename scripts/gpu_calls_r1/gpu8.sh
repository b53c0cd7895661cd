set -x
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 300 $TR --master-port 29511 scripts/mgpu_check.py > gpurun_out/mgpu_check.log 2>&1; echo "mgpu exit $?" >> gpurun_out/mgpu_check.log
timeout 300 $TR --master-port 29513 scripts/mgpu_check.py 64 100 12 >> gpurun_out/mgpu_check.log 2>&1; echo "mgpu exit $?" >> gpurun_out/mgpu_check.log
timeout 600 $TR --master-port 29512 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/bench_2gpu.log 2>&1; echo "bench2 exit $?" >> gpurun_out/bench_2gpu.log
TTV_P2P=0 timeout 600 $TR --master-port 29514 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/bench_2gpu_nccl.log 2>&1
grep -E "FAIL|PASS|exit|rror" gpurun_out/mgpu_check.log | head; tail -n 3 gpurun_out/bench_2gpu.log | cut -c1-200; tail -n 2 gpurun_out/bench_2gpu_nccl.log | cut -c1-200
