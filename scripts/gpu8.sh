set -x
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 scripts/mgpu_check.py > gpurun_out/mgpu_check.log 2>&1; echo "mgpu exit $?" >> gpurun_out/mgpu_check.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/bench_2gpu.log 2>&1; echo "bench2 exit $?" >> gpurun_out/bench_2gpu.log
timeout 600 python bench.py --gpus 1 --steps 10 --warmup 3 --no-cpu > gpurun_out/bench_1gpu.log 2>&1
timeout 900 python -m pytest tests -m gpu -q -x --timeout 600 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
grep -E "FAIL|PASS|exit" gpurun_out/mgpu_check.log; tail -n 3 gpurun_out/bench_2gpu.log | cut -c1-200; cut -c1-200 gpurun_out/bench_1gpu.log; tail -3 gpurun_out/pytest_gpu.log
