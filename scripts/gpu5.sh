set -x
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q -x --timeout 600 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
timeout 900 python bench.py --steps 10 --warmup 3 --no-cpu > gpurun_out/bench_512.log 2>&1; echo "bench512 exit $?" >> gpurun_out/bench_512.log
tail -n 15 gpurun_out/pytest_gpu.log; cut -c1-3000 gpurun_out/bench_512.log
