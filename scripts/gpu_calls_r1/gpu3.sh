set -x
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q --timeout 600 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
timeout 600 python scripts/fp_probe.py 512 128 3 > gpurun_out/fp_probe.log 2>&1; echo "probe exit $?" >> gpurun_out/fp_probe.log
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_512.log 2>&1; echo "bench512 exit $?" >> gpurun_out/bench_512.log
tail -n 4 gpurun_out/pytest_gpu.log gpurun_out/fp_probe.log; cut -c1-600 gpurun_out/bench_512.log
