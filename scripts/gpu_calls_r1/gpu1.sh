set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/gpu.txt 2>&1
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke exit $?" >> gpurun_out/smoke.log
timeout 900 python oracle/make_golden.py > gpurun_out/golden.log 2>&1; echo "golden exit $?" >> gpurun_out/golden.log
timeout 1200 python -m pytest tests -m gpu -q -x --timeout 600 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
timeout 600 python bench.py --size 256 --angles 90 --steps 10 --warmup 3 > gpurun_out/bench_256.log 2>&1; echo "bench256 exit $?" >> gpurun_out/bench_256.log
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_512.log 2>&1; echo "bench512 exit $?" >> gpurun_out/bench_512.log
tail -3 gpurun_out/smoke.log gpurun_out/golden.log gpurun_out/pytest_gpu.log gpurun_out/bench_256.log gpurun_out/bench_512.log
