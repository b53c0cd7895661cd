mkdir -p gpurun_out
timeout -k 10 600 python -m pytest tests -m gpu -x -q --timeout 120 > gpurun_out/r2e_pytest_gpu.log 2>&1; echo "pytest exit $?"
tail -n 12 gpurun_out/r2e_pytest_gpu.log | cut -c1-300
timeout -k 10 300 python bench.py --steps 20 --warmup 3 > gpurun_out/r2e_bench512.json 2> gpurun_out/r2e_bench512.err; echo "bench exit $?"
cut -c1-600 gpurun_out/r2e_bench512.json; tail -3 gpurun_out/r2e_bench512.err
