mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q --timeout 300 > gpurun_out/r1f_pytest_gpu.log 2>&1; echo "pytest exit $?"
tail -n 6 gpurun_out/r1f_pytest_gpu.log | cut -c1-300
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/r1f_bench512.json 2> gpurun_out/r1f_bench512.err; echo "bench exit $?"
cut -c1-400 gpurun_out/r1f_bench512.json
