mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q -x --timeout 600 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
timeout 300 python scripts/kprobe.py 512 512 180 10 tv fgp_sweep fgp_first fgp_final > gpurun_out/kprobe.log 2>&1
timeout 900 python bench.py --steps 20 --warmup 3 --no-cpu > gpurun_out/bench_512.log 2>&1; echo "bench512 exit $?" >> gpurun_out/bench_512.log
tail -n 3 gpurun_out/pytest_gpu.log; cat gpurun_out/kprobe.log; cut -c1-300 gpurun_out/bench_512.log
