mkdir -p gpurun_out
timeout -k 10 600 python -m pytest tests -m gpu -q --timeout 120 > gpurun_out/r2f_pytest_gpu.log 2>&1; echo "pytest exit $?"
tail -n 12 gpurun_out/r2f_pytest_gpu.log | cut -c1-300
timeout -k 5 120 python scripts/kernel_times.py > gpurun_out/r2f_times.log 2>&1; cat gpurun_out/r2f_times.log | tail -8
