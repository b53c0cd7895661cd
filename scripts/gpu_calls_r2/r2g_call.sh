mkdir -p gpurun_out
timeout -k 10 600 python -m pytest tests -m gpu -q --timeout 120 > gpurun_out/r2g_pytest_gpu.log 2>&1; echo "pytest exit $?"
tail -n 8 gpurun_out/r2g_pytest_gpu.log | cut -c1-300
