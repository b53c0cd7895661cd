mkdir -p gpurun_out
timeout -k 5 240 python -m pytest tests/test_gpu_table_kernels.py -m gpu -x -q --timeout 60 > gpurun_out/r2c_pytest_table.log 2>&1; echo "pytest table exit $?"
tail -n 15 gpurun_out/r2c_pytest_table.log | cut -c1-400
timeout -k 5 120 python scripts/kernel_times.py > gpurun_out/r2c_times.log 2>&1; cat gpurun_out/r2c_times.log | tail -8
