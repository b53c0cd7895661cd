mkdir -p gpurun_out
timeout -k 10 600 python -m pytest tests/test_gpu_table_kernels.py -m gpu -q --timeout 300 > gpurun_out/r2fx_pytest.log 2>&1; echo "pytest exit $?"
tail -n 4 gpurun_out/r2fx_pytest.log | cut -c1-400
timeout -k 5 200 python scripts/kernel_times.py 512 64 180 10 fp_variant=1:fp_sirt fp_variant=2:fp_sirt fp_xb=2:fp_sirt > gpurun_out/r2fx_times64.log 2>&1; tail -4 gpurun_out/r2fx_times64.log
timeout -k 5 200 python scripts/kernel_times.py 512 96 180 10 fp_variant=1:fp_sirt fp_variant=2:fp_sirt fp_xb=2:fp_sirt > gpurun_out/r2fx_times96.log 2>&1; tail -4 gpurun_out/r2fx_times96.log
timeout -k 5 200 python scripts/kernel_times.py 512 128 180 10 fp_variant=2:fp_sirt fp_xb=1:fp_sirt > gpurun_out/r2fx_times128.log 2>&1; tail -3 gpurun_out/r2fx_times128.log
timeout -k 5 200 python scripts/kernel_times.py 512 512 180 5 fp_variant=2:fp_sirt fp_xb=1:fp_sirt > gpurun_out/r2fx_times512.log 2>&1; tail -3 gpurun_out/r2fx_times512.log
