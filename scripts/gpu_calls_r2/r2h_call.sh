mkdir -p gpurun_out
timeout -k 10 300 python -m pytest tests/test_gpu_table_kernels.py -m gpu -q --timeout 60 > gpurun_out/r2h_pytest.log 2>&1; echo "pytest exit $?"
tail -n 8 gpurun_out/r2h_pytest.log | cut -c1-300
timeout -k 5 120 python scripts/kernel_times.py 512 512 180 5 bp_xb=8:bp_sirt bp_xb=2:bp_sirt > gpurun_out/r2h_times.log 2>&1; cat gpurun_out/r2h_times.log | tail -8
timeout -k 5 120 python scripts/kernel_times.py 512 64 180 10 bp_variant=1:bp_sirt bp_variant=2:bp_sirt fp_variant=1:fp_sirt fp_variant=2:fp_sirt > gpurun_out/r2h_times64.log 2>&1; cat gpurun_out/r2h_times64.log | tail -8
