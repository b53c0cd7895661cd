mkdir -p gpurun_out
timeout -k 5 240 python -m pytest tests/test_gpu_table_kernels.py -m gpu -x -q --timeout 60 > gpurun_out/r2d_pytest_table.log 2>&1; echo "pytest table exit $?"
tail -n 5 gpurun_out/r2d_pytest_table.log | cut -c1-400
timeout -k 5 120 python scripts/kernel_times.py > gpurun_out/r2d_times.log 2>&1; cat gpurun_out/r2d_times.log | tail -8
timeout -k 5 120 python scripts/ncu_variants.py 512 128 180 fp_variant=2:fp_sirt bp_variant=2:bp_sirt > gpurun_out/r2d_plain.log 2>&1 && \
timeout -k 10 600 ncu --set full --clock-control none --import-source on -k regex:'fpt_kernel|bpt_kernel' -o gpurun_out/r2d_table python scripts/ncu_variants.py 512 128 180 fp_variant=2:fp_sirt bp_variant=2:bp_sirt > gpurun_out/r2d_ncu.log 2>&1; echo "ncu exit $?"
cat gpurun_out/r2d_plain.log; tail -3 gpurun_out/r2d_ncu.log
