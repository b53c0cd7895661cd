# round-1 re-validation of the current tree: GPU parity tests, 512^3 bench, ncu launch list, --set full of the hot kernels
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q --timeout 300 > gpurun_out/r1b_pytest_gpu.log 2>&1; echo "pytest exit $?"
tail -n 3 gpurun_out/r1b_pytest_gpu.log
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/r1b_bench512.json 2> gpurun_out/r1b_bench512.err; echo "bench exit $?"
cut -c1-600 gpurun_out/r1b_bench512.json
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r1b_launches_bench512.csv python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/r1b_ncu_launch.log 2>&1; echo "ncu launches exit $?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'fgp_tile|fp_kernel|bp_kernel|tv_value|momentum' -s 30 -c 16 -o gpurun_out/r1b_full python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/r1b_ncu_full.log 2>&1; echo "ncu full exit $?"
ncu -i gpurun_out/r1b_full.ncu-rep --page details --csv > gpurun_out/r1b_full_details.csv 2>/dev/null
ls -la gpurun_out | head -20
