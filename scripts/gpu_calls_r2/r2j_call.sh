# Round-2 re-entry call (ONE GPU): full GPU suite, smoke, bench, projector variant times, launch list, ncu --set full of the defaults
mkdir -p gpurun_out
timeout -k 10 900 python -m pytest tests -m gpu -q --timeout 300 > gpurun_out/r2j_pytest_gpu.log 2>&1; echo "pytest exit $?"
tail -n 8 gpurun_out/r2j_pytest_gpu.log | cut -c1-300
timeout -k 10 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2j_smoke.log 2>&1; echo "smoke exit $?"; tail -n 2 gpurun_out/r2j_smoke.log
timeout -k 10 600 python bench.py --steps 20 --warmup 3 > gpurun_out/r2j_bench512.json 2> gpurun_out/r2j_bench512.err; echo "bench exit $?"
cut -c1-300 gpurun_out/r2j_bench512.json; tail -3 gpurun_out/r2j_bench512.err
timeout -k 5 200 python scripts/kernel_times.py > gpurun_out/r2j_times.log 2>&1; tail -8 gpurun_out/r2j_times.log
timeout -k 10 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2j_launches_bench512.csv python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/r2j_ncu_launch.log 2>&1; echo "ncu launches exit $?"
timeout -k 10 900 ncu --set full --clock-control none --import-source on -k regex:'fgp_tile|fpt_kernel|bpt_kernel|tv_value|momentum' -s 30 -c 16 -o gpurun_out/r2j_full python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/r2j_ncu_full.log 2>&1; echo "ncu full exit $?"
ncu -i gpurun_out/r2j_full.ncu-rep --page details --csv > gpurun_out/r2j_full_details.csv 2>/dev/null
ncu -i gpurun_out/r2j_full.ncu-rep --page raw --csv > gpurun_out/r2j_full_raw.csv 2>/dev/null
ncu -i gpurun_out/r2j_full.ncu-rep --page source --csv --kernel-name regex:fpt_kernel > gpurun_out/r2j_src_fpt.csv 2>/dev/null
ncu -i gpurun_out/r2j_full.ncu-rep --page source --csv --kernel-name regex:bpt_kernel > gpurun_out/r2j_src_bpt.csv 2>/dev/null
ls -la gpurun_out/ | head -30
rm -f gpurun_out/r2j_full.ncu-rep
