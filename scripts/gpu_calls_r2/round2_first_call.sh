# First GPU call of the next round (ONE GPU): validates everything changed after the last round-1 measurement
# (FFMA2 projector accumulators, 4V momentum, TV float partials, first-call all-reduce) and refreshes the
# evidence.  Every step has its own short timeout; nothing here can hold the box for long.
mkdir -p gpurun_out
timeout -k 10 900 python -m pytest tests -m gpu -x -q --timeout 600 > gpurun_out/r2a_pytest_gpu.log 2>&1; echo "pytest exit $?"
tail -n 5 gpurun_out/r2a_pytest_gpu.log | cut -c1-300
timeout -k 10 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2a_smoke.log 2>&1; echo "smoke exit $?"; tail -n 2 gpurun_out/r2a_smoke.log
timeout -k 10 600 python bench.py --steps 20 --warmup 3 > gpurun_out/r2a_bench512.json 2> gpurun_out/r2a_bench512.err; echo "bench exit $?"
cut -c1-400 gpurun_out/r2a_bench512.json
timeout -k 10 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2a_launches_bench512.csv python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/r2a_ncu_launch.log 2>&1; echo "ncu launches exit $?"
timeout -k 10 900 ncu --set full --clock-control none --import-source on -k regex:'fgp_tile|fp_kernel|bp_kernel|tv_value|momentum' -s 30 -c 16 -o gpurun_out/r2a_full python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/r2a_ncu_full.log 2>&1; echo "ncu full exit $?"
ncu -i gpurun_out/r2a_full.ncu-rep --page details --csv > gpurun_out/r2a_full_details.csv 2>/dev/null
ncu -i gpurun_out/r2a_full.ncu-rep --page raw --csv --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum > gpurun_out/r2a_dram_raw.csv 2>/dev/null
rm -f gpurun_out/r2a_full.ncu-rep
# secondary workloads (SIRT, SART pass, KL, CGLS, ASD-POCS iteration, FBP): first measurements of the non-headline paths
timeout -k 10 600 python scripts/bench_algos.py --steps 3 > gpurun_out/r2a_bench_algos.jsonl 2> gpurun_out/r2a_bench_algos.err; echo "bench_algos exit $?"
cut -c1-200 gpurun_out/r2a_bench_algos.jsonl
# experimental kernels, one variant per process (a faulting kernel must not take the others or the main suite down)
for v in fp1 fp8 bp5 bp8 bp9; do
  TTV_EXPERIMENTAL=1 timeout -k 10 300 python -m pytest tests/test_gpu_experimental.py -m gpu -q --timeout 200 -k $v > gpurun_out/r2a_pytest_exp_$v.log 2>&1; echo "experimental $v exit $?"
  tail -n 4 gpurun_out/r2a_pytest_exp_$v.log | cut -c1-300
done
# how fast are the experimental projector variants (only meaningful if the experimental tests above passed)
timeout -k 10 900 python scripts/variants_probe.py > gpurun_out/r2a_variants.log 2>&1; cat gpurun_out/r2a_variants.log
