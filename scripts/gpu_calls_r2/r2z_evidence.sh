# Evidence pass (ONE GPU): full GPU suite, smoke, bench (+ reference arm), launch list, ncu --set full of the default kernels
mkdir -p gpurun_out
timeout -k 10 900 python -m pytest tests -m gpu -q --timeout 300 > gpurun_out/r2z_pytest_gpu.log 2>&1; echo "pytest exit $?"
tail -n 5 gpurun_out/r2z_pytest_gpu.log | cut -c1-300
timeout -k 10 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2z_smoke.log 2>&1; echo "smoke exit $?"; tail -n 2 gpurun_out/r2z_smoke.log
timeout -k 10 600 python bench.py --steps 20 --warmup 3 > gpurun_out/r2z_bench512.json 2> gpurun_out/r2z_bench512.err; echo "bench exit $?"
cut -c1-200 gpurun_out/r2z_bench512.json
timeout -k 10 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r2z_bench512_ref.json 2> gpurun_out/r2z_bench512_ref.err; echo "bench ref exit $?"
cut -c1-300 gpurun_out/r2z_bench512_ref.json
timeout -k 10 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2z_launches_bench512.csv python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/r2z_ncu_launch.log 2>&1; echo "ncu launches exit $?"
timeout -k 10 900 ncu --set full --clock-control none --import-source on -k regex:'fgp2_kernel|fpt_kernel|bpt_kernel|tv_value|momentum' -s 24 -c 12 -o gpurun_out/r2z_full python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/r2z_ncu_full.log 2>&1; echo "ncu full exit $?"
ncu -i gpurun_out/r2z_full.ncu-rep --page details --csv > gpurun_out/r2z_full_details.csv 2>/dev/null
ncu -i gpurun_out/r2z_full.ncu-rep --page raw --csv > gpurun_out/r2z_full_raw.csv 2>/dev/null
rm -f gpurun_out/r2z_full.ncu-rep
timeout -k 10 300 python scripts/bench_algos.py --steps 3 > gpurun_out/r2z_bench_algos_512.jsonl 2> gpurun_out/r2z_bench_algos_512.err; echo "bench_algos 512 exit $?"
timeout -k 10 300 python scripts/bench_algos.py --size 256 --angles 71 --steps 5 > gpurun_out/r2z_bench_algos_256.jsonl 2> gpurun_out/r2z_bench_algos_256.err; echo "bench_algos 256 exit $?"
timeout -k 10 300 python bench.py --size 256 --angles 90 --steps 50 --warmup 3 --no-cpu > gpurun_out/r2z_bench256.json 2> gpurun_out/r2z_bench256.err; echo "bench 256 exit $?"
