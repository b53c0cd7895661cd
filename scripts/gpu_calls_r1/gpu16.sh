# round-1 third pass: GPU parity tests (incl. N=1024/2048), c5 slab on one GPU, 512^3 bench (both arms), ncu launch list + --set full
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q --timeout 600 > gpurun_out/r1h_pytest_gpu.log 2>&1; echo "pytest exit $?"
tail -n 5 gpurun_out/r1h_pytest_gpu.log | cut -c1-300
timeout 900 python scripts/c5_slab.py > gpurun_out/r1h_c5_slab.json 2> gpurun_out/r1h_c5_slab.err; echo "c5 slab exit $?"
cat gpurun_out/r1h_c5_slab.json; tail -n 3 gpurun_out/r1h_c5_slab.err
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/r1h_bench512.json 2> gpurun_out/r1h_bench512.err; echo "bench exit $?"
cut -c1-500 gpurun_out/r1h_bench512.json
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r1h_bench512_ref.json 2> gpurun_out/r1h_bench512_ref.err; echo "ref arm exit $?"
cat gpurun_out/r1h_bench512_ref.json
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r1h_launches_bench512.csv python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/r1h_ncu_launch.log 2>&1; echo "ncu launches exit $?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'fgp_tile|fp_kernel|bp_kernel|tv_value|momentum' -s 30 -c 16 -o gpurun_out/r1h_full python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/r1h_ncu_full.log 2>&1; echo "ncu full exit $?"
ncu -i gpurun_out/r1h_full.ncu-rep --page details --csv > gpurun_out/r1h_full_details.csv 2>/dev/null
ncu -i gpurun_out/r1h_full.ncu-rep --page raw --csv --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum > gpurun_out/r1h_dram_raw.csv 2>/dev/null
rm -f gpurun_out/r1h_full.ncu-rep
ls -la gpurun_out | tail -12
