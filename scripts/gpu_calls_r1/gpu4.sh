set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/gpu.txt 2>&1
timeout 1200 python -m pytest tests -m gpu -q --timeout 600 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_512.log 2>&1; echo "bench512 exit $?" >> gpurun_out/bench_512.log
timeout 600 python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/bench_short.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/ncu_list.log 2>&1
echo "ncu list exit $?" >> gpurun_out/ncu_list.log
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:'fgp_sweep|fp_kernel|bp_kernel|tv_value|momentum' -s 40 -c 8 -o gpurun_out/r1_full python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/ncu_full.log 2>&1
echo "ncu full exit $?" >> gpurun_out/ncu_full.log
tail -n 4 gpurun_out/pytest_gpu.log gpurun_out/ncu_list.log gpurun_out/ncu_full.log; cut -c1-400 gpurun_out/bench_512.log
