set -x
mkdir -p gpurun_out
timeout 900 python oracle/make_golden.py > gpurun_out/golden.log 2>&1; echo "golden exit $?" >> gpurun_out/golden.log
timeout 1200 python -m pytest tests -m gpu -q --timeout 600 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
timeout 600 python scripts/fp_probe.py 512 128 3 > gpurun_out/fp_probe.log 2>&1; echo "probe exit $?" >> gpurun_out/fp_probe.log
timeout 300 python scripts/fp_probe.py 256 64 1 > gpurun_out/fp_probe_small.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:fp_kernel -s 2 -c 2 -o gpurun_out/prof_fp python scripts/fp_probe.py 256 64 1 > gpurun_out/ncu_fp.log 2>&1
echo "ncu exit $?" >> gpurun_out/ncu_fp.log
tail -n 5 gpurun_out/pytest_gpu.log gpurun_out/fp_probe.log gpurun_out/ncu_fp.log
