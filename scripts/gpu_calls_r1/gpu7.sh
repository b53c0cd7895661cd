set -x
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q -x --timeout 600 -k "fgp or tv or fista or golden" > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
timeout 600 python scripts/kprobe.py 512 512 180 10 fgp_sweep fgp_first fgp_final > gpurun_out/kprobe.log 2>&1
tail -n 6 gpurun_out/pytest_gpu.log; cat gpurun_out/kprobe.log
