mkdir -p gpurun_out/golden
timeout 900 python oracle/make_golden.py --out gpurun_out/golden --only cgls,fbp > gpurun_out/r1d_golden.log 2>&1; echo "golden exit $?"
grep -v "^#" gpurun_out/r1d_golden.log | tail -n 40
