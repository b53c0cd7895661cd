mkdir -p gpurun_out
timeout -k 10 300 compute-sanitizer --tool memcheck python -m pytest tests/test_gpu_parity.py -m gpu -q -x --timeout 280 -k "two_sweeps and (33-63-4 or 1-5-2 or 65-58-7)" > gpurun_out/r2l_sanitizer.log 2>&1; echo "sanitizer exit $?"
grep -E "ERROR SUMMARY|Invalid|at 0x|by thread" gpurun_out/r2l_sanitizer.log | head -12
timeout -k 10 600 python -m pytest tests/test_gpu_parity.py -m gpu -q --timeout 300 -k "two_sweeps or inverse_sqrt or goldens or fgp" > gpurun_out/r2l_pytest.log 2>&1; echo "pytest exit $?"
tail -n 15 gpurun_out/r2l_pytest.log | cut -c1-400
timeout -k 5 200 python scripts/kernel_times.py 512 512 180 10 :fgp_first :fgp2_first :fgp_sweep :fgp2_sweep :fgp_final :fgp2_final > gpurun_out/r2l_times.log 2>&1; tail -8 gpurun_out/r2l_times.log
timeout -k 10 600 python bench.py --steps 20 --warmup 3 --no-cpu > gpurun_out/r2l_bench512.json 2> gpurun_out/r2l_bench512.err; echo "bench exit $?"
python - <<'PY'
import json
d = json.loads(open("gpurun_out/r2l_bench512.json").read().strip().splitlines()[-1])
print(round(d["value"], 2), "it/s", round(d["ms_per_step"], 3), "ms e2e", round(d["e2e"]["value"], 2), {k: (round(v["ms_per_launch"], 3), v["launches_per_step"]) for k, v in d["kernels"].items()})
PY
tail -3 gpurun_out/r2l_bench512.err
