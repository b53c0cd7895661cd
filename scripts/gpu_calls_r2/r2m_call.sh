mkdir -p gpurun_out
timeout -k 10 600 python -m pytest tests/test_gpu_parity.py -m gpu -q --timeout 300 -k "two_sweeps or inverse_sqrt or goldens or fgp" > gpurun_out/r2m_pytest.log 2>&1; echo "pytest exit $?"
tail -n 6 gpurun_out/r2m_pytest.log | cut -c1-400
timeout -k 5 200 python scripts/kernel_times.py 512 512 180 10 :fgp2_first :fgp2_sweep :fgp2_final > gpurun_out/r2m_times.log 2>&1; tail -4 gpurun_out/r2m_times.log
timeout -k 10 600 python bench.py --steps 20 --warmup 3 --no-cpu > gpurun_out/r2m_bench512.json 2> gpurun_out/r2m_bench512.err; echo "bench exit $?"
python - <<'PY'
import json
d = json.loads(open("gpurun_out/r2m_bench512.json").read().strip().splitlines()[-1])
print(round(d["value"], 2), "it/s", round(d["ms_per_step"], 3), "ms e2e", round(d["e2e"]["value"], 2), {k: (round(v["ms_per_launch"], 3), v["launches_per_step"]) for k, v in d["kernels"].items()})
PY
tail -3 gpurun_out/r2m_bench512.err
