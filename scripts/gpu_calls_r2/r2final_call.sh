# final verification of the committed tree (ONE GPU): full GPU suite, smoke, bench
mkdir -p gpurun_out
timeout -k 10 900 python -m pytest tests -m gpu -q --timeout 300 > gpurun_out/r2final_pytest_gpu.log 2>&1; echo "pytest exit $?"
tail -n 4 gpurun_out/r2final_pytest_gpu.log | cut -c1-300
timeout -k 10 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2final_smoke.log 2>&1; echo "smoke exit $?"; tail -n 2 gpurun_out/r2final_smoke.log
timeout -k 10 600 python bench.py --steps 20 --warmup 3 > gpurun_out/r2final_bench512.json 2> gpurun_out/r2final_bench512.err; echo "bench exit $?"
python - <<'PY'
import json
d = json.loads(open("gpurun_out/r2final_bench512.json").read().strip().splitlines()[-1])
print(round(d["value"], 2), "it/s e2e", round(d["e2e"]["value"], 2), d["recon_sha256"][:12], d["tv_fgp_sha256"][:12], d["roofline"]["frac"], d["roofline"].get("traffic"), d["roofline"].get("traffic_frac_of_peak"), d["gpu_launches"], d["clocks"])
PY
