mkdir -p gpurun_out
timeout -k 10 900 python -m pytest tests -m gpu -q --timeout 300 > gpurun_out/r2o_pytest.log 2>&1; echo "pytest exit $?"
tail -n 12 gpurun_out/r2o_pytest.log | cut -c1-600
timeout -k 10 600 python scripts/bench_algos.py --steps 3 --only sart_pass,asd_pocs > gpurun_out/r2o_bench_algos_512.jsonl 2> gpurun_out/r2o_bench_algos_512.err; echo "bench_algos 512 exit $?"
timeout -k 10 600 python scripts/bench_algos.py --size 256 --angles 71 --steps 5 --only sart_pass,asd_pocs > gpurun_out/r2o_bench_algos_256.jsonl 2> gpurun_out/r2o_bench_algos_256.err; echo "bench_algos 256 exit $?"
python - <<'PY'
import json
for f in ("gpurun_out/r2o_bench_algos_512.jsonl", "gpurun_out/r2o_bench_algos_256.jsonl"):
    for l in open(f):
        d = json.loads(l)
        print(d["size"], d["workload"], round(d["ms_per_step"], 3), "ms", round(d["steps_per_s"], 2), "/s", {k: round(v["ms_per_launch"] * v["launches_per_step"], 3) for k, v in d["kernels"].items()})
PY
tail -3 gpurun_out/r2o_bench_algos_512.err
