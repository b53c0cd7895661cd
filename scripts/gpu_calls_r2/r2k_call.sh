# goldens for geometry C (+ gd10, spread), parity table, secondary workloads, full-size c2 test
mkdir -p gpurun_out
timeout -k 10 900 python oracle/make_golden.py --out gpurun_out/golden > gpurun_out/r2k_make_golden.log 2>&1; echo "make_golden exit $?"
tail -n 3 gpurun_out/r2k_make_golden.log | cut -c1-300
timeout -k 10 600 python -m pytest tests/test_gpu_fullsize.py -m gpu -q --timeout 600 > gpurun_out/r2k_pytest_fullsize.log 2>&1; echo "pytest fullsize exit $?"
tail -n 5 gpurun_out/r2k_pytest_fullsize.log | cut -c1-300
timeout -k 10 900 python scripts/bench_algos.py --steps 3 > gpurun_out/r2k_bench_algos_512.jsonl 2> gpurun_out/r2k_bench_algos_512.err; echo "bench_algos 512 exit $?"
timeout -k 10 600 python scripts/bench_algos.py --size 256 --angles 71 --steps 5 > gpurun_out/r2k_bench_algos_256.jsonl 2> gpurun_out/r2k_bench_algos_256.err; echo "bench_algos 256 exit $?"
python - <<'PY'
import json
for f in ("gpurun_out/r2k_bench_algos_512.jsonl", "gpurun_out/r2k_bench_algos_256.jsonl"):
    for l in open(f):
        d = json.loads(l)
        print(d["size"], d["workload"], round(d["ms_per_step"], 3), "ms", round(d["steps_per_s"], 2), "/s", {k: round(v["ms_per_launch"] * v["launches_per_step"], 3) for k, v in d["kernels"].items()})
PY
