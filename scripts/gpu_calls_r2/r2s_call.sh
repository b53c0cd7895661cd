mkdir -p gpurun_out
timeout -k 10 600 python -m pytest tests/test_gpu_parity.py -m gpu -q --timeout 300 -k "sart or goldens" > gpurun_out/r2s_pytest.log 2>&1; echo "pytest exit $?"
tail -n 4 gpurun_out/r2s_pytest.log | cut -c1-400
timeout -k 10 600 python scripts/bench_c3.py > gpurun_out/r2s_c3_n1.jsonl 2> gpurun_out/r2s_c3_n1.err; echo "c3 n1 exit $?"
python - <<'PY'
import json
for l in open("gpurun_out/r2s_c3_n1.jsonl"):
    try:
        d = json.loads(l); print(d["workload"], d["n_gpus"], round(d["ms_per_step"], 3), "ms", d["recon_sha256"][:16], {k: round(v["ms_per_step"], 3) for k, v in d["kernels"].items()})
    except Exception as e:
        print("c3 line unreadable", e)
PY
tail -2 gpurun_out/r2s_c3_n1.err
