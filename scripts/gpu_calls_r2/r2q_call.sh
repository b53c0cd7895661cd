# 2-GPU call: multi-rank tests (engine, API, fusion), bench at N=1 and N=2 with the reconstruction hash, c3 workload at N=2
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout -k 10 900 python -m pytest tests/test_gpu_fullsize.py -m gpu -q --timeout 600 > gpurun_out/r2q_pytest_fullsize.log 2>&1; echo "pytest fullsize exit $?"
tail -n 6 gpurun_out/r2q_pytest_fullsize.log | cut -c1-1500
timeout -k 5 300 python bench.py --steps 20 --warmup 3 --no-cpu --max-seconds 250 > gpurun_out/r2q_bench512_n1.json 2> gpurun_out/r2q_bench512_n1.err; echo "bench n1 exit $?"
timeout -k 5 300 $TR --nproc-per-node 2 --master-port 29512 bench.py --gpus 2 --steps 40 --warmup 3 --no-cpu --max-seconds 250 > gpurun_out/r2q_bench512_n2.json 2> gpurun_out/r2q_bench512_n2.err; echo "bench n2 exit $?"
python - <<'PY'
import json
for f in ("gpurun_out/r2q_bench512_n1.json", "gpurun_out/r2q_bench512_n2.json"):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        print(f, d["n_gpus"], round(d["value"], 2), "it/s  e2e", round(d["e2e"]["value"], 2), d["recon_sha256"][:16], d["recon_tv"],
              {k: round(v["ms_per_launch"] * v["launches_per_step"], 3) for k, v in d["kernels"].items()})
    except Exception as e:
        print(f, "unreadable", e)
PY
tail -3 gpurun_out/r2q_bench512_n2.err | cut -c1-300
