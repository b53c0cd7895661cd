# 8-GPU call (charged 8x): multi-rank correctness on 8 GPUs, headline bench at N=8 (hash must equal N=1), config c5 (2048^3), config c3
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout -k 5 200 $TR --nproc-per-node 8 --master-port 29521 scripts/mgpu_check.py 300 96 24 > gpurun_out/r2r_mgpu8.log 2>&1; echo "mgpu_check8 exit $?"
grep -E "MGPU" gpurun_out/r2r_mgpu8.log | cut -c1-120
timeout -k 5 200 $TR --nproc-per-node 8 --master-port 29522 bench.py --gpus 8 --steps 20 --warmup 3 --no-cpu --max-seconds 150 > gpurun_out/r2r_bench512_n8.json 2> gpurun_out/r2r_bench512_n8.err; echo "bench n8 exit $?"
timeout -k 5 200 $TR --nproc-per-node 4 --master-port 29523 bench.py --gpus 4 --steps 20 --warmup 3 --no-cpu --max-seconds 150 > gpurun_out/r2r_bench512_n4.json 2> gpurun_out/r2r_bench512_n4.err; echo "bench n4 exit $?"
timeout -k 5 330 $TR --nproc-per-node 8 --master-port 29524 bench.py --gpus 8 --size 2048 --angles 180 --steps 5 --warmup 3 --phantom-period 16 --no-cpu --max-seconds 300 > gpurun_out/r2r_bench_c5_n8.json 2> gpurun_out/r2r_bench_c5_n8.err; echo "bench c5 n8 exit $?"
timeout -k 5 200 $TR --nproc-per-node 8 --master-port 29525 scripts/bench_c3.py > gpurun_out/r2r_c3_n8.jsonl 2> gpurun_out/r2r_c3_n8.err; echo "c3 n8 exit $?"
python - <<'PY'
import json
for f in ("gpurun_out/r2r_bench512_n8.json", "gpurun_out/r2r_bench512_n4.json", "gpurun_out/r2r_bench_c5_n8.json"):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        print(f, d["n_gpus"], round(d["value"], 2), "it/s  e2e", round(d["e2e"]["value"], 2), d["recon_sha256"][:16], d["recon_tv"],
              {k: round(v["ms_per_launch"] * v["launches_per_step"], 3) for k, v in d["kernels"].items()})
    except Exception as e:
        print(f, "unreadable", e)
for l in open("gpurun_out/r2r_c3_n8.jsonl"):
    try:
        d = json.loads(l); print(d["workload"], d["n_gpus"], round(d["ms_per_step"], 3), "ms", d["recon_sha256"][:16], {k: round(v["ms_per_step"], 3) for k, v in d["kernels"].items()})
    except Exception as e:
        print("c3 line unreadable", e)
PY
tail -n 3 gpurun_out/r2r_bench_c5_n8.err | cut -c1-300
