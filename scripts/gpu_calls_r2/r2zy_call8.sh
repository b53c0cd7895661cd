# after fpt_kernel<EPI,1>: headline bench at N = 8 again (and mgpu_check on 8 GPUs with the final tree)
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout -k 5 200 $TR --nproc-per-node 8 --master-port 29538 bench.py --gpus 8 --steps 20 --warmup 3 --no-cpu --max-seconds 150 > gpurun_out/r2zy_bench512_n8.json 2> gpurun_out/r2zy_bench512_n8.err; echo "bench n8 exit $?"
timeout -k 5 200 $TR --nproc-per-node 8 --master-port 29521 scripts/mgpu_check.py 300 96 24 > gpurun_out/r2zy_mgpu8.log 2>&1; echo "mgpu_check8 exit $?"
grep -E "MGPU" gpurun_out/r2zy_mgpu8.log | cut -c1-120
python - <<'PY'
import json
d = json.loads(open("gpurun_out/r2zy_bench512_n8.json").read().strip().splitlines()[-1])
print(8, round(d["value"], 2), "it/s  e2e", round(d["e2e"]["value"], 2), d["recon_sha256"][:12], d["tv_fgp_sha256"][:12],
      {k: round(v["ms_per_launch"] * v["launches_per_step"], 3) for k, v in d["kernels"].items()})
PY
