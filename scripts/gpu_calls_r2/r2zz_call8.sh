# final 8-GPU pass: headline bench at N = 1, 2, 4, 8 on the same box (equal recon / tv_fgp hashes), config c3 at N = 8
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout -k 5 200 python bench.py --steps 20 --warmup 3 --no-cpu --max-seconds 150 > gpurun_out/r2zz_bench512_n1.json 2> gpurun_out/r2zz_bench512_n1.err; echo "bench n1 exit $?"
for n in 2 4 8; do
timeout -k 5 200 $TR --nproc-per-node $n --master-port 2953$n bench.py --gpus $n --steps 20 --warmup 3 --no-cpu --max-seconds 150 > gpurun_out/r2zz_bench512_n$n.json 2> gpurun_out/r2zz_bench512_n$n.err; echo "bench n$n exit $?"
done
timeout -k 5 200 $TR --nproc-per-node 8 --master-port 29525 scripts/bench_c3.py > gpurun_out/r2zz_c3_n8.jsonl 2> gpurun_out/r2zz_c3_n8.err; echo "c3 n8 exit $?"
python - <<'PY'
import json
for n in (1, 2, 4, 8):
    f = f"gpurun_out/r2zz_bench512_n{n}.json"
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        print(n, round(d["value"], 2), "it/s  e2e", round(d["e2e"]["value"], 2), d["recon_sha256"][:12], d["tv_fgp_sha256"][:12],
              {k: round(v["ms_per_launch"] * v["launches_per_step"], 3) for k, v in d["kernels"].items()})
    except Exception as e:
        print(f, "unreadable", e)
for l in open("gpurun_out/r2zz_c3_n8.jsonl"):
    try:
        d = json.loads(l); print(d["workload"], d["n_gpus"], round(d["ms_per_step"], 3), "ms", d["recon_sha256"][:16], {k: round(v["ms_per_step"], 3) for k, v in d["kernels"].items()})
    except Exception as e:
        print("c3 line unreadable", e)
PY
