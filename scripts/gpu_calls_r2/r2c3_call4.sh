# config c3 (SIRT + ASD-POCS) at N = 2 and N = 4 on one 4-GPU box
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
for n in 2 4; do
timeout -k 5 200 $TR --nproc-per-node $n --master-port 2954$n scripts/bench_c3.py > gpurun_out/r2c3_n$n.jsonl 2> gpurun_out/r2c3_n$n.err; echo "c3 n$n exit $?"
done
python - <<'PY'
import json
for n in (2, 4):
    for l in open(f"gpurun_out/r2c3_n{n}.jsonl"):
        try:
            d = json.loads(l); print(d["workload"], d["n_gpus"], round(d["ms_per_step"], 3), "ms", d["recon_sha256"][:16], {k: round(v["ms_per_step"], 3) for k, v in d["kernels"].items()})
        except Exception as e:
            pass
PY
