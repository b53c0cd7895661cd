mkdir -p gpurun_out
timeout -k 10 900 python -m pytest tests -m gpu -q --timeout 300 > gpurun_out/r2p_pytest.log 2>&1; echo "pytest exit $?"
tail -n 8 gpurun_out/r2p_pytest.log | cut -c1-400
grep -E "^E  " gpurun_out/r2p_pytest.log | cut -c1-200 | head -10
timeout -k 10 600 python scripts/bench_algos.py --steps 3 --only sart_pass,asd_pocs > gpurun_out/r2p_bench_algos_512.jsonl 2> gpurun_out/r2p_bench_algos_512.err; echo "bench_algos 512 exit $?"
timeout -k 10 600 python scripts/bench_algos.py --size 256 --angles 71 --steps 5 --only sart_pass,asd_pocs > gpurun_out/r2p_bench_algos_256.jsonl 2> gpurun_out/r2p_bench_algos_256.err; echo "bench_algos 256 exit $?"
python - <<'PY'
import json
for f in ("gpurun_out/r2p_bench_algos_512.jsonl", "gpurun_out/r2p_bench_algos_256.jsonl"):
    for l in open(f):
        d = json.loads(l)
        print(d["size"], d["workload"], round(d["ms_per_step"], 3), "ms", round(d["steps_per_s"], 2), "/s", {k: round(v["ms_per_launch"] * v["launches_per_step"], 3) for k, v in d["kernels"].items()})
PY
timeout -k 5 200 python scripts/kernel_times.py 512 512 180 5 :tvgd_grad :tvgd_grad2 > gpurun_out/r2p_times.log 2>&1; tail -3 gpurun_out/r2p_times.log
timeout -k 10 600 ncu --set full --clock-control none --import-source on -k regex:sartc_kernel -c 1 -o gpurun_out/r2p_sartc python scripts/bench_algos.py --size 512 --angles 180 --steps 1 --only sart_pass > gpurun_out/r2p_ncu.log 2>&1; echo "ncu exit $?"
ncu -i gpurun_out/r2p_sartc.ncu-rep --page details --csv > gpurun_out/r2p_sartc_details.csv 2>/dev/null
ncu -i gpurun_out/r2p_sartc.ncu-rep --page source --csv > gpurun_out/r2p_sartc_src.csv 2>/dev/null
rm -f gpurun_out/r2p_sartc.ncu-rep
