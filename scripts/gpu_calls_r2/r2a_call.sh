# Round-2 first GPU call (ONE GPU): validate state after round 1's last edits, time secondary workloads, run experimental variants.
mkdir -p gpurun_out
timeout -k 10 900 python -m pytest tests -m gpu -x -q --timeout 600 > gpurun_out/r2a_pytest_gpu.log 2>&1; echo "pytest exit $?"
tail -n 5 gpurun_out/r2a_pytest_gpu.log | cut -c1-300
timeout -k 10 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2a_smoke.log 2>&1; echo "smoke exit $?"; tail -n 2 gpurun_out/r2a_smoke.log
timeout -k 10 600 python bench.py --steps 20 --warmup 3 > gpurun_out/r2a_bench512.json 2> gpurun_out/r2a_bench512.err; echo "bench exit $?"
cut -c1-400 gpurun_out/r2a_bench512.json
timeout -k 10 600 python scripts/bench_algos.py --steps 3 > gpurun_out/r2a_bench_algos.jsonl 2> gpurun_out/r2a_bench_algos.err; echo "bench_algos exit $?"
cut -c1-200 gpurun_out/r2a_bench_algos.jsonl
for v in fp1 fp8 bp5 bp8 bp9; do
  TTV_EXPERIMENTAL=1 timeout -k 10 300 python -m pytest tests/test_gpu_experimental.py -m gpu -q --timeout 200 -k $v > gpurun_out/r2a_pytest_exp_$v.log 2>&1; echo "experimental $v exit $?"
  tail -n 4 gpurun_out/r2a_pytest_exp_$v.log | cut -c1-300
done
timeout -k 10 900 python scripts/variants_probe.py > gpurun_out/r2a_variants.log 2>&1; cat gpurun_out/r2a_variants.log
