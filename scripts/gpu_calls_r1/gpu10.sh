mkdir -p gpurun_out
NG=${NG:-8}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1"
timeout 300 $TR --master-port 29511 scripts/mgpu_check.py 100 96 20 > gpurun_out/mgpu_check_$NG.log 2>&1; echo "mgpu exit $?" >> gpurun_out/mgpu_check_$NG.log
timeout 600 $TR --master-port 29512 bench.py --gpus $NG --steps 20 --warmup 3 > gpurun_out/bench_${NG}gpu.log 2>&1; echo "bench exit $?" >> gpurun_out/bench_${NG}gpu.log
grep -E "FAIL|PASS|exit|rror" gpurun_out/mgpu_check_$NG.log | head; tail -n 2 gpurun_out/bench_${NG}gpu.log | cut -c1-250
