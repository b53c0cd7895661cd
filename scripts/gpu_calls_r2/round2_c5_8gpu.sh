# Config c5 (2048^3 x 180 on 8 GPUs).  An 8-GPU call is charged 8x: the whole script is bounded to ~6 minutes
# (bench.py's own watchdog 240 s, `timeout -k` behind it) -- the round-1 attempt hung after a rank failed and
# burnt the round's budget.  Run with: gpurun --gpus 8 --timeout 420 -- 'bash scripts/round2_c5_8gpu.sh'
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout -k 5 300 $TR --nproc-per-node 8 --master-port 29524 bench.py --gpus 8 --size 2048 --angles 180 --steps 5 --warmup 3 \
    --phantom-period 16 --no-cpu --max-seconds 240 > gpurun_out/r2_bench_c5_n8.json 2> gpurun_out/r2_bench_c5_n8.err; echo "bench c5 n8 exit $?"
cut -c1-400 gpurun_out/r2_bench_c5_n8.json; tail -n 4 gpurun_out/r2_bench_c5_n8.err | cut -c1-300
