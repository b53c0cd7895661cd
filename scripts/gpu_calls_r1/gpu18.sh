# 8-GPU pass: 512^3 strong-scaling bench at N=8 and N=4, sharded parity at N=8, config c5 (2048^3 x 180) on 8 GPUs
mkdir -p gpurun_out
free -g | head -2
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 600 $TR --nproc-per-node 8 --master-port 29521 bench.py --gpus 8 --steps 20 --warmup 3 > gpurun_out/r1j_bench512_n8.json 2> gpurun_out/r1j_bench512_n8.err; echo "bench n8 exit $?"
cut -c1-330 gpurun_out/r1j_bench512_n8.json
timeout 600 $TR --nproc-per-node 4 --master-port 29522 bench.py --gpus 4 --steps 20 --warmup 3 > gpurun_out/r1j_bench512_n4.json 2> gpurun_out/r1j_bench512_n4.err; echo "bench n4 exit $?"
cut -c1-330 gpurun_out/r1j_bench512_n4.json
timeout 600 $TR --nproc-per-node 8 --master-port 29523 scripts/mgpu_check.py 300 96 20 > gpurun_out/r1j_mgpu8.log 2>&1; echo "mgpu_check n8 exit $?"
grep -E "MGPU|FAIL" gpurun_out/r1j_mgpu8.log | cut -c1-200
timeout 1200 $TR --nproc-per-node 8 --master-port 29524 bench.py --gpus 8 --size 2048 --angles 180 --steps 5 --warmup 3 --phantom-period 16 --no-cpu > gpurun_out/r1j_bench_c5_n8.json 2> gpurun_out/r1j_bench_c5_n8.err; echo "bench c5 n8 exit $?"
cut -c1-330 gpurun_out/r1j_bench_c5_n8.json; tail -n 4 gpurun_out/r1j_bench_c5_n8.err | cut -c1-300
