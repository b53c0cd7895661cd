# 2-GPU pass: sharded-vs-whole parity (engine + TomoGPU on multigpuengine), 512^3 bench at N=2
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 scripts/mgpu_check.py > gpurun_out/r1i_mgpu2.log 2>&1; echo "mgpu_check exit $?"
grep -E "mgpu\[|MGPU" gpurun_out/r1i_mgpu2.log | cut -c1-200; tail -n 5 gpurun_out/r1i_mgpu2.log | cut -c1-300
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 scripts/mgpu_check.py 96 160 24 > gpurun_out/r1i_mgpu2b.log 2>&1; echo "mgpu_check(96,160,24) exit $?"
grep -E "MGPU" gpurun_out/r1i_mgpu2b.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus 2 --steps 20 --warmup 3 > gpurun_out/r1i_bench512_n2.json 2> gpurun_out/r1i_bench512_n2.err; echo "bench n2 exit $?"
cut -c1-400 gpurun_out/r1i_bench512_n2.json; tail -n 3 gpurun_out/r1i_bench512_n2.err
