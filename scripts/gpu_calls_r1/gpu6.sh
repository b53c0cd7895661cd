set -x
mkdir -p gpurun_out
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'fgp_tile' -s 1 -c 1 -o gpurun_out/r1_fgp_sweep python scripts/kprobe.py 512 512 180 2 fgp_sweep > gpurun_out/ncu_fgp.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'fgp_tile' -s 1 -c 1 -o gpurun_out/r1_fgp_first python scripts/kprobe.py 512 512 180 2 fgp_first >> gpurun_out/ncu_fgp.log 2>&1
tail -3 gpurun_out/ncu_fgp.log
