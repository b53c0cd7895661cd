#!/bin/bash
# scan of the split point n2 of the mixed forward-projector launch (fp_xb = 3, fp_mix_n2) at 512 / 256 / 128 / 96 / 64 slices
mkdir -p gpurun_out
L=gpurun_out/r2my_scan.log; : > $L
scan() { nx=$1; shift; args="fp_xb=2:fp_sirt fp_xb=1:fp_sirt fp_xb=3:fp_sirt"; for v in "$@"; do args="$args fp_mix_n2=$v:fp_sirt"; done
  echo "== $nx slices" >> $L; timeout 300 python scripts/kernel_times.py 512 $nx 180 10 $args >> $L 2>&1; }
scan 512 1036 1184 1258 1332 1369 1406 1443 1480 1508
scan 256 296 444 518 592 629 666 703 740
scan 128 74 111 148 185 222 259 296 333
scan 96 74 111 148 185 222 259 296 333
scan 64 37 74 111 130 148 170
cat $L
