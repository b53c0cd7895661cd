mkdir -p gpurun_out
timeout -k 10 600 ncu --set full --clock-control none --import-source on -k regex:fgp2_kernel -c 2 -o gpurun_out/r2n_fgp2 python scripts/kprobe.py 512 512 180 2 fgp2_sweep > gpurun_out/r2n_ncu.log 2>&1; echo "ncu exit $?"
ncu -i gpurun_out/r2n_fgp2.ncu-rep --page details --csv > gpurun_out/r2n_fgp2_details.csv 2>/dev/null
ncu -i gpurun_out/r2n_fgp2.ncu-rep --page source --csv > gpurun_out/r2n_fgp2_src.csv 2>/dev/null
ncu -i gpurun_out/r2n_fgp2.ncu-rep --page raw --csv > gpurun_out/r2n_fgp2_raw.csv 2>/dev/null
rm -f gpurun_out/r2n_fgp2.ncu-rep
tail -3 gpurun_out/r2n_ncu.log
