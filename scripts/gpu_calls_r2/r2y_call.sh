mkdir -p gpurun_out
timeout -k 10 600 ncu --set full --clock-control none --import-source on -k regex:tvgd_gradient_march -c 1 -o gpurun_out/r2y_tvgd python scripts/kprobe.py 512 512 180 1 tvgd_grad2 > gpurun_out/r2y_ncu.log 2>&1; echo "ncu exit $?"
ncu -i gpurun_out/r2y_tvgd.ncu-rep --page details --csv > gpurun_out/r2y_tvgd_details.csv 2>/dev/null
ncu -i gpurun_out/r2y_tvgd.ncu-rep --page source --csv > gpurun_out/r2y_tvgd_src.csv 2>/dev/null
rm -f gpurun_out/r2y_tvgd.ncu-rep
