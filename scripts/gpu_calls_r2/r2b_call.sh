mkdir -p gpurun_out
python scripts/ncu_variants.py > gpurun_out/r2b_plain.log 2>&1 && \
timeout -k 10 1200 ncu --set full --clock-control none --import-source on -k regex:'fp_kernel|fp8_kernel|bp_kernel|bp8_kernel' -o gpurun_out/r2b_variants python scripts/ncu_variants.py > gpurun_out/r2b_ncu.log 2>&1; echo "ncu exit $?"
cat gpurun_out/r2b_plain.log; tail -3 gpurun_out/r2b_ncu.log; ls -la gpurun_out/r2b_variants.ncu-rep
