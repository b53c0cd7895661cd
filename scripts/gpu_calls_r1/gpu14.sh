mkdir -p gpurun_out/golden
timeout 900 python oracle/make_golden.py --out gpurun_out/golden > gpurun_out/r1e_golden.log 2>&1; echo "golden exit $?"
timeout 600 python oracle/make_golden_fusion.py --out gpurun_out/golden > gpurun_out/r1e_golden_fusion.log 2>&1; echo "golden fusion exit $?"
grep -c "ref~cuda" gpurun_out/golden/parity_table.txt gpurun_out/golden/parity_table_fusion.txt
grep -E "FAILED" gpurun_out/r1e_golden.log gpurun_out/r1e_golden_fusion.log | head
