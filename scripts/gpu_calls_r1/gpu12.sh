mkdir -p gpurun_out/golden
timeout 600 python oracle/make_golden_fusion.py --out gpurun_out/golden > gpurun_out/r1c_golden_fusion.log 2>&1; echo "golden exit $?"
tail -n 40 gpurun_out/r1c_golden_fusion.log
timeout 600 python -m pytest tests/test_gpu_fusion.py -m gpu -x -q --timeout 300 > gpurun_out/r1c_pytest_fusion.log 2>&1; echo "pytest exit $?"
tail -n 15 gpurun_out/r1c_pytest_fusion.log
timeout 600 python -m pytest tests -m gpu -x -q --timeout 300 --deselect tests/test_gpu_fusion.py > gpurun_out/r1c_pytest_gpu.log 2>&1; echo "pytest all exit $?"
tail -n 3 gpurun_out/r1c_pytest_gpu.log
