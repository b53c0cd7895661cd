mkdir -p gpurun_out
for i in 1 2 3 4 5 6; do
timeout 100 python -m pytest tests -m gpu -q --timeout 90 -k "peer_memory" > gpurun_out/p2p_test_$i.log 2>&1; echo "run $i exit $?"; grep -E "passed|failed" gpurun_out/p2p_test_$i.log | tail -1; grep -oE "HDR[^\"]*" gpurun_out/p2p_test_$i.log | sort | uniq -c | cut -c1-330
done
