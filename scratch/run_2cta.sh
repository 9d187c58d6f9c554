set -x
timeout 120 python scratch/ncu_top3.py 0 1 2 3 > gpurun_out/sp_probe.log 2>&1; echo "probe rc=$?" >> gpurun_out/sp_probe.log
cat gpurun_out/sp_probe.log | tail -n 6
timeout 300 python -m pytest tests/test_match_gpu.py -x -q > gpurun_out/sp_match.log 2>&1; echo "match rc=$?" >> gpurun_out/sp_match.log
tail -n 5 gpurun_out/sp_match.log
