timeout 600 python -m pytest tests/test_match_gpu.py -x -q > gpurun_out/sp_match.log 2>&1; echo "match rc=$?" >> gpurun_out/sp_match.log
tail -n 25 gpurun_out/sp_match.log
