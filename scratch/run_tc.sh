python -m pytest tests/test_match_gpu.py tests/test_sinkhorn_gpu.py -x -q > gpurun_out/tc_match.log 2>&1; echo "rc=$?" >> gpurun_out/tc_match.log
tail -n 4 gpurun_out/tc_match.log
python bench_kernels.py --only float > gpurun_out/tc_bk.jsonl 2>/dev/null; grep -o '"kernel": "[^"]*", "ms": [0-9.]*' gpurun_out/tc_bk.jsonl
