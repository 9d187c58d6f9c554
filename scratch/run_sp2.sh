set -x
timeout 900 python -m pytest tests/test_match_gpu.py -x -q > gpurun_out/sp_match.log 2>&1; echo "match rc=$?" >> gpurun_out/sp_match.log
python scratch/ncu_top3.py 0 1 2 3 4 5 6 > gpurun_out/sp_probe.log 2>&1
timeout 600 python bench.py --headline-only --no-cpu-baseline > gpurun_out/sp_bench.json 2> gpurun_out/sp_bench.err; echo "bench rc=$?"
tail -n 3 gpurun_out/sp_match.log; cat gpurun_out/sp_probe.log
