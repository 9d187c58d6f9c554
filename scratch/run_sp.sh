set -x
timeout 600 python -m pytest tests/test_match_gpu.py -x -q > gpurun_out/sp_match.log 2>&1; echo "match rc=$?" >> gpurun_out/sp_match.log
timeout 600 python bench.py --headline-only > gpurun_out/sp_bench.json 2> gpurun_out/sp_bench.err; echo "bench rc=$?"
FS_TUNE_FLOAT_SINGLE_PASS=0 timeout 600 python bench.py --headline-only --no-verify --no-cpu-baseline > gpurun_out/sp_bench_off.json 2> gpurun_out/sp_bench_off.err; echo "bench rc=$?"
tail -n 5 gpurun_out/sp_match.log
