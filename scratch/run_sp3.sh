set -x
timeout 900 python -m pytest tests/test_match_gpu.py -x -q > gpurun_out/sp_match.log 2>&1; echo "match rc=$?" >> gpurun_out/sp_match.log
timeout 600 python bench.py --headline-only --no-cpu-baseline > gpurun_out/sp_bench.json 2> gpurun_out/sp_bench.err; echo "bench rc=$?"
timeout 600 python bench.py --workload float4096 --no-cpu-baseline > gpurun_out/sp_bench_f4096.json 2> gpurun_out/sp_bench_f4096.err; echo "bench rc=$?"
tail -n 3 gpurun_out/sp_match.log
