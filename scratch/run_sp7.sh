set -x
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "all rc=$?" >> gpurun_out/pytest_gpu.log
tail -n 3 gpurun_out/pytest_gpu.log
timeout 300 python bench.py --headline-only --no-cpu-baseline > gpurun_out/sp_bench.json 2> gpurun_out/sp_bench.err; echo "bench rc=$?"
timeout 300 python bench.py --workload stereo --headline-only --no-cpu-baseline --no-verify > gpurun_out/sp_bench_stereo.json 2>/dev/null
python bench_kernels.py --only prepare > gpurun_out/tc_bk.jsonl 2>/dev/null
