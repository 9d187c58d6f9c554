set -x
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "all rc=$?" >> gpurun_out/pytest_gpu.log
timeout 600 python bench.py --workload float4096 --no-cpu-baseline > gpurun_out/sp_bench_f4096.json 2> gpurun_out/sp_bench_f4096.err; echo "bench rc=$?"
for L in 4 6 8; do timeout 300 python bench.py --headline-only --no-cpu-baseline --no-verify --lanes $L > gpurun_out/sp_lanes$L.json 2>/dev/null; done
for C in 40 56; do FS_TUNE_E_FIRST_CHUNK=$C timeout 300 python bench.py --headline-only --no-cpu-baseline --no-verify > gpurun_out/sp_chunk$C.json 2>/dev/null; done
tail -n 3 gpurun_out/pytest_gpu.log
