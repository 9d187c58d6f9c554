set -x
timeout 120 python scratch/ncu_top3.py 0 1 2 3 > gpurun_out/sp_probe.log 2>&1; tail -n 4 gpurun_out/sp_probe.log
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "all rc=$?" >> gpurun_out/pytest_gpu.log
timeout 900 python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "bench rc=$?"
tail -n 3 gpurun_out/pytest_gpu.log
