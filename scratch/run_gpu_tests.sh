set -x
python -m pytest tests/test_zz_no_alloc.py -x -q > gpurun_out/zz_alone.log 2>&1; echo "alone rc=$?" >> gpurun_out/zz_alone.log
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "all rc=$?" >> gpurun_out/pytest_gpu.log
python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "bench rc=$?"
tail -3 gpurun_out/zz_alone.log gpurun_out/pytest_gpu.log
