set -x
timeout 1200 python bench_kernels.py > gpurun_out/r02p_bench_kernels.jsonl 2> gpurun_out/r02p_bench_kernels.err; echo "bk rc=$?"
tail -n 3 gpurun_out/r02p_bench_kernels.err
