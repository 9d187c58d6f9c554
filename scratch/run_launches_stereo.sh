python bench.py --workload stereo --steps 2 --warmup 1 --headline-only --no-cpu-baseline --no-verify > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/r02y_launches_stereo512.csv python bench.py --workload stereo --steps 2 --warmup 1 --headline-only --no-cpu-baseline --no-verify > gpurun_out/ncu.log 2>&1
echo rc=$?
