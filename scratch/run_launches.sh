set -x
python bench.py --steps 2 --warmup 1 --headline-only --no-cpu-baseline --no-verify > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/r02n_launches_mono512.csv python bench.py --steps 2 --warmup 1 --headline-only --no-cpu-baseline --no-verify > gpurun_out/ncu.log 2>&1
tail -n 2 gpurun_out/plain.log | cut -c1-300
