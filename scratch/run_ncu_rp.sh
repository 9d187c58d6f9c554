python bench.py --steps 1 --warmup 1 --headline-only --no-cpu-baseline --no-verify > gpurun_out/plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_rp_cheirality -s 1 -c 1 -o gpurun_out/r02za_rp_cheirality python bench.py --steps 1 --warmup 1 --headline-only --no-cpu-baseline --no-verify > gpurun_out/ncu.log 2>&1
echo rc=$?
