python bench.py --workload stereo --steps 1 --warmup 1 --headline-only --no-cpu-baseline --no-verify > gpurun_out/plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_rg_solve -s 0 -c 2 -o gpurun_out/r02x_pnp_solve python bench.py --workload stereo --steps 1 --warmup 1 --headline-only --no-cpu-baseline --no-verify > gpurun_out/ncu.log 2>&1
echo rc=$?
