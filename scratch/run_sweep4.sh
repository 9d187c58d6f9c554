for C in 128 256 320; do FS_TUNE_P_FIRST_CHUNK=$C timeout 300 python bench.py --workload stereo --headline-only --no-cpu-baseline --no-verify --steps 20 > gpurun_out/sw4_c$C.json 2>/dev/null; done
for W in 4 5; do FS_TUNE_P_SOLVE_WARPS=$W timeout 300 python bench.py --workload stereo --headline-only --no-cpu-baseline --no-verify --steps 20 > gpurun_out/sw4_w$W.json 2>/dev/null; done
for L in 3 8; do timeout 300 python bench.py --workload stereo --headline-only --no-cpu-baseline --no-verify --steps 20 --lanes $L > gpurun_out/sw4_l$L.json 2>/dev/null; done
echo done
