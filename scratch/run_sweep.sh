for C in 24 32; do FS_TUNE_E_FIRST_CHUNK=$C timeout 300 python bench.py --headline-only --no-cpu-baseline --no-verify > gpurun_out/sw_e$C.json 2>/dev/null; done
for C in 96 128 160 256; do FS_TUNE_P_FIRST_CHUNK=$C timeout 300 python bench.py --workload stereo --headline-only --no-cpu-baseline --no-verify > gpurun_out/sw_p$C.json 2>/dev/null; done
timeout 300 python bench.py --workload stereo --headline-only --no-cpu-baseline --no-verify > gpurun_out/sw_p192.json 2>/dev/null
echo done
