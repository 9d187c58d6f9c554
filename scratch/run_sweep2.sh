for W in 7 8; do for C in 32 40; do FS_TUNE_E_SOLVE_WARPS=$W FS_TUNE_E_FIRST_CHUNK=$C timeout 300 python bench.py --pairs 1024 --headline-only --no-cpu-baseline --no-verify > gpurun_out/sw2_w${W}_c$C.json 2>/dev/null; done; done
for W in 5 6; do FS_TUNE_E_SOLVE_WARPS=$W timeout 300 python bench.py --headline-only --no-cpu-baseline --no-verify > gpurun_out/sw2_512_w${W}.json 2>/dev/null; done
timeout 300 python bench.py --pairs 768 --headline-only --no-cpu-baseline --no-verify > gpurun_out/sw2_768.json 2>/dev/null
echo done
