for L in 8 12; do for C in 16 24 32; do FS_TUNE_E_FIRST_CHUNK=$C timeout 300 python bench.py --headline-only --no-cpu-baseline --no-verify --lanes $L --steps 36 > gpurun_out/sw3_l${L}_c$C.json 2>/dev/null; done; done
echo done
