#!/bin/bash
# measurement sweep of the RANSAC engine's per-policy knobs (FS_TUNE_*); prints one line per setting
out=gpurun_out/r01zh_ransac_tuning.txt
: > $out
for w in 4 6 8; do for f in 64 128 192 256; do
  v=$(FS_TUNE_P_SOLVE_WARPS=$w FS_TUNE_P_FIRST_CHUNK=$f timeout 200 python bench.py --workload stereo --steps 6 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(round(d['value']), round(d['roofline']['stage_ms_per_step']['pnp_ransac'],2))")
  echo "P warps=$w first=$f -> $v" >> $out
done; done
for w in 4 8; do for f in 32 64 128; do
  v=$(FS_TUNE_E_SOLVE_WARPS=$w FS_TUNE_E_FIRST_CHUNK=$f timeout 200 python bench.py --steps 6 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(round(d['value']), round(d['roofline']['stage_ms_per_step']['essential_ransac'],2))")
  echo "E warps=$w first=$f -> $v" >> $out
done; done
cat $out
