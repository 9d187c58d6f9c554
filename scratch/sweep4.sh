#!/bin/bash
out=gpurun_out/r01zr_pnp_warps5.txt
: > $out
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -2 >> $out
for w in 4 5; do for f in 128 160 192 224; do
  v=$(FS_TUNE_P_SOLVE_WARPS=$w FS_TUNE_P_FIRST_CHUNK=$f timeout 200 python bench.py --workload stereo --steps 8 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(round(d['value']), round(d['roofline']['stage_ms_per_step']['pnp_ransac'],2))")
  echo "P(32-thread CTAs) warps=$w first=$f -> $v" >> $out
done; done
cat $out
