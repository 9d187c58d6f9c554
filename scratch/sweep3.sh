#!/bin/bash
out=gpurun_out/r01zj_pnp_smem.txt
: > $out
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -2 >> $out
for w in 2 4; do for f in 128 192 256; do
  v=$(FS_TUNE_P_SOLVE_WARPS=$w FS_TUNE_P_FIRST_CHUNK=$f timeout 200 python bench.py --workload stereo --steps 6 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(round(d['value']), round(d['roofline']['stage_ms_per_step']['pnp_ransac'],2))")
  echo "P warps=$w first=$f -> $v" >> $out
done; done
cat $out
