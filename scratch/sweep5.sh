#!/bin/bash
out=gpurun_out/r01zv_first_chunk_lanes5.txt
: > $out
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -2 >> $out
for f in 32 48 64 96; do for rep in 1 2; do
  v=$(FS_TUNE_E_FIRST_CHUNK=$f timeout 200 python bench.py --steps 10 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(round(d['value']), round(d['roofline']['stage_ms_per_step']['essential_ransac'],2))")
  echo "E first=$f -> $v" >> $out
done; done
cat $out
