#!/bin/bash
out=gpurun_out/r01zx_p_first_chunk_lanes5.txt
: > $out
for f in 128 160 192 224 256; do
  v=$(FS_TUNE_P_FIRST_CHUNK=$f timeout 200 python bench.py --workload stereo --steps 10 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(round(d['value']), round(d['roofline']['stage_ms_per_step']['pnp_ransac'],2))")
  echo "P first=$f -> $v" >> $out
done
cat $out
