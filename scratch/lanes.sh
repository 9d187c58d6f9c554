#!/bin/bash
out=gpurun_out/r01zo_lanes.txt
: > $out
for l in 3 4 5 6; do
  v=$(timeout 200 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --lanes $l 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(round(d['value']), round(d['e2e']['value']))")
  echo "mono lanes=$l -> $v" >> $out
  v=$(timeout 200 python bench.py --workload stereo --steps 10 --warmup 3 --no-cpu-baseline --lanes $l 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(round(d['value']), round(d['e2e']['value']))")
  echo "stereo lanes=$l -> $v" >> $out
done
cat $out
