#!/bin/bash
# A/B of library variants on killeroo (bench.py) and C4 (bench.py --workload c4); run on the GPU box.
mkdir -p gpurun_out
: > gpurun_out/abc4.log
for cfg in "$@"; do
  echo "== $cfg" >> gpurun_out/abc4.log
  for wl in killeroo c4; do
    env $cfg python bench.py --workload $wl --steps 3 --warmup 3 --no-cpu-baseline 2>>gpurun_out/abc4.err | python -c "
import sys, json
for l in sys.stdin:
    l = l.strip()
    if not l.startswith('{'): continue
    j = json.loads(l)
    print('$wl', round(j['value'], 1), 'Msamples/s', round(j['ms_per_step'], 2), 'ms  trace launch', round(j['roofline']['avg_launch_ms'], 3), 'ms')
" >> gpurun_out/abc4.log
  done
done
cat gpurun_out/abc4.log
