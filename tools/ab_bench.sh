#!/bin/bash
# A/B runs of bench.py over the trace-kernel tuning knobs (run on the GPU box via gpurun).
# usage: tools/ab_bench.sh "VAR=val VAR2=val" "VAR=val" ...   -> gpurun_out/ab.log
mkdir -p gpurun_out
: > gpurun_out/ab.log
for cfg in "$@"; do
  echo "== $cfg" >> gpurun_out/ab.log
  env $cfg python bench.py --steps 3 --warmup 3 --no-cpu-baseline 2>>gpurun_out/ab.err | python -c "
import sys, json
for l in sys.stdin:
    l = l.strip()
    if not l.startswith('{'): continue
    j = json.loads(l)
    print(json.dumps({'value': round(j['value'], 1), 'ms': round(j['ms_per_step'], 2), 'e2e': round(j['e2e']['value'], 1), 'frac': round(j['roofline']['frac'], 3), 'trace_ms': round(j['roofline']['avg_launch_ms'], 3), 'share': {k: round(v, 3) for k, v in j['roofline']['share_of_step'].items()}}))
" >> gpurun_out/ab.log
done
cat gpurun_out/ab.log
