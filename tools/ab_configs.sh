#!/bin/bash
# A/B of library variants over bench.py (killeroo) and tools/bench_configs.py (Cornell, portal room); run on the GPU box.
# usage: tools/ab_configs.sh "VAR=val ..." "VAR=val ..."   -> gpurun_out/abc.log
mkdir -p gpurun_out
: > gpurun_out/abc.log
for cfg in "$@"; do
  echo "== $cfg" >> gpurun_out/abc.log
  env $cfg python bench.py --steps 3 --warmup 3 --no-cpu-baseline 2>>gpurun_out/abc.err | python -c "
import sys, json
for l in sys.stdin:
    l = l.strip()
    if not l.startswith('{'): continue
    j = json.loads(l)
    print('killeroo', round(j['value'], 1), 'Msamples/s', round(j['ms_per_step'], 2), 'ms  trace launch', round(j['roofline']['avg_launch_ms'], 3), 'ms')
" >> gpurun_out/abc.log
  env $cfg python tools/bench_configs.py 2>>gpurun_out/abc.err | python -c "
import sys, json
for l in sys.stdin:
    l = l.strip()
    if not l.startswith('{'): continue
    j = json.loads(l)
    print(j['config'][:44].ljust(44), j['Msamples_per_s'], 'Msamples/s', j['kernel_ms'])
" >> gpurun_out/abc.log
done
cat gpurun_out/abc.log
