#!/bin/bash
# r02zb: wave buffers staggered inside their allocations (PBRTGPU_WAVE_STAGGER)
mkdir -p gpurun_out
: > gpurun_out/r02zb_ab.log
run() {
  echo "== $*" >> gpurun_out/r02zb_ab.log
  env "$@" python bench.py --steps 6 --warmup 3 --no-cpu-baseline --no-other-configs 2>>gpurun_out/r02zb_ab.err | python -c "
import sys, json
for l in sys.stdin:
    l = l.strip()
    if not l.startswith('{'): continue
    j = json.loads(l)
    print('killeroo', round(j['value'], 1), 'Msamples/s', round(j['ms_per_step'], 2), 'ms  e2e', round(j['e2e']['value'], 1), {k: round(v, 2) for k, v in j['roofline']['kernel_ms_in_profile_step'].items()})
" >> gpurun_out/r02zb_ab.log
}
for s in 0 4352 69632 1114112 0 4352 33024 557056; do run PBRTGPU_WAVE_STAGGER=$s; done
cat gpurun_out/r02zb_ab.log
