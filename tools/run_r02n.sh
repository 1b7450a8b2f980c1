#!/bin/bash
# r02n: refill threshold / phase bias sweep with the claim-ahead refill (variant `ahead`) and without
mkdir -p gpurun_out
V=pbrt-v3-light-portals_b200/lib/variants
: > gpurun_out/r02n_ab.log
run() {
  echo "== $*" >> gpurun_out/r02n_ab.log
  env "$@" python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-other-configs 2>>gpurun_out/r02n_ab.err | python -c "
import sys, json
for l in sys.stdin:
    l = l.strip()
    if not l.startswith('{'): continue
    j = json.loads(l)
    print('killeroo', round(j['value'], 1), 'Msamples/s', round(j['ms_per_step'], 2), 'ms  trace launch', round(j['roofline']['avg_launch_ms'], 3), 'ms', {k: round(v, 2) for k, v in j['roofline']['kernel_ms_in_profile_step'].items()})
" >> gpurun_out/r02n_ab.log
}
A=PBRTGPU_LIBRARY=$PWD/$V/libpbrtgpu_ahead.so
for t in 4 6 8 10 12 16; do run $A PBRTGPU_FETCH_THRESHOLD=$t; done
for t in 8 12; do run PBRTGPU_FETCH_THRESHOLD=$t; done
for b in 10 13 20 24; do run $A PBRTGPU_PHASE_BIAS=$b; done
cat gpurun_out/r02n_ab.log
