#!/bin/bash
# r02za: evict-first LOADS only on the wavefront arrays (stores plain), against no hints and against loads + stores
mkdir -p gpurun_out
V=pbrt-v3-light-portals_b200/lib/variants
: > gpurun_out/r02za_ab.log
run() {
  echo "== $*" >> gpurun_out/r02za_ab.log
  env "$@" python bench.py --steps 6 --warmup 3 --no-cpu-baseline --no-other-configs 2>>gpurun_out/r02za_ab.err | python -c "
import sys, json
for l in sys.stdin:
    l = l.strip()
    if not l.startswith('{'): continue
    j = json.loads(l)
    print('killeroo', round(j['value'], 1), 'Msamples/s', round(j['ms_per_step'], 2), 'ms  e2e', round(j['e2e']['value'], 1), {k: round(v, 2) for k, v in j['roofline']['kernel_ms_in_profile_step'].items()})
" >> gpurun_out/r02za_ab.log
  env "$@" python tools/bench_configs.py "c1 cornell_dielectric path 512" "strategy light" 2>>gpurun_out/r02za_ab.err | python -c "
import sys, json
for l in sys.stdin:
    l = l.strip()
    if not l.startswith('{'): continue
    j = json.loads(l)
    print(j['config'][:44].ljust(44), j['Msamples_per_s'], 'Msamples/s', j['kernel_ms'])
" >> gpurun_out/r02za_ab.log
}
run PBRTGPU_LIBRARY=$PWD/$V/libpbrtgpu_prev.so
run PBRTGPU_NOP=1
run PBRTGPU_LIBRARY=$PWD/$V/libpbrtgpu_ldst.so
run PBRTGPU_LIBRARY=$PWD/$V/libpbrtgpu_prev.so
run PBRTGPU_NOP=1
cat gpurun_out/r02za_ab.log
