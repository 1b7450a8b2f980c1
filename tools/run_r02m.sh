#!/bin/bash
# r02m on the GPU box: A/B of trace-kernel variants (tools/build_variants.py): next-record prefetch, 64-bit top of stack,
# queue positions claimed ahead with their rays prefetched.
mkdir -p gpurun_out
V=pbrt-v3-light-portals_b200/lib/variants
: > gpurun_out/r02m_ab.log
for name in base pf pop2 pop2pf ahead aheadpf all3; do
  lib=$PWD/$V/libpbrtgpu_$name.so
  [ "$name" = base ] && lib=$PWD/pbrt-v3-light-portals_b200/lib/libpbrtgpu.so
  echo "== $name" >> gpurun_out/r02m_ab.log
  PBRTGPU_LIBRARY=$lib python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-other-configs 2>>gpurun_out/r02m_ab.err | python -c "
import sys, json
for l in sys.stdin:
    l = l.strip()
    if not l.startswith('{'): continue
    j = json.loads(l)
    print('killeroo', round(j['value'], 1), 'Msamples/s', round(j['ms_per_step'], 2), 'ms  trace launch', round(j['roofline']['avg_launch_ms'], 3), 'ms', {k: round(v, 2) for k, v in j['roofline']['kernel_ms_in_profile_step'].items()})
" >> gpurun_out/r02m_ab.log
  PBRTGPU_LIBRARY=$lib python tools/bench_configs.py "c1 cornell_dielectric path 512" "strategy light" 2>>gpurun_out/r02m_ab.err | python -c "
import sys, json
for l in sys.stdin:
    l = l.strip()
    if not l.startswith('{'): continue
    j = json.loads(l)
    print(j['config'][:44].ljust(44), j['Msamples_per_s'], 'Msamples/s', j['kernel_ms'])
" >> gpurun_out/r02m_ab.log
done
cat gpurun_out/r02m_ab.log
PBRTGPU_LIBRARY=$PWD/$V/libpbrtgpu_all3.so python -m pytest tests -m gpu -x -q -k "rays or oracle or golden or killeroo or no_alloc" > gpurun_out/r02m_tests_all3.log 2>&1; tail -2 gpurun_out/r02m_tests_all3.log
