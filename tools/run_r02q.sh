#!/bin/bash
# r02q: 64-byte queued rays (invDir and shear stored by the producer): refill threshold sweep, then the tier-1 / image tests
mkdir -p gpurun_out
: > gpurun_out/r02q_ab.log
run() {
  echo "== $*" >> gpurun_out/r02q_ab.log
  env "$@" python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-other-configs 2>>gpurun_out/r02q_ab.err | python -c "
import sys, json
for l in sys.stdin:
    l = l.strip()
    if not l.startswith('{'): continue
    j = json.loads(l)
    print('killeroo', round(j['value'], 1), 'Msamples/s', round(j['ms_per_step'], 2), 'ms  e2e', round(j['e2e']['value'], 1), ' trace launch', round(j['roofline']['avg_launch_ms'], 3), 'ms', {k: round(v, 2) for k, v in j['roofline']['kernel_ms_in_profile_step'].items()})
" >> gpurun_out/r02q_ab.log
}
for t in 12 8 6 4 3 2; do run PBRTGPU_FETCH_THRESHOLD=$t; done
cat gpurun_out/r02q_ab.log
for t in 12 4; do
PBRTGPU_FETCH_THRESHOLD=$t python tools/bench_configs.py "c1 cornell_dielectric path 512" "strategy light" 2>>gpurun_out/r02q_ab.err | python -c "
import sys, json
for l in sys.stdin:
    l = l.strip()
    if not l.startswith('{'): continue
    j = json.loads(l)
    print('thr $t', j['config'][:44].ljust(44), j['Msamples_per_s'], 'Msamples/s', j['kernel_ms'])
" | tee -a gpurun_out/r02q_ab.log
done
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r02q_tests.log 2>&1; tail -3 gpurun_out/r02q_tests.log
