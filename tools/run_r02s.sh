#!/bin/bash
# r02s: hit queue ordered by the cell of the hit point (PBRTGPU_SORT_CELLS): image tests with it forced on, C4 with and without
mkdir -p gpurun_out
PBRTGPU_SORT_CELLS=1 timeout 900 python -m pytest tests -m gpu -x -q -k "reference_image or baseline or knobs or analytic or portal or shards" > gpurun_out/r02s_tests_sortcells.log 2>&1; tail -3 gpurun_out/r02s_tests_sortcells.log
: > gpurun_out/r02s_ab.log
for v in 0 1; do
  echo "== PBRTGPU_SORT_CELLS=$v" >> gpurun_out/r02s_ab.log
  PBRTGPU_SORT_CELLS=$v python bench.py --workload c4 --steps 2 --warmup 1 --no-cpu-baseline --no-other-configs 2>>gpurun_out/r02s_ab.err | python -c "
import sys, json
for l in sys.stdin:
    l = l.strip()
    if not l.startswith('{'): continue
    j = json.loads(l)
    print('c4', round(j['value'], 1), 'Msamples/s', round(j['ms_per_step'], 1), 'ms  e2e', round(j['e2e']['value'], 1), {k: round(v, 1) for k, v in j['roofline']['kernel_ms_in_profile_step'].items()})
" >> gpurun_out/r02s_ab.log
done
for v in 0 1; do
  echo "== killeroo PBRTGPU_SORT_CELLS=$v" >> gpurun_out/r02s_ab.log
  PBRTGPU_SORT_CELLS=$v python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-other-configs 2>>gpurun_out/r02s_ab.err | python -c "
import sys, json
for l in sys.stdin:
    l = l.strip()
    if not l.startswith('{'): continue
    j = json.loads(l)
    print('killeroo', round(j['value'], 1), 'Msamples/s', round(j['ms_per_step'], 2), 'ms', {k: round(v, 2) for k, v in j['roofline']['kernel_ms_in_profile_step'].items()})
" >> gpurun_out/r02s_ab.log
done
cat gpurun_out/r02s_ab.log
