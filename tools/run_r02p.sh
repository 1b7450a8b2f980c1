#!/bin/bash
# r02p: two rays per lane (variant `dual`) against the default library
mkdir -p gpurun_out /tmp/prof
V=pbrt-v3-light-portals_b200/lib/variants
: > gpurun_out/r02p_ab.log
run() {
  echo "== $*" >> gpurun_out/r02p_ab.log
  env "$@" python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-other-configs 2>>gpurun_out/r02p_ab.err | python -c "
import sys, json
for l in sys.stdin:
    l = l.strip()
    if not l.startswith('{'): continue
    j = json.loads(l)
    print('killeroo', round(j['value'], 1), 'Msamples/s', round(j['ms_per_step'], 2), 'ms  trace launch', round(j['roofline']['avg_launch_ms'], 3), 'ms', {k: round(v, 2) for k, v in j['roofline']['kernel_ms_in_profile_step'].items()})
" >> gpurun_out/r02p_ab.log
  env "$@" python tools/bench_configs.py "c1 cornell_dielectric path 512" "strategy light" 2>>gpurun_out/r02p_ab.err | python -c "
import sys, json
for l in sys.stdin:
    l = l.strip()
    if not l.startswith('{'): continue
    j = json.loads(l)
    print(j['config'][:44].ljust(44), j['Msamples_per_s'], 'Msamples/s', j['kernel_ms'])
" >> gpurun_out/r02p_ab.log
}
D=PBRTGPU_LIBRARY=$PWD/$V/libpbrtgpu_dual.so
PBRTGPU_LIBRARY=$PWD/$V/libpbrtgpu_dual.so timeout 600 python -m pytest tests -m gpu -x -q -k "rays or oracle or golden or killeroo or no_alloc" > gpurun_out/r02p_tests_dual.log 2>&1; tail -3 gpurun_out/r02p_tests_dual.log
run PBRTGPU_NOP=1
run $D
run $D PBRTGPU_FETCH_THRESHOLD=6
run $D PBRTGPU_FETCH_THRESHOLD=20
run $D PBRTGPU_PHASE_BIAS=24
run $D PBRTGPU_PERSIST_BLOCKS=6
cat gpurun_out/r02p_ab.log
B="python bench.py --steps 1 --warmup 1 --no-other-configs --no-cpu-baseline"
PBRTGPU_LIBRARY=$PWD/$V/libpbrtgpu_dual.so ncu --set full --clock-control none --import-source on -k regex:"k_ptrace_fused" -s 3 -c 1 -f -o /tmp/prof/r02p_trace_src $B > gpurun_out/r02p_ncu_trace.log 2>&1
python tools/ncu_summary.py /tmp/prof/r02p_trace_src.ncu-rep > gpurun_out/r02p_trace_src_summary.txt 2>&1
ncu -i /tmp/prof/r02p_trace_src.ncu-rep --page source --csv > gpurun_out/r02p_trace_src_source.csv 2>/dev/null
cat gpurun_out/r02p_trace_src_summary.txt
