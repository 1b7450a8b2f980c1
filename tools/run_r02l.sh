#!/bin/bash
# r02l on the GPU box: GPU tests, A/B of the three-run hit queue + all-specular shade instantiation and of the shared-memory
# scene staging (Cornell, portal room), a full bench line, source-level ncu of k_ptrace_fused.
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r02l_tests.log 2>&1; tail -3 gpurun_out/r02l_tests.log
: > gpurun_out/r02l_ab.log
for cfg in "PBRTGPU_SORT_HITS=1 PBRTGPU_TRACE_SHARED_SCENE=0" "PBRTGPU_SORT_HITS=2 PBRTGPU_TRACE_SHARED_SCENE=0" "PBRTGPU_SORT_HITS=2 PBRTGPU_TRACE_SHARED_SCENE=1"; do
  echo "== $cfg" >> gpurun_out/r02l_ab.log
  env $cfg python tools/bench_configs.py "c1 cornell_dielectric path 512" "strategy light" "strategy portal" 2>>gpurun_out/r02l_ab.err | python -c "
import sys, json
for l in sys.stdin:
    l = l.strip()
    if not l.startswith('{'): continue
    j = json.loads(l)
    print(j['config'][:44].ljust(44), j['Msamples_per_s'], 'Msamples/s', j['kernel_ms'])
" >> gpurun_out/r02l_ab.log
done
cat gpurun_out/r02l_ab.log
python bench.py --no-cpu-baseline > gpurun_out/r02l_bench.jsonl 2> gpurun_out/r02l_bench.err
python -c "
import json
l=[x for x in open('gpurun_out/r02l_bench.jsonl') if x.startswith('{')][-1]; j=json.loads(l); print(j['value'], j['ms_per_step'], j['e2e']['value'], j['roofline']['kernel_ms_in_profile_step']); print({k[:28]:round(v['value'],1) for k,v in j['other_configs'].items()})"
mkdir -p /tmp/prof
B="python bench.py --steps 1 --warmup 1 --no-other-configs --no-cpu-baseline"
ncu --set full --clock-control none --import-source on -k regex:"k_ptrace_fused" -s 3 -c 1 -f -o /tmp/prof/r02l_trace_src $B > gpurun_out/r02l_ncu_trace.log 2>&1
python tools/ncu_summary.py /tmp/prof/r02l_trace_src.ncu-rep > gpurun_out/r02l_trace_src_summary.txt 2>&1
ncu -i /tmp/prof/r02l_trace_src.ncu-rep --page source --csv > gpurun_out/r02l_trace_src_source.csv 2>/dev/null
ls -la gpurun_out | tail -12
