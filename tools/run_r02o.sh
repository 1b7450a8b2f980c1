#!/bin/bash
# r02o: GPU tests with the claim-ahead refill as the default, source-level ncu of the shade kernels, C4 with the round-2 kernels
mkdir -p gpurun_out /tmp/prof
python -m pytest tests -m gpu -x -q > gpurun_out/r02o_tests.log 2>&1; tail -2 gpurun_out/r02o_tests.log
B="python bench.py --steps 1 --warmup 1 --no-other-configs --no-cpu-baseline"
ncu --set full --clock-control none --import-source on -k regex:"k_shade_sync" -s 2 -c 2 -f -o /tmp/prof/r02o_shade_src $B > gpurun_out/r02o_ncu_shade.log 2>&1
python tools/ncu_summary.py /tmp/prof/r02o_shade_src.ncu-rep > gpurun_out/r02o_shade_src_summary.txt 2>&1
ncu -i /tmp/prof/r02o_shade_src.ncu-rep --page source --csv > gpurun_out/r02o_shade_src_source.csv 2>/dev/null
( time python bench.py --workload c4 --steps 2 --warmup 1 --no-cpu-baseline --no-other-configs > gpurun_out/r02o_bench_c4.jsonl 2> gpurun_out/r02o_bench_c4.err ) 2> gpurun_out/r02o_c4_time.txt
python -c "
import json
l=[x for x in open('gpurun_out/r02o_bench_c4.jsonl') if x.startswith('{')][-1]; j=json.loads(l); print('c4', j['value'], j['ms_per_step'], j['e2e'], j['roofline']['kernel_ms_in_profile_step'], j['roofline']['node_tests_per_ray'])"
ncu --set full --clock-control none -k regex:"k_ptrace_fused" -s 1 -c 1 -f -o /tmp/prof/r02o_c4_trace python bench.py --workload c4 --steps 1 --warmup 1 --no-other-configs --no-cpu-baseline > gpurun_out/r02o_ncu_c4.log 2>&1
python tools/ncu_summary.py /tmp/prof/r02o_c4_trace.ncu-rep > gpurun_out/r02o_c4_trace_summary.txt 2>&1
cat gpurun_out/r02o_c4_trace_summary.txt gpurun_out/r02o_shade_src_summary.txt
