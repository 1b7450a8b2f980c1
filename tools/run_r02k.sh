python -m pytest tests -m gpu -x -q -k "rays or oracle or golden or killeroo or no_alloc" > gpurun_out/r02k_tests.log 2>&1; tail -2 gpurun_out/r02k_tests.log
python bench.py --no-cpu-baseline > gpurun_out/r02k_bench.jsonl 2> gpurun_out/r02k_bench.err
python -c "
import json
l=[x for x in open('gpurun_out/r02k_bench.jsonl') if x.startswith('{')][-1]; j=json.loads(l); print(j['value'], j['ms_per_step'], j['e2e']['value'], j['roofline']['kernel_ms_in_profile_step']); print({k[:28]:round(v['value'],1) for k,v in j['other_configs'].items()})"
mkdir -p /tmp/prof
B="python bench.py --steps 1 --warmup 1 --no-other-configs --no-cpu-baseline"
ncu --set full --clock-control none --import-source on -k regex:"k_shade_sync" -s 2 -c 2 -f -o /tmp/prof/r02k_shade_src $B > gpurun_out/r02k_ncu_shade.log 2>&1
python tools/ncu_summary.py /tmp/prof/r02k_shade_src.ncu-rep > gpurun_out/r02k_shade_src_summary.txt 2>&1
ncu -i /tmp/prof/r02k_shade_src.ncu-rep --page source --csv > gpurun_out/r02k_shade_src_source.csv 2>/dev/null
ls -la gpurun_out | tail -8
