#!/bin/bash
# r02x: the round's final evidence -- all GPU tests, the full bench line (other configs, CPU baseline), the reference arm, ncu
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r02x_tests.log 2>&1; tail -2 gpurun_out/r02x_tests.log
python bench.py > gpurun_out/r02x_bench.jsonl 2> gpurun_out/r02x_bench.err
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r02x_bench_reference.jsonl 2> gpurun_out/r02x_bench_reference.err
python -c "
import json
l=[x for x in open('gpurun_out/r02x_bench.jsonl') if x.startswith('{')][-1]; j=json.loads(l); print(j['value'], j['ms_per_step'], j['e2e']['value'], j['roofline']['kernel_ms_in_profile_step'], j['cpu_baseline']); print({k[:40]:round(v['value'],1) for k,v in j['other_configs'].items()})"
tail -1 gpurun_out/r02x_bench_reference.jsonl | cut -c1-400
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02x_smoke.log 2>&1; tail -2 gpurun_out/r02x_smoke.log
bash tools/profile_r02x.sh
