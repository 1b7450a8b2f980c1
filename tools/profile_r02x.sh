#!/bin/bash
# Final ncu evidence of round 2 (one gpurun call): launch list of the bench command, `--set full` of every kernel of one
# bench step (-> profiles/r02x_ncu_bench_step_summary.txt, which bench.py reads for issue_frac / lanes_per_inst), the Cornell
# box's kernels (all-specular shade instantiation, shared-memory traversal).  Summaries and raw CSVs come back, reports stay.
out=gpurun_out; mkdir -p $out /tmp/prof
export LD_LIBRARY_PATH=pbrt-v3-light-portals_b200/lib:${LD_LIBRARY_PATH:-}
cli=pbrt-v3-light-portals_b200/lib/pbrt_gpu
B="python bench.py --steps 1 --warmup 1 --no-other-configs --no-cpu-baseline"
$B > $out/r02x_plain.log 2>&1 || { echo "plain bench run failed"; tail -5 $out/r02x_plain.log; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file $out/r02x_launches.csv $B > $out/r02x_ncu_launches.log 2>&1
ncu --set full --clock-control none -s 48 -c 42 -f -o /tmp/prof/r02x_ncu_bench_step $B > $out/r02x_ncu_step.log 2>&1
ncu --set full --clock-control none -k regex:"k_miss|k_shade_sync|k_ptrace" -s 2 -c 10 -f -o /tmp/prof/r02x_ncu_cornell $cli --quiet --spp 32 --outfile /tmp/prof/c1.pfm oracle/_ref/scenes/cornell-path.pbrt > $out/r02x_ncu_cornell.log 2>&1
for r in /tmp/prof/r02x_*.ncu-rep; do
  b=$(basename $r .ncu-rep)
  python tools/ncu_summary.py $r > $out/${b}_summary.txt 2>&1
  ncu -i $r --page raw --csv > $out/${b}_raw.csv 2>/dev/null
done
ls -la $out | tail -12
