#!/bin/bash
# ncu evidence of round 2 (run on the GPU box, one gpurun call): launch list of the bench command, `--set full` of every
# kernel of one bench step, and of the kernels the bench scene does not reach (k_miss, k_shade_direct, the EXT shade
# instantiation, k_pintersect_*, the HLBVH build and layout kernels).  Reports go to gpurun_out/r02_*.ncu-rep;
# The .ncu-rep files stay on the box (a full-set report is ~3 MB per launch and gpurun_out/ carries 64 MiB): each is turned
# into the per-kernel summary of tools/ncu_summary.py and the raw metric table (CSV) there, and those come back.
out=gpurun_out; mkdir -p $out /tmp/prof
export LD_LIBRARY_PATH=pbrt-v3-light-portals_b200/lib:${LD_LIBRARY_PATH:-}
cli=pbrt-v3-light-portals_b200/lib/pbrt_gpu
B="python bench.py --steps 1 --warmup 1 --no-other-configs --no-cpu-baseline"
$B > $out/r02_plain.log 2>&1 || { echo "plain bench run failed"; tail -5 $out/r02_plain.log; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file $out/r02_launches.csv $B > $out/r02_ncu_launches.log 2>&1
# one step's worth of launches after the uploads and the warm-up step (7 + 37 launches)
ncu --set full --clock-control none -s 48 -c 42 -f -o /tmp/prof/r02_bench_step $B > $out/r02_ncu_step.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_ptrace_fused -s 8 -c 1 -f -o /tmp/prof/r02_trace_src $B > $out/r02_ncu_trace_src.log 2>&1
# kernels the bench scene does not reach
python tools/gen_c3.py /tmp/prof/c3_mis3.pbrt --strategy solidangle_mis3 --spp 16 --res 512 > /dev/null
ncu --set full --clock-control none -k regex:"k_shade_sync|k_nee_resolve" -s 2 -c 4 -f -o /tmp/prof/r02_c3_mis3 $cli --quiet --outfile /tmp/prof/c3.pfm /tmp/prof/c3_mis3.pbrt > $out/r02_ncu_c3.log 2>&1
ncu --set full --clock-control none -k regex:"k_miss|k_shade_sync" -s 1 -c 5 -f -o /tmp/prof/r02_cornell $cli --quiet --spp 32 --outfile /tmp/prof/c1.pfm oracle/_ref/scenes/cornell-path.pbrt > $out/r02_ncu_cornell.log 2>&1
ncu --set full --clock-control none -k regex:"k_shade_direct|k_ptrace_fused" -s 1 -c 4 -f -o /tmp/prof/r02_lamp_direct $cli --quiet --spp 16 --outfile /tmp/prof/lamp.pfm oracle/_ref/scenes/lamp/lamp.pbrt > $out/r02_ncu_lamp.log 2>&1
ncu --set full --clock-control none -k regex:"k_pintersect" -c 4 -f -o /tmp/prof/r02_rays python tools/bench_rays.py --rays 4000000 --reps 1 --check-rays 0 > $out/r02_ncu_rays.log 2>&1
ncu --set full --clock-control none -k regex:"k_centroid|k_morton|k_radix|k_scan|k_treelet|k_emit_level|k_up_level|k_down_level|k_interior|k_wide_nodes|k_prim_records|k_tag_last" -c 40 -f -o /tmp/prof/r02_build python tools/bench_build.py --quads 1000 --repeat 1 > $out/r02_ncu_build.log 2>&1
for r in /tmp/prof/r02_*.ncu-rep; do
  b=$(basename $r .ncu-rep)
  python tools/ncu_summary.py $r > $out/${b}_summary.txt 2>&1
  ncu -i $r --page raw --csv > $out/${b}_raw.csv 2>/dev/null
done
ncu -i /tmp/prof/r02_trace_src.ncu-rep --page source --csv > $out/r02_trace_src_source.csv 2>/dev/null
ls -la /tmp/prof/*.ncu-rep $out | tail -40
du -sh $out
