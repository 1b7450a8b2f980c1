# A/B on the GPU box: block shape of the Lambert shade instantiation (and the instantiation off), killeroo + Cornell
for v in "PBRTGPU_SHADE_LAMBERT=0" "PBRTGPU_LAMBERT_BLOCK=5122" "PBRTGPU_LAMBERT_BLOCK=5121" "PBRTGPU_LAMBERT_BLOCK=3842" "PBRTGPU_LAMBERT_BLOCK=2563" "PBRTGPU_LAMBERT_BLOCK=2562"; do
  echo "== $v"
  env $v python bench.py --steps 6 --warmup 2 --no-cpu-baseline --no-other-configs | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); r=d['roofline']
print('killeroo %.1f Msamples/s  %.2f ms/step  shade %.2f ms  trace %.2f ms' % (d['value'], d['ms_per_step'], r['kernel_ms_in_profile_step']['shade'], r['kernel_ms_in_profile_step']['trace_closest']))"
  env $v python tools/bench_configs.py "512x512 64spp" | python -c "
import json,sys
for l in sys.stdin:
    d=json.loads(l); print('%-52s %.1f Msamples/s  shade %.2f trace %.2f' % (d['config'], d['Msamples_per_s'], d['kernel_ms']['shade'], d['kernel_ms']['trace_closest']))"
done
