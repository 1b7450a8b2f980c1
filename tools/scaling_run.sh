#!/bin/bash
# bench.py at N = 1, 2, 4, 8 GPUs of one box (or the Ns given), launched the way the driver does (torchrun for N > 1).
# usage (on the GPU box): tools/scaling_run.sh outfile [N ...]   -> one JSON line per N
out=${1:-gpurun_out/scaling.jsonl}
shift
ns=${@:-1 2 4 8}
mkdir -p "$(dirname "$out")"
: > "$out"
for n in $ns; do
  if [ "$n" = 1 ]; then
    python bench.py --gpus 1 --steps 8 --warmup 3 --no-cpu-baseline 2>>"$out.err" | grep '^{' >> "$out"
  else
    python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500 + n)) \
      bench.py --gpus $n --steps 8 --warmup 3 --no-cpu-baseline 2>>"$out.err" | grep '^{' >> "$out"
  fi
done
python - "$out" <<'PY'
import json, sys
for l in open(sys.argv[1]):
    j = json.loads(l)
    print(j["n_gpus"], "GPUs:", round(j["value"], 1), j["unit"], round(j["ms_per_step"], 2), "ms/step, e2e", round(j["e2e"]["value"], 1))
PY
