#!/bin/bash
# N-GPU A/B of environment settings with the quick bench.  Usage: bash scripts/gpu_multi_ab.sh TAG N "ENV1" "ENV2" ...
TAG=$1; N=$2; shift 2
OUT=gpurun_out/$TAG; mkdir -p $OUT
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29519"
i=0
for E in "$@"; do
  i=$((i+1))
  env $E timeout 300 $RUN bench.py --gpus $N --steps 20 --warmup 5 --quick --no-cpu-baseline > $OUT/ab_$i.json 2> $OUT/ab_$i.err; echo "[$E] rc=$?"
  python - <<PY
import json
try:
    d = json.load(open("$OUT/ab_$i.json"))
    print("   $E:", round(d["ms_per_step"], 4), "ms/step", round(d["value"], 1), {k: round(v["ms_per_step"], 3) for k, v in d["kernels"].items()})
except Exception as e:
    print("   unreadable", e)
PY
done
