#!/bin/bash
# One GPU-box visit of round 2.  Usage (under gpurun, repo root):  bash scripts/gpu_r02.sh TAG [pytest -k expr] [bench args]
TAG=${1:-r02}; KEXPR=${2:-}; shift 2
OUT=gpurun_out/$TAG
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw,memory.total --format=csv > $OUT/smi.csv 2>&1
if [ -n "$KEXPR" ]; then
  timeout 1500 python -m pytest tests -m gpu -x -q -s -k "$KEXPR" > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/summary.txt
else
  timeout 1700 python -m pytest tests -m gpu -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/summary.txt
fi
tail -30 $OUT/pytest_gpu.log
timeout 900 python bench.py --steps 20 --warmup 5 "$@" > $OUT/bench_train.json 2> $OUT/bench_train.err; echo "bench train rc=$?" | tee -a $OUT/summary.txt
tail -5 $OUT/bench_train.err
cut -c1-1500 $OUT/bench_train.json
if [ -n "$AB_ENV" ]; then   # A/B of one environment switch inside the same box visit:  AB_ENV="HPC_SIDE_WGRAD=0"
  env $AB_ENV timeout 600 python bench.py --steps 20 --warmup 5 --quick --no-cpu-baseline > $OUT/bench_train_ab.json 2> $OUT/bench_train_ab.err
  echo "A/B ($AB_ENV):"; python - <<PY
import json
for f in ("bench_train.json", "bench_train_ab.json"):
    d = json.load(open("$OUT/" + f)); print(f, round(d["ms_per_step"], 4), "ms/step", round(d["value"], 1), {k: round(v["ms_per_step"], 4) for k, v in d["kernels"].items()})
PY
fi
