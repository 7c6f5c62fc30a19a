#!/bin/bash
# quick GPU visit: selected tests + train bench.  Usage: bash scripts/gpu_quick.sh TAG "pytest args"
TAG=${1:-q}
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 1200 python -m pytest tests -m gpu -x -q ${2:-} > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -25 $OUT/pytest_gpu.log
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > $OUT/bench_train.json 2> $OUT/bench_train.err; echo "bench rc=$?"
tail -3 $OUT/bench_train.err; cat $OUT/bench_train.json
