#!/bin/bash
# eval bench with and without CUDA graphs + the graph tests.  Usage: bash scripts/gpu_eval.sh TAG
TAG=${1:-ev}
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 900 python -m pytest tests/test_gpu_graphs.py -x -q > $OUT/pytest_graphs.log 2>&1; echo "pytest rc=$?"
tail -15 $OUT/pytest_graphs.log
for v in "" "--no-graphs"; do
  timeout 600 python bench.py --workload eval --steps 50 --warmup 10 --no-cpu-baseline $v > $OUT/bench_eval$v.json 2> $OUT/bench_eval$v.err; echo "bench $v rc=$?"
  tail -3 $OUT/bench_eval$v.err; python -c "
import json,sys; d=json.load(open('$OUT/bench_eval$v.json')); print({k:d[k] for k in ('value','ms_per_step','gpu_launches')}, d['e2e'])"
done
