#!/bin/bash
# A/B of programmatic dependent launch and of the per-kernel event pairs.  Usage: bash scripts/gpu_pdl.sh TAG
TAG=${1:-pdl}; OUT=gpurun_out/$TAG; mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 $OUT/pytest_gpu.log
for pdl in 0 1; do for ev in timed after; do
  HPC_PDL=$pdl timeout 300 python bench.py --steps 30 --warmup 5 --no-cpu-baseline --kernel-events $ev > $OUT/bench_pdl${pdl}_$ev.json 2> $OUT/bench_pdl${pdl}_$ev.err
  echo "pdl=$pdl events=$ev rc=$? $(python -c "import json,sys; d=json.load(open('$OUT/bench_pdl${pdl}_$ev.json')); print(d['ms_per_step'], d['value'], d['e2e']['value'])")"
done; done
