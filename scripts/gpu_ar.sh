#!/bin/bash
# A/B of the early all-reduce at N ranks.  Usage: bash scripts/gpu_ar.sh TAG N "0 1"
TAG=${1:-ar}; N=${2:-2}; OUT=gpurun_out/$TAG; mkdir -p $OUT
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517"
timeout 120 $RUN tests/nccl_grad_check.py > $OUT/nccl_grad_check.log 2>&1; echo "grad check rc=$?"; tail -4 $OUT/nccl_grad_check.log
for ov in ${3:-0 1}; do
  HPC_AR_OVERLAP=$ov timeout 200 $RUN bench.py --gpus $N --steps 40 --warmup 5 > $OUT/bench_n${N}_ov$ov.json 2> $OUT/bench_n${N}_ov$ov.err
  echo "overlap=$ov rc=$? $(python -c "import json; d=json.load(open('$OUT/bench_n${N}_ov$ov.json')); print(d['ms_per_step'], d['value'], d['e2e']['value'])")"
done
