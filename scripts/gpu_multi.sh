#!/bin/bash
# multi-GPU visit (gpurun --gpus N): the driver's launch line for N ranks, train + eval + reference arm.
# Usage: bash scripts/gpu_multi.sh TAG N
TAG=${1:-m}; N=${2:-2}
OUT=gpurun_out/$TAG
mkdir -p $OUT
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517"
timeout 600 $RUN bench.py --gpus $N --steps 20 --warmup 5 > $OUT/bench_train_n$N.json 2> $OUT/bench_train_n$N.err; echo "train n=$N rc=$?"
cut -c1-500 $OUT/bench_train_n$N.json; tail -3 $OUT/bench_train_n$N.err
timeout 600 $RUN bench.py --gpus $N --workload eval --steps 30 --warmup 5 > $OUT/bench_eval_n$N.json 2> $OUT/bench_eval_n$N.err; echo "eval n=$N rc=$?"
cut -c1-400 $OUT/bench_eval_n$N.json; tail -3 $OUT/bench_eval_n$N.err
timeout 600 $RUN bench.py --impl reference --gpus $N --steps 1 --warmup 1 > $OUT/bench_reference_n$N.json 2> $OUT/bench_reference_n$N.err; echo "reference n=$N rc=$?"
cut -c1-300 $OUT/bench_reference_n$N.json
timeout 600 $RUN tests/nccl_grad_check.py > $OUT/nccl_grad_check_n$N.log 2>&1; echo "nccl grad check rc=$?"; tail -3 $OUT/nccl_grad_check_n$N.log
