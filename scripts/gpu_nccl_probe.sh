#!/bin/bash
# Usage (under gpurun --gpus N): bash scripts/gpu_nccl_probe.sh TAG N
TAG=$1; N=$2; OUT=gpurun_out/$TAG; mkdir -p $OUT
run() { env "$@" timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 scripts/nccl_ar_probe.py 2>$OUT/err_$RANDOM.log | tail -1 | tee -a $OUT/probe.jsonl; }
run NCCL_DEBUG=WARN
run NCCL_ALGO=Ring
run NCCL_ALGO=Tree
run NCCL_ALGO=NVLS
run NCCL_NVLS_ENABLE=0
NCCL_DEBUG=INFO NCCL_DEBUG_SUBSYS=INIT,TUNING timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29534 scripts/nccl_ar_probe.py 2>&1 | grep -i "nvls\|algo\|channel\|Connected" | head -40 > $OUT/nccl_info.log
tail -5 $OUT/nccl_info.log
