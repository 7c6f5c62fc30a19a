#!/bin/bash
# multi-GPU visit (gpurun --gpus N): NCCL gradient check, then the driver's launch line for N ranks.
# Usage: bash scripts/gpu_multi.sh TAG N [extra env for an A/B run, e.g. HPC_AR_OVERLAP=0]
TAG=${1:-m}; N=${2:-2}; AB=${3:-}
OUT=gpurun_out/$TAG
mkdir -p $OUT
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517"
timeout 600 $RUN tests/nccl_grad_check.py > $OUT/nccl_grad_check_n$N.log 2>&1; echo "nccl grad check rc=$?"; grep -v "^W\|^\[" $OUT/nccl_grad_check_n$N.log | tail -8
timeout 900 $RUN bench.py --gpus $N --steps 20 --warmup 5 > $OUT/bench_train_n$N.json 2> $OUT/bench_train_n$N.err; echo "train n=$N rc=$?"
tail -3 $OUT/bench_train_n$N.err
if [ -n "$AB" ]; then
  env $AB timeout 600 $RUN bench.py --gpus $N --steps 20 --warmup 5 --quick > $OUT/bench_train_n${N}_ab.json 2> $OUT/bench_train_n${N}_ab.err; echo "train A/B ($AB) rc=$?"
fi
timeout 300 python bench.py --gpus 1 --steps 20 --warmup 5 --quick --no-cpu-baseline > $OUT/bench_train_n1.json 2> $OUT/bench_train_n1.err; echo "train n=1 rc=$?"
python - <<PY
import json, glob
for f in sorted(glob.glob("$OUT/bench_train_n*.json")):
    try:
        d = json.load(open(f))
    except Exception as e:
        print(f, "unreadable", e); continue
    ex = {k: (round(v["value"], 1), round(v["ms_per_step"], 3)) for k, v in d.get("extra", {}).items()}
    print(f, "n_gpus", d["n_gpus"], round(d["value"], 1), d["unit"], round(d["ms_per_step"], 4), "ms/step e2e", round(d["e2e"]["value"], 1), "sustained", d.get("sustained", {}).get("ms_per_step"), ex)
PY
