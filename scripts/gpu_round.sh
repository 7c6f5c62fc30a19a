#!/bin/bash
# One GPU-box visit: parity tests, bench lines, ncu launch list, one full capture of the dominant kernel.
# Usage (from the repo root, under gpurun):  bash scripts/gpu_round.sh [tag]
TAG=${1:-r01}
OUT=gpurun_out/$TAG
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw,memory.total --format=csv > $OUT/smi.csv 2>&1
python -c "import __graft_entry__ as g; g.smoke()" > $OUT/smoke.log 2>&1; echo "smoke rc=$?" | tee -a $OUT/summary.txt
timeout 1500 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/summary.txt
tail -5 $OUT/pytest_gpu.log
timeout 600 python bench.py --steps 20 --warmup 5 > $OUT/bench_train.json 2> $OUT/bench_train.err; echo "bench train rc=$?" | tee -a $OUT/summary.txt
cat $OUT/bench_train.json
timeout 600 python bench.py --workload eval --steps 20 --warmup 5 --no-cpu-baseline > $OUT/bench_eval.json 2> $OUT/bench_eval.err; echo "bench eval rc=$?" | tee -a $OUT/summary.txt
cat $OUT/bench_eval.json
timeout 600 python bench.py --workload train_ssim --steps 10 --warmup 3 --no-cpu-baseline > $OUT/bench_ssim.json 2> $OUT/bench_ssim.err; echo "bench ssim rc=$?" | tee -a $OUT/summary.txt
cat $OUT/bench_ssim.json
# launch list of the same command (per-launch times are cold-cache and serialised: shares, not absolutes)
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file $OUT/launches.csv \
    python bench.py --steps 2 --warmup 1 --no-cpu-baseline > $OUT/ncu_list.log 2>&1; echo "ncu list rc=$?" | tee -a $OUT/summary.txt
# one full capture of the dominant kernel (a few launches of the UMMA forward kernel, g_a conv1 shape among them)
timeout 900 ncu --set full --clock-control none --import-source on -k regex:igemm_fwd_umma -s 20 -c 6 -o $OUT/igemm_fwd_full -f \
    python bench.py --steps 1 --warmup 1 --no-cpu-baseline > $OUT/ncu_full.log 2>&1; echo "ncu full rc=$?" | tee -a $OUT/summary.txt
timeout 900 ncu --set full --clock-control none --import-source on -k regex:igemm_wgrad_umma -s 4 -c 3 -o $OUT/igemm_wgrad_full -f \
    python bench.py --steps 1 --warmup 1 --no-cpu-baseline > $OUT/ncu_full2.log 2>&1; echo "ncu wgrad rc=$?" | tee -a $OUT/summary.txt
ls -la $OUT
