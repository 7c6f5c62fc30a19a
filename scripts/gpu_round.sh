#!/bin/bash
# One GPU-box visit: smoke, parity tests, bench lines (ours + reference arm), ncu launch list, DRAM traffic list,
# full captures of the dominant kernels.  Usage (from the repo root, under gpurun):  bash scripts/gpu_round.sh [tag]
TAG=${1:-r01}
OUT=gpurun_out/$TAG
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw,memory.total --format=csv > $OUT/smi.csv 2>&1
python -c "import __graft_entry__ as g; g.smoke()" > $OUT/smoke.log 2>&1; echo "smoke rc=$?" | tee -a $OUT/summary.txt
timeout 1500 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/summary.txt
tail -3 $OUT/pytest_gpu.log
timeout 600 python bench.py --steps 20 --warmup 5 > $OUT/bench_train.json 2> $OUT/bench_train.err; echo "bench train rc=$?" | tee -a $OUT/summary.txt
cut -c1-400 $OUT/bench_train.json
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > $OUT/bench_reference.json 2> $OUT/bench_reference.err; echo "bench reference rc=$?" | tee -a $OUT/summary.txt
cut -c1-300 $OUT/bench_reference.json
timeout 600 python bench.py --workload eval --steps 50 --warmup 10 --no-cpu-baseline > $OUT/bench_eval.json 2> $OUT/bench_eval.err; echo "bench eval rc=$?" | tee -a $OUT/summary.txt
cut -c1-300 $OUT/bench_eval.json
timeout 600 python bench.py --workload train_ssim --steps 10 --warmup 3 --no-cpu-baseline > $OUT/bench_ssim.json 2> $OUT/bench_ssim.err; echo "bench ssim rc=$?" | tee -a $OUT/summary.txt
cut -c1-300 $OUT/bench_ssim.json
# launch list of the same command (per-launch times are cold-cache and serialised: shares, not absolutes)
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file $OUT/launches.csv \
    python bench.py --steps 2 --warmup 1 --no-cpu-baseline > $OUT/ncu_list.log 2>&1; echo "ncu list rc=$?" | tee -a $OUT/summary.txt
# DRAM bytes of every launch of the same command -> per-kernel average traffic per launch (roofline.traffic)
timeout 900 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 3000 --csv --log-file $OUT/traffic.csv \
    python bench.py --steps 2 --warmup 1 --no-cpu-baseline > $OUT/ncu_traffic.log 2>&1; echo "ncu traffic rc=$?" | tee -a $OUT/summary.txt
# full captures of the dominant kernels (one training step's worth of each)
for spec in "igemm_fwd_umma:24:8" "igemm_wgrad_umma:12:4" "gdn_bwd_dual:6:3" "gdn_fwd_umma:6:3" "conv_c3_umma:2:2" "wgrad_c3_umma_kernel:1:1" "deconv_c3_umma:1:1" "conditional_:2:2" "adamw_step:3:1"; do
  IFS=: read RE SKIP CNT <<< "$spec"
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:$RE -s $SKIP -c $CNT -o $OUT/${RE}_full -f \
      python bench.py --steps 1 --warmup 1 --no-cpu-baseline > $OUT/ncu_full_$RE.log 2>&1; echo "ncu full $RE rc=$?" | tee -a $OUT/summary.txt
done
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"ssim_level_rb_kernel|ssim_bwd_filter_rb" -c 3 -o $OUT/ssim_rb_full -f \
    python bench.py --workload train_ssim --steps 1 --warmup 0 --no-cpu-baseline > $OUT/ncu_full_ssim_rb.log 2>&1; echo "ncu full ssim rc=$?" | tee -a $OUT/summary.txt
ls -la $OUT
