#!/bin/bash
# Round-2 evidence visit (one GPU): smoke, the driver's default bench line, the reference arm, then ncu on the SAME
# command in its --quick form: launch list, DRAM traffic list, full captures of the dominant kernels.
# Usage (under gpurun, repo root):  bash scripts/gpu_prof_r02.sh TAG
TAG=${1:-r02p}
OUT=gpurun_out/$TAG
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw,memory.total --format=csv > $OUT/smi.csv 2>&1
python -c "import __graft_entry__ as g; g.smoke()" > $OUT/smoke.log 2>&1; echo "smoke rc=$?" | tee -a $OUT/summary.txt; tail -1 $OUT/smoke.log
timeout 900 python bench.py --steps 20 --warmup 5 > $OUT/bench_train.json 2> $OUT/bench_train.err; echo "bench train rc=$?" | tee -a $OUT/summary.txt
timeout 900 python bench.py --impl reference --steps 3 --warmup 1 > $OUT/bench_reference.json 2> $OUT/bench_reference.err; echo "bench reference rc=$?" | tee -a $OUT/summary.txt
cut -c1-300 $OUT/bench_reference.json
BENCH="python bench.py --steps 2 --warmup 1 --quick --no-cpu-baseline"
# launch list (per-launch times are cold-cache and serialised: shares, not absolutes)
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file $OUT/launches.csv $BENCH > $OUT/ncu_list.log 2>&1; echo "ncu list rc=$?" | tee -a $OUT/summary.txt
# DRAM bytes of every launch of the same command -> per-kernel average traffic per launch (roofline.traffic)
timeout 900 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 4000 --csv --log-file $OUT/traffic.csv $BENCH > $OUT/ncu_traffic.log 2>&1; echo "ncu traffic rc=$?" | tee -a $OUT/summary.txt
# full captures: regex : launches to skip (warm-up step) : launches to capture
for spec in "igemm_fwd_umma:24:6" "igemm_wgrad_umma:12:3" "gdn_bwd_dual:6:2" "gdn_fwd_umma:6:2" "conv_c3_umma:2:2" "wgrad_c3_umma_kernel:1:1" "deconv_c3_umma:1:1" "factorized_:2:2" "conditional_:2:2" "pack_multi:1:1"; do
  IFS=: read RE SKIP CNT <<< "$spec"
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:$RE -s $SKIP -c $CNT -o $OUT/${RE}_full -f $BENCH > $OUT/ncu_full_$RE.log 2>&1; echo "ncu full $RE rc=$?" | tee -a $OUT/summary.txt
done
ls -la $OUT | head -40
