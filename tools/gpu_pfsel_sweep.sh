#!/usr/bin/env bash
# SGM stage time for several per-direction prefetch-variant choices (developer sweep).
OUT=gpurun_out; mkdir -p $OUT
for sel in ${@:-"--------" "00220000" "11220000" "10220000" "01220000"}; do
  echo -n "$sel "; STM_DEBUG_SGM_PFSEL=$sel python tools/stage_bench.py 1080 1920 128 --stages sgm --reps 10 2>&1 | tail -1
done | tee $OUT/pfsel_sweep.txt
