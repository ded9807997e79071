#!/usr/bin/env bash
OUT=gpurun_out; mkdir -p $OUT
for f in ${@:-00000000 00001000 00000100 00000010 00000001 00001100 00001111 11000000 00110000}; do
  echo -n "flip=$f "; STM_DEBUG_SGM_FLIP=$f python tools/stage_bench.py 1080 1920 128 --stages sgm --reps 10 2>&1 | tail -1
done | tee $OUT/flip_sweep.txt
