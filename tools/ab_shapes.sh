#!/usr/bin/env bash
# SGM stage times for a list of shapes once per variant library.  bash tools/ab_shapes.sh "c48 c384" "375 450 128" "480 640 64" ...
VARS=$1; shift
for shape in "$@"; do
  for v in $VARS; do
    echo -n "$v $shape: "
    STM_B200_LIB=tools/bin/libstm_$v.so timeout 300 python tools/stage_bench.py $shape --stages sgm,sgm_wta 2>&1 | tail -1
  done
done
