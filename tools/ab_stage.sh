#!/usr/bin/env bash
# Stage times once per variant library on this box.   bash tools/ab_stage.sh "ssd" ssd_nh7 ssd_nh11 ...
STAGES=$1; shift
for v in "$@"; do
  echo "== $v"
  STM_B200_LIB=tools/bin/libstm_$v.so timeout 300 python tools/stage_bench.py 1080 1920 128 --stages $STAGES 2>&1 | tail -1
  STM_B200_LIB=tools/bin/libstm_$v.so timeout 300 python tools/stage_bench.py 1080 1920 192 --stages $STAGES 2>&1 | tail -1
done
