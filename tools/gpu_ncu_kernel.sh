#!/usr/bin/env bash
# One gpurun call: `ncu --set full` of one kernel (regex) of one stage at 1080p D=128.
#   bash tools/gpu_ncu_kernel.sh <tag> <stage> <kernel-regex> [skip] [count]
# ncu runs only after the identical command exited 0 without it.
set -u
TAG=$1; STAGE=$2; KRE=$3; SKIP=${4:-3}; CNT=${5:-1}
OUT=gpurun_out
mkdir -p $OUT
CMD="python tools/stage_bench.py 1080 1920 128 --stages $STAGE --reps 1"
timeout 300 $CMD > $OUT/${TAG}_plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:$KRE -s $SKIP -c $CNT \
    -f -o $OUT/${TAG}_full $CMD > $OUT/${TAG}_ncu.log 2>&1
echo "ncu exit=$?"; tail -3 $OUT/${TAG}_ncu.log
