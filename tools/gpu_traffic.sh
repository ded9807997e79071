#!/usr/bin/env bash
# One gpurun call: measured DRAM bytes per kernel for every stage (1080p D=128) and an
# `ncu --set full` report of the dominant kernel.  ncu runs only after the same command exited 0.
set -u
TAG=${1:-traffic}
OUT=gpurun_out
mkdir -p $OUT
CMD="python tools/stage_bench.py 1080 1920 128 --stages ssd,bt,sgm,wta,dp --reps 1"
timeout 300 $CMD > $OUT/${TAG}_plain.log 2>&1 &&
timeout 900 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum \
    --clock-control none -c 200 --csv --log-file $OUT/${TAG}_dram.csv $CMD > $OUT/${TAG}_ncu1.log 2>&1
echo "ncu dram pass exit=$?"
timeout 300 $CMD > $OUT/${TAG}_plain2.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:sgm_vsweep_kernel -c 2 \
    -f -o $OUT/${TAG}_sgm_vsweep_full $CMD > $OUT/${TAG}_ncu2.log 2>&1
echo "ncu full pass exit=$?"; tail -3 $OUT/${TAG}_ncu2.log
