#!/usr/bin/env bash
# Stage times and the per-pass ncu launch list of the SGM stage for one shape (developer tool).
#   bash tools/gpu_sgm_passes.sh <tag> H W D
TAG=$1; H=$2; W=$3; D=$4
OUT=gpurun_out; mkdir -p $OUT
python tools/stage_bench.py $H $W $D --stages ssd,sgm,wta --reps 10 | tee $OUT/${TAG}_stages.json
CMD="python tools/stage_bench.py $H $W $D --stages sgm --reps 1"
timeout 300 $CMD > $OUT/${TAG}_plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 80 --csv \
    --log-file $OUT/${TAG}_launches.csv $CMD > $OUT/${TAG}_ncu.log 2>&1
echo "ncu exit=$?"
