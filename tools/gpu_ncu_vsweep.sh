#!/usr/bin/env bash
# ncu --set full of the fused-sweep kernel (first two launches of the probe = down and up sweep).
#   STM_DEBUG_VSWEEP=<k> bash tools/gpu_ncu_vsweep.sh <tag>
TAG=${1:-vsweep}; OUT=gpurun_out; mkdir -p $OUT
CMD="python tools/vsweep_probe.py 1080 1920 128 --reps 1"
timeout 200 $CMD > $OUT/${TAG}_plain.log 2>&1 || { echo "plain run failed"; tail -3 $OUT/${TAG}_plain.log; exit 86; }
timeout 800 ncu ${NCU_SET:---set full --import-source on} --clock-control none -k regex:sgm_vsweep -c 2 \
    -o $OUT/${TAG}_prof -f $CMD > $OUT/${TAG}_ncu.log 2>&1
echo "ncu exit=$?"; tail -3 $OUT/${TAG}_ncu.log
