#!/usr/bin/env bash
# Stage timings of the other BASELINE.json configs.  (compute-sanitizer is closed on this pool.)
set -u
OUT=gpurun_out; mkdir -p $OUT; TAG=${1:-cfg}
echo "== 1080x1920 D=192 (config 5 frame)"; timeout 300 python tools/stage_bench.py 1080 1920 192 > $OUT/${TAG}_stages_1080p_d192.json 2> $OUT/${TAG}_d192.err; echo "exit=$?"; cat $OUT/${TAG}_stages_1080p_d192.json
echo "== 375x450 D=64 (configs 1-2; L2-resident)"; timeout 300 python tools/stage_bench.py 375 450 64 > $OUT/${TAG}_stages_teddy_d64.json 2> $OUT/${TAG}_teddy.err; echo "exit=$?"; cat $OUT/${TAG}_stages_teddy_d64.json
