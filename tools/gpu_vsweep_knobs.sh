#!/usr/bin/env bash
# Fused-sweep probe with the developer knob STM_DEBUG_VSWEEP (0 = product; 1 = importer does not
# poll; 3 = no waits at all -- wrong results, timing only).  bash tools/gpu_vsweep_knobs.sh <tag>
TAG=${1:-knobs}; OUT=gpurun_out; mkdir -p $OUT
for k in ${KNOBS:-0 1 3}; do
  STM_DEBUG_VSWEEP=$k timeout 200 python tools/vsweep_probe.py 1080 1920 128 --reps 5 > $OUT/${TAG}_dbg$k.json 2> $OUT/${TAG}_dbg$k.err
  echo "dbg=$k rc=$?"; python - <<PY
import json
d=json.load(open("$OUT/${TAG}_dbg$k.json"))
print({k:d[k] for k in ("max_rel_err","fused_ms","fused_wta_ms","fused_passes")})
PY
done
