#!/usr/bin/env bash
# Runs a probe once per variant library (tools/variant_libs.py) on this box: A/B on identical hardware.
#   bash tools/ab_probe.sh "python tools/sgm_overlap_probe.py 128" v0_new v5_allold ...
CMD=$1; shift
for v in "$@"; do
  echo "== $v"
  STM_B200_LIB=tools/bin/libstm_$v.so timeout 300 $CMD 2>&1 | tail -1 | python -c "
import json,sys
r=json.loads(sys.stdin.readline())
print({k:r[k] for k in ('plain_passes','plain_disparity_only_passes','overlap_disparity_only_ms','overlap_disparity_only_passes')})"
done
