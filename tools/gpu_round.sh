#!/usr/bin/env bash
# One gpurun call: GPU parity tests, per-stage timings, the bench line, then an ncu launch list.
#   gpurun --timeout 1500 -- 'bash tools/gpu_round.sh <tag>'
# Outputs land in gpurun_out/<tag>_*.  ncu runs only after the identical command exited 0.
set -u
TAG=${1:-run}
OUT=gpurun_out
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $OUT/${TAG}_smi.csv 2>&1

echo "== pytest -m gpu"
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/${TAG}_pytest.log 2>&1
PYTEST_RC=$?
echo "pytest exit=$PYTEST_RC" >> $OUT/${TAG}_pytest.log
tail -25 $OUT/${TAG}_pytest.log
if [ $PYTEST_RC -ne 0 ]; then
  echo "parity tests failed: skipping benches and ncu (never profile a faulting program)"
  exit $PYTEST_RC
fi

echo "== smoke()"
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $OUT/${TAG}_smoke.log 2>&1; echo "smoke exit=$?"; tail -1 $OUT/${TAG}_smoke.log

echo "== stage bench"
timeout 300 python tools/stage_bench.py 1080 1920 128 > $OUT/${TAG}_stages_1080p_d128.json 2> $OUT/${TAG}_stages.err
echo "stage_bench exit=$?"; cat $OUT/${TAG}_stages_1080p_d128.json

echo "== bench"
timeout 600 python bench.py --steps 5 --warmup 3 > $OUT/${TAG}_bench.json 2> $OUT/${TAG}_bench.err
echo "bench exit=$?"; cat $OUT/${TAG}_bench.json; tail -3 $OUT/${TAG}_bench.err

echo "== ncu launch list (only if the plain run exits 0)"
CMD="python bench.py --steps 2 --warmup 1 --frames 1 --no-cpu-baseline"
timeout 300 $CMD > $OUT/${TAG}_ncu_plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
    --log-file $OUT/${TAG}_launches.csv $CMD > $OUT/${TAG}_ncu.log 2>&1
echo "ncu exit=$?"
