#!/usr/bin/env bash
# One gpurun call of round 2: GPU parity tests (named shapes + reference-CUDA pins included),
# smoke(), the bench line, the ncu launch list of the bench and `ncu --set full` of every kernel
# of the path at 1080p D=128.  ncu runs only after the identical command exited 0 without it.
#   gpurun --timeout 2400 -- 'bash tools/gpu_r02_round.sh <tag> [skip-tests]'
set -u
TAG=${1:-r02}
OUT=gpurun_out
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $OUT/${TAG}_smi.csv 2>&1
nproc > $OUT/${TAG}_nproc.txt; free -g >> $OUT/${TAG}_nproc.txt

if [ "${2:-}" != "skip-tests" ]; then
  echo "== pytest -m gpu"
  timeout 1500 python -m pytest tests -m gpu -x -q -s --durations=15 > $OUT/${TAG}_pytest.log 2>&1
  PYTEST_RC=$?
  echo "pytest exit=$PYTEST_RC" >> $OUT/${TAG}_pytest.log
  grep -E "mismatch|max rel err|agree|rows < D|passed|failed|Error|error" $OUT/${TAG}_pytest.log | tail -60
  tail -30 $OUT/${TAG}_pytest.log
  if [ $PYTEST_RC -ne 0 ]; then
    echo "parity tests failed: skipping benches and ncu (never profile a faulting program)"
    exit $PYTEST_RC
  fi
  echo "== smoke()"
  timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $OUT/${TAG}_smoke.log 2>&1; echo "smoke exit=$?"; tail -1 $OUT/${TAG}_smoke.log
fi

echo "== bench"
timeout 900 python bench.py --steps 10 --warmup 3 > $OUT/${TAG}_bench.json 2> $OUT/${TAG}_bench.err
echo "bench exit=$?"; cat $OUT/${TAG}_bench.json; tail -3 $OUT/${TAG}_bench.err

echo "== stage bench D=128 / D=192 / teddy"
timeout 300 python tools/stage_bench.py 1080 1920 128 > $OUT/${TAG}_stages_1080p_d128.json 2> $OUT/${TAG}_stages.err; cat $OUT/${TAG}_stages_1080p_d128.json
timeout 300 python tools/stage_bench.py 1080 1920 192 > $OUT/${TAG}_stages_1080p_d192.json 2>> $OUT/${TAG}_stages.err; cat $OUT/${TAG}_stages_1080p_d192.json

echo "== ncu launch list (only if the plain run exits 0)"
CMD="python bench.py --steps 2 --warmup 1 --frames 2 --no-cpu-baseline --no-configs"
timeout 300 $CMD > $OUT/${TAG}_ncu_plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
    --log-file $OUT/${TAG}_launches.csv $CMD > $OUT/${TAG}_ncu.log 2>&1
echo "ncu launch list exit=$?"

echo "== ncu --set full of every kernel of the path (stage_bench, one launch per stage)"
CMD="python tools/stage_bench.py 1080 1920 128 --reps 1 --warm 0"
timeout 300 $CMD > $OUT/${TAG}_full_plain.log 2>&1 &&
timeout 1500 ncu --set full --clock-control none --import-source on \
    -k regex:'sgm_vsweep|sgm_row_kernel|ssd_rows_kernel|ssd_tiled_kernel|birchfield_row|dp_row|wta_vec_kernel' \
    -c 32 -f -o $OUT/${TAG}_full $CMD > $OUT/${TAG}_full_ncu.log 2>&1
echo "ncu full exit=$?"; tail -3 $OUT/${TAG}_full_ncu.log; ls -la $OUT/${TAG}_full.ncu-rep
