set -u
mkdir -p gpurun_out
timeout 300 python tools/sweep_probe.py > gpurun_out/r01e_probe.jsonl 2> gpurun_out/r01e_probe.err; echo "probe exit=$?"; cat gpurun_out/r01e_probe.jsonl; tail -3 gpurun_out/r01e_probe.err
CMD="python tools/sweep_probe.py --shapes 1080,1920,128 --reps 1 --modes sweep"
timeout 300 $CMD > gpurun_out/r01e_ncu_plain.log 2>&1 &&
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:sgm_sweep -s 2 -c 2 -f -o gpurun_out/r01e_sweep_prof $CMD > gpurun_out/r01e_ncu.log 2>&1
echo "ncu exit=$?"; tail -5 gpurun_out/r01e_ncu.log
