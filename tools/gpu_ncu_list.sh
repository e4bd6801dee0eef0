#!/bin/bash
# launch list (per-kernel device time) of one bench step; shares only, never a bench value
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 3 --frames ${FRAMES:-4} --kind ${KIND:-U} --no-cpu --no-e2e"
$CMD > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_${KIND:-U}.csv $CMD > gpurun_out/ncu.log 2>&1
echo "exit $?"; tail -3 gpurun_out/plain.log | cut -c1-600; tail -3 gpurun_out/ncu.log | cut -c1-300
