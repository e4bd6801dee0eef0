#!/bin/bash
# every bench workload once on one GPU (short); logs under gpurun_out/
mkdir -p gpurun_out
for wl in 1080p labels 16k; do
  timeout 900 python bench.py --workload $wl --steps ${STEPS:-3} --warmup 3 ${EXTRA} > gpurun_out/bench_$wl.log 2> gpurun_out/bench_$wl.err
  echo "$wl exit $?"; cut -c1-900 gpurun_out/bench_$wl.log; tail -3 gpurun_out/bench_$wl.err
done
