#!/bin/bash
# One gpurun call: parity tests, smoke, bench (U headline, N, slot sweeps), launch lists, one
# ncu --set full capture of every kernel of one launch group.  Logs under gpurun_out/.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/gpu.txt 2>&1
timeout 1200 python -m pytest tests -m gpu -q --tb=short --timeout 600 -p no:cacheprovider > gpurun_out/pytest_gpu.log 2>&1
echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -8 gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke exit $?" >> gpurun_out/smoke.log
tail -2 gpurun_out/smoke.log
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/bench.log 2> gpurun_out/bench.err; echo "bench exit $?"
cut -c1-1500 gpurun_out/bench.log; tail -5 gpurun_out/bench.err
for k in U N; do for s in 1 2 8; do
  timeout 300 python bench.py --steps 3 --warmup 3 --kind $k --slots $s --no-cpu --no-e2e > gpurun_out/bench_${k}_s$s.log 2> gpurun_out/bench_${k}_s$s.err
  echo "kind $k slots $s exit $?"
  grep -o '"value": [0-9.]*' gpurun_out/bench_${k}_s$s.log | head -1
  grep -o '"kernels": {.*"phase_share_of_step": {[^}]*}' gpurun_out/bench_${k}_s$s.log
done; done
for k in U N; do
CMD="python bench.py --steps 1 --warmup 3 --frames 4 --kind $k --no-cpu --no-e2e"
$CMD > gpurun_out/plain_$k.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_$k.csv $CMD > gpurun_out/ncu_list_$k.log 2>&1
echo "ncu list $k exit $?"
done
for k in U N; do
CMD="python bench.py --steps 1 --warmup 3 --frames 4 --kind $k --no-cpu --no-e2e"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"^k_" -s 39 -c 13 -f -o gpurun_out/prof_r01_v3_$k $CMD > gpurun_out/ncu_full_$k.log 2>&1
echo "ncu full $k exit $?"; tail -2 gpurun_out/ncu_full_$k.log | cut -c1-300
done
ls -la gpurun_out
