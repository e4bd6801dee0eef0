#!/bin/bash
# apply-gather fix + partition rank trick: parity, A/B bench, launch list, ncu --set full of one group
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q --tb=short --timeout 600 -p no:cacheprovider -x > gpurun_out/pytest_gpu.log 2>&1
echo "pytest all exit $?"; tail -8 gpurun_out/pytest_gpu.log
for k in U N; do for cfg in "1 1" "1 2" "1 3" "1 0"; do
  set -- $cfg
  timeout 300 python bench.py --steps 3 --warmup 3 --kind $k --front $1 --apply $2 --no-cpu --no-e2e > gpurun_out/bench_${k}_f$1a$2.log 2> gpurun_out/bench_${k}_f$1a$2.err
  echo "kind $k front $1 apply $2 exit $?"
  grep -o '"value": [0-9.]*' gpurun_out/bench_${k}_f$1a$2.log | head -1
  grep -o '"kernels": {.*"phase_share_of_step": {[^}]*}' gpurun_out/bench_${k}_f$1a$2.log
  tail -2 gpurun_out/bench_${k}_f$1a$2.err
done; done
for k in U; do
CMD="python bench.py --steps 1 --warmup 3 --frames 4 --kind $k --no-cpu --no-e2e"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_$k.csv $CMD > gpurun_out/ncu_list_$k.log 2>&1
echo "ncu list $k exit $?"
done
CMD="python bench.py --steps 1 --warmup 3 --frames 4 --kind U --no-cpu --no-e2e"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"k_apply_warp" -s 3 -c 1 -f -o gpurun_out/prof_r01_v6_apply_U $CMD > gpurun_out/ncu_full_U.log 2>&1
echo "ncu full exit $?"; tail -2 gpurun_out/ncu_full_U.log | cut -c1-200
