#!/bin/bash
# A/B of the apply kernels (RCT_OPT_APPLY) and the match traversal (RCT_OPT_MATCH) + launch list
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q --tb=short --timeout 600 -p no:cacheprovider -x > gpurun_out/pytest_gpu.log 2>&1
echo "pytest all exit $?"; tail -8 gpurun_out/pytest_gpu.log
run() {  # kind apply match
  timeout 300 python bench.py --steps 3 --warmup 3 --kind $1 --apply $2 --match $3 --no-cpu --no-e2e > gpurun_out/bench_$1_a$2m$3.log 2> gpurun_out/bench_$1_a$2m$3.err
  echo "kind $1 apply $2 match $3 exit $?"
  grep -o '"value": [0-9.]*' gpurun_out/bench_$1_a$2m$3.log | head -1
  grep -o '"kernels": {.*"phase_share_of_step": {[^}]*}' gpurun_out/bench_$1_a$2m$3.log | cut -c1-700
  tail -2 gpurun_out/bench_$1_a$2m$3.err
}
for k in U N; do
  for ap in 1 2 4 0; do run $k $ap 1; done
  run $k 0 0
done
CMD="python bench.py --steps 1 --warmup 3 --frames 4 --kind U --no-cpu --no-e2e"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_U.csv $CMD > gpurun_out/ncu_list_U.log 2>&1
echo "ncu list exit $?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"k_apply_warp|k_match" -s 6 -c 2 -f -o gpurun_out/prof_r01_v7_U $CMD > gpurun_out/ncu_full_U.log 2>&1
echo "ncu full exit $?"; tail -2 gpurun_out/ncu_full_U.log | cut -c1-200
