#!/bin/bash
# new front end + bulk-async apply: targeted parity tests first (hang-guarded), then the full
# GPU suite, A/B bench runs and a launch list.
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q --tb=short --timeout 240 -p no:cacheprovider -x \
  -k "front_ends or partition_front or batch_apply" > gpurun_out/pytest_new.log 2>&1
echo "pytest new exit $?"; tail -25 gpurun_out/pytest_new.log
timeout 1200 python -m pytest tests -m gpu -q --tb=short --timeout 600 -p no:cacheprovider > gpurun_out/pytest_gpu.log 2>&1
echo "pytest all exit $?"; tail -8 gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke exit $?"; tail -2 gpurun_out/smoke.log
for k in U N; do for cfg in "1 1" "0 1" "1 0" "0 0"; do
  set -- $cfg
  timeout 300 python bench.py --steps 3 --warmup 3 --kind $k --front $1 --apply $2 --no-cpu --no-e2e > gpurun_out/bench_${k}_f$1a$2.log 2> gpurun_out/bench_${k}_f$1a$2.err
  echo "kind $k front $1 apply $2 exit $?"
  grep -o '"value": [0-9.]*' gpurun_out/bench_${k}_f$1a$2.log | head -1
  grep -o '"kernels": {.*"phase_share_of_step": {[^}]*}' gpurun_out/bench_${k}_f$1a$2.log
  tail -2 gpurun_out/bench_${k}_f$1a$2.err
done; done
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/bench.log 2> gpurun_out/bench.err; echo "bench exit $?"
cut -c1-1800 gpurun_out/bench.log; tail -5 gpurun_out/bench.err
timeout 300 python bench.py --workload 1080p --steps 3 --warmup 3 --no-cpu --no-e2e > gpurun_out/bench_1080p.log 2>&1; echo "1080p exit $?"; grep -o '"value": [0-9.]*' gpurun_out/bench_1080p.log | head -1
for k in U N; do
CMD="python bench.py --steps 1 --warmup 3 --frames 4 --kind $k --no-cpu --no-e2e"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_$k.csv $CMD > gpurun_out/ncu_list_$k.log 2>&1
echo "ncu list $k exit $?"
done
