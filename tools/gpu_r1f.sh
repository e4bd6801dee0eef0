#!/bin/bash
# apply kernel: L2 eviction hints / 5 CTAs per SM A/B; new per-kernel timing in bench.py
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q --tb=short --timeout 600 -p no:cacheprovider -x > gpurun_out/pytest_gpu.log 2>&1
echo "pytest all exit $?"; tail -8 gpurun_out/pytest_gpu.log
run() {  # kind apply
  timeout 300 python bench.py --steps 3 --warmup 3 --kind $1 --apply $2 --no-cpu --no-e2e > gpurun_out/bench_$1_a$2.log 2> gpurun_out/bench_$1_a$2.err
  echo "kind $1 apply $2 exit $?"
  grep -o '"value": [0-9.]*' gpurun_out/bench_$1_a$2.log | head -1
  grep -o '"pixel_kernels": {.*"all_kernels_ms_per_launch": {[^}]*}' gpurun_out/bench_$1_a$2.log | cut -c1-1500
  tail -2 gpurun_out/bench_$1_a$2.err
}
for k in U N; do for ap in 1 5 2 0; do run $k $ap; done; done
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/bench.log 2> gpurun_out/bench.err; echo "bench exit $?"
cat gpurun_out/bench.log; tail -5 gpurun_out/bench.err
