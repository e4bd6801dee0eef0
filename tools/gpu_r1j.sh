#!/bin/bash
# memcheck of the staged-resample apply kernel; scatter (shared-memory tables) parity + timing with the rs kernel disabled
mkdir -p gpurun_out
timeout 600 compute-sanitizer --tool memcheck --print-limit 5 python -m pytest tests/test_gpu_parity.py -q -x -k "test_tma_staged_resample_apply and 200-300" -p no:cacheprovider > gpurun_out/memcheck_rs.log 2>&1
echo "memcheck exit $?"; grep -m 12 -A12 "========= Invalid\|========= Error\|Misaligned\|=========     at" gpurun_out/memcheck_rs.log | head -60
RCT_DISABLE_RS=1 timeout 1200 python -m pytest tests -m gpu -q --tb=short --timeout 600 -p no:cacheprovider > gpurun_out/pytest_gpu.log 2>&1
echo "pytest (rs disabled) exit $?"; tail -6 gpurun_out/pytest_gpu.log
for k in U N; do
timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu --no-e2e --kind $k > gpurun_out/b_${k}_scat.log 2> gpurun_out/b_${k}_scat.err
echo "bench $k exit $?"; python - <<PY
import json
d=json.loads(open("gpurun_out/b_${k}_scat.log").read().strip().splitlines()[-1])
print(round(d["value"]), d["roofline"]["all_kernels_ms_per_launch"])
PY
done
