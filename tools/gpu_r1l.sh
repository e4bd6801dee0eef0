#!/bin/bash
# L2 fetch granularity experiment, adjacent-kernel timings, new video test
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_video.py -m gpu -q --tb=short -p no:cacheprovider > gpurun_out/pytest_video.log 2>&1
echo "pytest video exit $?"; tail -3 gpurun_out/pytest_video.log
timeout 300 python tools/bench_adjacent.py > gpurun_out/adjacent.json 2> gpurun_out/adjacent.err; echo "adjacent exit $?"; cat gpurun_out/adjacent.json; tail -3 gpurun_out/adjacent.err
for g in 0 32 128; do for k in U N; do
  if [ $g = 0 ]; then unset RCT_L2_FETCH; else export RCT_L2_FETCH=$g; fi
  timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu --no-e2e --kind $k > gpurun_out/b_${k}_l2f$g.log 2> gpurun_out/b_${k}_l2f$g.err
  echo "l2fetch $g kind $k exit $?"; grep librct gpurun_out/b_${k}_l2f$g.err | head -1
  python - <<PY
import json
d=json.loads(open("gpurun_out/b_${k}_l2f$g.log").read().strip().splitlines()[-1])
print(" ", round(d["value"]), d["roofline"]["all_kernels_ms_per_launch"])
PY
done; done
