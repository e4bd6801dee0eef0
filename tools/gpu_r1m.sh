#!/bin/bash
# palette LUT (RCT_OPT_LUT) parity + A/B, vectorised warp/blend + softmax fast path, ncu of the U apply kernel
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q --tb=short --timeout 600 -p no:cacheprovider > gpurun_out/pytest_gpu.log 2>&1
echo "pytest exit $?" | tee -a gpurun_out/pytest_gpu.log; tail -8 gpurun_out/pytest_gpu.log
timeout 300 python tools/bench_adjacent.py > gpurun_out/adjacent.json 2> gpurun_out/adjacent.err; echo "adjacent exit $?"
python - <<'PY'
import json
d=json.load(open('gpurun_out/adjacent.json'))
for k,v in d['kernels'].items(): print(" ",k, round(v['ms'],3), round(v['achieved_GBps']), round(v['frac'],3))
PY
for k in U N; do for l in 1 0; do
  timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu --no-e2e --kind $k --lut $l > gpurun_out/b_${k}_lut$l.log 2> gpurun_out/b_${k}_lut$l.err
  echo "kind $k lut $l exit $?"
  python - <<PY
import json
d=json.loads(open("gpurun_out/b_${k}_lut$l.log").read().strip().splitlines()[-1])
print(" ", round(d["value"]), d["roofline"]["all_kernels_ms_per_launch"])
PY
done; done
CMD="python bench.py --steps 1 --warmup 3 --frames 4 --kind U --no-cpu --no-e2e"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"k_apply_warp|k_pal|k_match" -s 18 -c 6 -f -o gpurun_out/prof_r01_v10_U_pal $CMD > gpurun_out/ncu_U_pal.log 2>&1
echo "ncu exit $?"
