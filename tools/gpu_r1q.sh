#!/bin/bash
# launch-configuration variants, bench only (no tests): VARIANTS="_a _b"
mkdir -p gpurun_out
for v in ${VARIANTS:-""}; do
  [ "$v" = "base" ] && v=""
  export RCT_LIB=$PWD/applied_image_processing_b200/librct$v.so
  echo "== variant '$v'"
  for k in U N; do
    timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu --no-e2e --kind $k > gpurun_out/b_${k}_var$v.log 2> gpurun_out/b_${k}_var$v.err
    python - <<PY
import json
d=json.loads(open("gpurun_out/b_${k}_var$v.log").read().strip().splitlines()[-1])
m=d["roofline"]["all_kernels_ms_per_launch"]
print("  $k", round(d["value"]), {x:m[x] for x in ("k_scatter","k_part_sort","k_match")})
PY
  done
done
