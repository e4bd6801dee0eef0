#!/bin/bash
# compile-time tuning A/B: part size (RCT_PARTT) and scatter entries per thread (RCT_SCATPER)
mkdir -p gpurun_out
for v in ${VARIANTS:-"" _p3072}; do
  export RCT_LIB=$PWD/applied_image_processing_b200/librct$v.so
  echo "== variant '$v'"
  timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "pipeline_vs or integer_artefacts or sort_global or match_traversals or 4k_properties or front_ends" -p no:cacheprovider 2>&1 | tail -2
  for k in U N; do
    timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu --no-e2e --kind $k > gpurun_out/b_${k}_var$v.log 2> gpurun_out/b_${k}_var$v.err
    python - <<PY
import json
d=json.loads(open("gpurun_out/b_${k}_var$v.log").read().strip().splitlines()[-1])
m=d["roofline"]["all_kernels_ms_per_launch"]
print("  $k", round(d["value"]), {x:m[x] for x in ("k_keys","k_front_part","k_front_count","k_fine_hist","k_apply")})
PY
  done
done
