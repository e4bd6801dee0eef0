#!/bin/bash
# A/B of tuning builds in ONE gpurun call: every build/variants/librct_<name>.so (built with -D knobs, see the RCT_*
# macros in csrc/) runs the same short bench through RCT_LIB; one line per variant with the step time and the kernels
# named in KERNELS.  ARGS: bench.py arguments.  The in-tree librct.so runs first as "default".
mkdir -p gpurun_out
TAG=${TAG:-var}
ARGS=${ARGS:---gpus 1 --steps 10 --warmup 3 --no-cpu --no-e2e --files 0}
KERNELS=${KERNELS:-k_match k_lab_moments}
run() {
  name=$1; lib=$2
  RCT_LIB=$lib timeout 300 python bench.py $ARGS > gpurun_out/var_${TAG}_$name.log 2> gpurun_out/var_${TAG}_$name.err
  rc=$?
  python - "$name" "$rc" gpurun_out/var_${TAG}_$name.log $KERNELS <<'PY'
import json, sys
name, rc, path, kernels = sys.argv[1], sys.argv[2], sys.argv[3], sys.argv[4:]
try:
    b = json.loads(open(path).read().strip().split("\n")[-1])
    ms = b["roofline"]["all_kernels_ms_per_launch"]
    print(name, "rc", rc, "MP/s %.0f" % b["value"], "ms/step %.3f" % b["ms_per_step"], " ".join("%s %.3f" % (k, ms.get(k, -1)) for k in kernels))
except Exception as e:
    print(name, "rc", rc, "failed:", e)
PY
}
run default ""
for lib in build/variants/librct_*.so; do
  name=$(basename $lib .so); name=${name#librct_}
  run $name $PWD/$lib
done
