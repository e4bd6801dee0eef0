#!/bin/bash
# fused label pass parity + A/B, gather dedupe on the resampled workload
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q --tb=short --timeout 600 -p no:cacheprovider > gpurun_out/pytest_gpu.log 2>&1
echo "pytest exit $?" | tee -a gpurun_out/pytest_gpu.log; tail -12 gpurun_out/pytest_gpu.log | cut -c1-300
pk() { python - "$1" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    r=d["roofline"]; ap=[(k,round(v["ms_per_launch"],4),round(v["frac"],3)) for k,v in r["pixel_kernels"].items()]
    print("  value",round(d["value"]),"e2e",d["e2e"] and round(d["e2e"]["value"]),"ms/step",round(d["ms_per_step"],3),"launches",d["gpu_launches"],ap)
    print("  ",r["all_kernels_ms_per_launch"])
except Exception as e: print("  parse failed",e)
PY
}
run() { tag=$1; shift; timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu "$@" > gpurun_out/b_$tag.log 2> gpurun_out/b_$tag.err; echo "$tag exit $?"; pk gpurun_out/b_$tag.log; tail -2 gpurun_out/b_$tag.err; }
run labels_fused --workload labels --labels-fused 1
run labels_seq --workload labels --labels-fused 0 --no-e2e
run labels_fused_U --workload labels --labels-fused 1 --kind U --no-e2e
run N_rs --kind N --sty-short 512 --no-e2e
run U_rs --kind U --sty-short 512 --no-e2e
