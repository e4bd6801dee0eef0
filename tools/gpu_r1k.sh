#!/bin/bash
# bulk-store apply flavour (option 7) A/B, opt-in staged-resample kernel (option 8) parity + A/B, full GPU suite
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q --tb=short --timeout 600 -p no:cacheprovider > gpurun_out/pytest_gpu.log 2>&1
echo "pytest exit $?" | tee -a gpurun_out/pytest_gpu.log; tail -8 gpurun_out/pytest_gpu.log
pk() { python - "$1" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    r=d["roofline"]; ap=[(k,round(v["ms_per_launch"],4),round(v["frac"],3)) for k,v in r["pixel_kernels"].items()]
    print("  value",round(d["value"]),"e2e",d["e2e"] and round(d["e2e"]["value"]),"ms/step",round(d["ms_per_step"],3),ap)
except Exception as e: print("  parse failed",e)
PY
}
run() {  # tag, args...
  tag=$1; shift
  timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu --no-e2e "$@" > gpurun_out/b_$tag.log 2> gpurun_out/b_$tag.err
  echo "$tag exit $?"; pk gpurun_out/b_$tag.log; tail -2 gpurun_out/b_$tag.err
}
for k in U N; do
  run ${k}_a1 --kind $k --apply 1
  run ${k}_a7 --kind $k --apply 7
  run ${k}_rs_a8 --kind $k --sty-short 512 --apply 8
done
CMD="python bench.py --steps 1 --warmup 3 --frames 4 --kind N --apply 7 --no-cpu --no-e2e"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"k_apply_warp" -s 3 -c 1 -f -o gpurun_out/prof_r01_v9_N_apply7 $CMD > gpurun_out/ncu_N_apply7.log 2>&1
echo "ncu N apply7 exit $?"
