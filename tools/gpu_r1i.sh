#!/bin/bash
# TMA-staged resample apply (k_apply_warp_rs) parity + A/B, 3-stage apply A/B, new e2e defaults, ncu of the apply kernels
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q --tb=short --timeout 600 -p no:cacheprovider > gpurun_out/pytest_gpu.log 2>&1
echo "pytest exit $?" | tee -a gpurun_out/pytest_gpu.log; tail -12 gpurun_out/pytest_gpu.log
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
  timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu "$@" > gpurun_out/b_$tag.log 2> gpurun_out/b_$tag.err
  echo "$tag exit $?"; pk gpurun_out/b_$tag.log; tail -2 gpurun_out/b_$tag.err
}
run U_default --kind U
run N_default --kind N
for k in U N; do
  run ${k}_a6 --kind $k --apply 6 --no-e2e
  run ${k}_rs_a1 --kind $k --sty-short 512 --apply 1
  run ${k}_rs_a0 --kind $k --sty-short 512 --apply 0 --no-e2e
done
CMD="python bench.py --steps 1 --warmup 3 --frames 4 --kind N --no-cpu --no-e2e"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"k_apply_warp" -s 3 -c 1 -f -o gpurun_out/prof_r01_v9_N_apply $CMD > gpurun_out/ncu_N_apply.log 2>&1
echo "ncu N apply exit $?"
CMD="python bench.py --steps 1 --warmup 3 --frames 4 --kind N --sty-short 512 --no-cpu --no-e2e"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"k_apply_warp_rs" -s 3 -c 1 -f -o gpurun_out/prof_r01_v9_N_rs $CMD > gpurun_out/ncu_N_rs.log 2>&1
echo "ncu N rs exit $?"
