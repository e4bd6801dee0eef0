#!/bin/bash
# One gpurun call, parameterised by environment variables (replaces the one-off per-session scripts):
#   TESTS=1      run the GPU suite (TEST_ARGS adds pytest arguments; RCT_TEST_FAST=1 skips the live 16K oracle run)
#   SMOKE=1      __graft_entry__.smoke()
#   BENCH="..."  semicolon-separated bench.py argument lists, each run once; logs gpurun_out/bench_<i>.log
#   JPEG=1       tools/bench_jpeg.py --parallel
#   LIST="..."   bench.py arguments for one ncu launch list (gpu__time_duration), after the same command ran clean
#   FULL="..."   bench.py arguments for one ncu --set full capture of KREGEX (NSKIP / NCAP select launches)
#   TAG          suffix of the output files
mkdir -p gpurun_out
TAG=${TAG:-r02}
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/gpu_$TAG.txt 2>&1
( lscpu | egrep 'Model name|^CPU\(s\)|NUMA|Socket|Thread'; free -g | head -2; nvidia-smi topo -m ) > gpurun_out/host_$TAG.txt 2>&1
if [ "${TESTS:-0}" = 1 ]; then
  timeout ${TEST_TIMEOUT:-1500} python -m pytest tests -m gpu -q --tb=short --timeout 900 -p no:cacheprovider ${TEST_ARGS} > gpurun_out/pytest_$TAG.log 2>&1
  echo "pytest exit $?" | tee -a gpurun_out/pytest_$TAG.log; tail -${TEST_TAIL:-25} gpurun_out/pytest_$TAG.log | cut -c1-300
fi
if [ "${SMOKE:-0}" = 1 ]; then
  timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke_$TAG.log 2>&1; echo "smoke exit $?"; tail -2 gpurun_out/smoke_$TAG.log
fi
i=0
IFS=';' read -ra BL <<< "${BENCH:-}"
for args in "${BL[@]}"; do
  [ -z "$args" ] && continue
  i=$((i+1))
  timeout 900 python bench.py $args > gpurun_out/bench_${TAG}_$i.log 2> gpurun_out/bench_${TAG}_$i.err
  echo "bench [$args] exit $?"; cut -c1-${BENCH_CUT:-1800} gpurun_out/bench_${TAG}_$i.log; tail -3 gpurun_out/bench_${TAG}_$i.err | cut -c1-400
done
if [ "${JPEG:-0}" = 1 ]; then
  timeout 600 python tools/bench_jpeg.py --parallel > gpurun_out/bench_jpeg_$TAG.json 2> gpurun_out/bench_jpeg_$TAG.err
  echo "bench_jpeg exit $?"; cut -c1-1500 gpurun_out/bench_jpeg_$TAG.json; tail -3 gpurun_out/bench_jpeg_$TAG.err | cut -c1-300
fi
if [ -n "${LIST:-}" ]; then
  CMD="python bench.py $LIST"
  $CMD > gpurun_out/list_plain_$TAG.log 2>&1 &&
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c ${LIST_N:-600} --csv --log-file gpurun_out/launches_$TAG.csv $CMD > gpurun_out/list_ncu_$TAG.log 2>&1
  echo "launch list exit $?"; tail -2 gpurun_out/list_ncu_$TAG.log | cut -c1-300
fi
if [ -n "${FULL:-}" ]; then
  CMD="python bench.py $FULL"
  $CMD > gpurun_out/full_plain_$TAG.log 2>&1 &&
  timeout 1200 ncu --set full --clock-control none --import-source on -k regex:"${KREGEX:-k_apply}" -s ${NSKIP:-0} -c ${NCAP:-4} -f -o gpurun_out/prof_$TAG $CMD > gpurun_out/full_ncu_$TAG.log 2>&1
  echo "ncu full exit $?"; tail -3 gpurun_out/full_ncu_$TAG.log | cut -c1-300
fi
