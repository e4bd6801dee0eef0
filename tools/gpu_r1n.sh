#!/bin/bash
# round-1 closing run on one GPU: full GPU suite, smoke, default bench (U) + N + the other workloads, adjacent kernels,
# launch list + ncu --set full of the timed step (U)
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q --tb=short --timeout 600 -p no:cacheprovider > gpurun_out/pytest_gpu.log 2>&1
echo "pytest exit $?" | tee -a gpurun_out/pytest_gpu.log; tail -6 gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke exit $?"; tail -1 gpurun_out/smoke.log
timeout 300 python tools/bench_adjacent.py > gpurun_out/adjacent.json 2> gpurun_out/adjacent.err; echo "adjacent exit $?"
python - <<'PY'
import json
d=json.load(open('gpurun_out/adjacent.json'))
for k,v in d['kernels'].items(): print(" ",k, round(v['ms'],3), round(v['achieved_GBps']), round(v['frac'],3))
PY
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/bench.log 2> gpurun_out/bench.err; echo "bench exit $?"; cut -c1-700 gpurun_out/bench.log
timeout 300 python bench.py --steps 5 --warmup 3 --kind N --no-cpu > gpurun_out/bench_N.log 2> gpurun_out/bench_N.err; echo "bench N exit $?"; cut -c1-300 gpurun_out/bench_N.log
timeout 300 python bench.py --steps 5 --warmup 3 --kind N --sty-short 512 --no-cpu > gpurun_out/bench_N_rs.log 2> gpurun_out/bench_N_rs.err; echo "bench N rs exit $?"; cut -c1-300 gpurun_out/bench_N_rs.log
for wl in 1080p labels 16k; do
  timeout 600 python bench.py --workload $wl --steps 3 --warmup 3 --no-cpu > gpurun_out/bench_$wl.log 2> gpurun_out/bench_$wl.err
  echo "$wl exit $?"; cut -c1-260 gpurun_out/bench_$wl.log; grep -o '"e2e": {[^}]*}' gpurun_out/bench_$wl.log; tail -2 gpurun_out/bench_$wl.err
done
CMD="python bench.py --steps 1 --warmup 3 --frames 4 --kind U --no-cpu --no-e2e"
$CMD > gpurun_out/plain_U.log 2>&1 && timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_U.csv $CMD > gpurun_out/ncu_list_U.log 2>&1
echo "ncu list exit $?"
