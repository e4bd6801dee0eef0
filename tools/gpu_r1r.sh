#!/bin/bash
# v11 closing run: full GPU suite, smoke, default bench (cpu + e2e), N, reference arm, launch list, ncu --set full of the timed step
mkdir -p gpurun_out
unset RCT_LIB
timeout 1200 python -m pytest tests -m gpu -q --tb=short --timeout 600 -p no:cacheprovider > gpurun_out/pytest_gpu.log 2>&1
echo "pytest exit $?" | tee -a gpurun_out/pytest_gpu.log; tail -6 gpurun_out/pytest_gpu.log | cut -c1-200
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke exit $?"; tail -1 gpurun_out/smoke.log
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/bench.log 2> gpurun_out/bench.err; echo "bench exit $?"; cut -c1-200 gpurun_out/bench.log
timeout 300 python bench.py --steps 5 --warmup 3 --kind N --no-cpu > gpurun_out/bench_N.log 2> gpurun_out/bench_N.err; echo "bench N exit $?"
timeout 300 python bench.py --steps 5 --warmup 3 --kind N --sty-short 512 --no-cpu > gpurun_out/bench_N_rs.log 2> gpurun_out/bench_N_rs.err; echo "bench N rs exit $?"
for wl in 1080p labels; do
  timeout 600 python bench.py --workload $wl --steps 3 --warmup 3 --no-cpu > gpurun_out/bench_$wl.log 2> gpurun_out/bench_$wl.err; echo "$wl exit $?"
done
timeout 400 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.log 2> gpurun_out/bench_ref.err; echo "ref exit $?"
CMD="python bench.py --steps 1 --warmup 3 --frames 4 --kind U --no-cpu --no-e2e"
$CMD > gpurun_out/plain_U.log 2>&1 || exit 1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_U.csv $CMD > gpurun_out/ncu_list_U.log 2>&1
echo "ncu list exit $?"
# the ncu --set full pass (about a minute of box time); skip with NO_NCU_FULL=1
if [ -z "$NO_NCU_FULL" ]; then
  timeout 1200 ncu --set full --clock-control none --import-source on -k regex:"k_" -s 45 -c 15 -f -o gpurun_out/prof_r01_full_U $CMD > gpurun_out/ncu_full_U.log 2>&1
  echo "ncu full exit $?"
fi
