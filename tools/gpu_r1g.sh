#!/bin/bash
# Re-validation of HEAD: GPU parity tests (incl. the 8f adjacent rows), smoke, default bench (with
# cpu_baseline + e2e), N bench, reference arm, launch list and ncu --set full of every k_ kernel.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/gpu.txt 2>&1
timeout 1200 python -m pytest tests -m gpu -q --tb=short --timeout 600 -p no:cacheprovider > gpurun_out/pytest_gpu.log 2>&1
echo "pytest exit $?" | tee -a gpurun_out/pytest_gpu.log; tail -12 gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke exit $?" | tee -a gpurun_out/smoke.log
tail -2 gpurun_out/smoke.log
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/bench.log 2> gpurun_out/bench.err; echo "bench exit $?"
cat gpurun_out/bench.log; tail -3 gpurun_out/bench.err
timeout 300 python bench.py --steps 5 --warmup 3 --kind N --no-cpu > gpurun_out/bench_N.log 2> gpurun_out/bench_N.err; echo "bench N exit $?"
cut -c1-1200 gpurun_out/bench_N.log
timeout 400 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.log 2> gpurun_out/bench_ref.err; echo "ref exit $?"
cat gpurun_out/bench_ref.log
for K in U N; do
CMD="python bench.py --steps 1 --warmup 3 --frames 4 --kind $K --no-cpu --no-e2e"
$CMD > gpurun_out/plain_$K.log 2>&1 || { echo "plain $K failed"; continue; }
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_$K.csv $CMD > gpurun_out/ncu_list_$K.log 2>&1
echo "ncu list $K exit $?"
done
CMD="python bench.py --steps 1 --warmup 3 --frames 4 --kind U --no-cpu --no-e2e"
N1=$(grep -c '"k_\|rct::k_' gpurun_out/launches_U.csv)
echo "k launches in list: $N1"
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:"k_" -s ${NSKIP:-45} -c ${NCAP:-15} -f -o gpurun_out/prof_r01_v8_U $CMD > gpurun_out/ncu_full_U.log 2>&1
echo "ncu full exit $?"; tail -2 gpurun_out/ncu_full_U.log | cut -c1-200
