#!/bin/bash
# One gpurun call: GPU parity tests, smoke, bench, then ONE ncu --set full capture of the two
# per-pixel kernels (after the same command exited 0 without ncu).  Logs under gpurun_out/.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/gpu.txt 2>&1
timeout 1200 python -m pytest tests -m gpu -q --tb=short --timeout 600 -p no:cacheprovider > gpurun_out/pytest_gpu.log 2>&1
echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke exit $?" >> gpurun_out/smoke.log
tail -3 gpurun_out/smoke.log
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/bench.log 2> gpurun_out/bench.err; echo "bench exit $?"
cat gpurun_out/bench.log; tail -5 gpurun_out/bench.err
CMD="python bench.py --steps 1 --warmup 3 --frames 2 --kind ${KIND:-U} --no-cpu --no-e2e"
$CMD > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"${KREGEX:-k_hist|k_apply}" -s ${NSKIP:-6} -c ${NCAP:-2} -f -o gpurun_out/prof_${TAG:-v1} $CMD > gpurun_out/ncu_full.log 2>&1
echo "ncu exit $?"; tail -3 gpurun_out/ncu_full.log | cut -c1-300
