#!/bin/bash
# f-4 GPU parity, staging-ring A/B of the host entry point (e2e), raw PCIe floor
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q --tb=short --timeout 600 -p no:cacheprovider > gpurun_out/pytest_gpu.log 2>&1
echo "pytest exit $?" | tee -a gpurun_out/pytest_gpu.log; tail -12 gpurun_out/pytest_gpu.log
python tools/pcie_bw.py 2>&1 | tee gpurun_out/pcie.log
run() {  # sets e2e-slots kind
  timeout 300 python bench.py --steps 5 --warmup 3 --kind $3 --stage-sets $1 --e2e-slots $2 --no-cpu > gpurun_out/e2e_$3_s$1_g$2.log 2> gpurun_out/e2e_$3_s$1_g$2.err
  echo "kind $3 sets $1 e2e-slots $2 exit $? $(grep -o '"e2e": {[^}]*}' gpurun_out/e2e_$3_s$1_g$2.log)"
}
for k in U N; do
  run 2 2 $k; run 3 2 $k; run 4 2 $k; run 3 1 $k; run 4 1 $k; run 3 4 $k
done
