#!/bin/bash
# First GPU call of the next session: everything that was written after the GPU budget of round 1 ran out.
#   1. the JPEG decoder through the C ABI (tests/test_zz_jpeg_decode_gpu.py: kernels already bit-exact via the checker)
#   2. the whole GPU suite, smoke, the default bench
#   3. tools/bench_jpeg.py (encoder roofline line, decoder times) and the RCT_JDB=4 build of the encoder on all geometries
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zz_jpeg_decode_gpu.py -m gpu -q --tb=short -p no:cacheprovider > gpurun_out/pytest_jpeg_decode.log 2>&1
echo "decode pytest exit $?"; tail -3 gpurun_out/pytest_jpeg_decode.log | cut -c1-200
timeout 1500 python -m pytest tests -m gpu -q --tb=short --timeout 600 -p no:cacheprovider > gpurun_out/pytest_gpu.log 2>&1
echo "pytest exit $?"; tail -4 gpurun_out/pytest_gpu.log | cut -c1-200
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke exit $?"
timeout 300 python tools/bench_jpeg.py > gpurun_out/bench_jpeg.json 2> gpurun_out/bench_jpeg.err; echo "bench_jpeg exit $?"; cut -c1-400 gpurun_out/bench_jpeg.json
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/bench.log 2> gpurun_out/bench.err; echo "bench exit $?"; cut -c1-200 gpurun_out/bench.log
