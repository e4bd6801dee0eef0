#!/bin/bash
# First GPU call of the next session: everything that was written after the GPU budget of round 1 ran out.
#   1. the JPEG decoder through the C ABI (tests/test_zz_jpeg_decode_gpu.py: kernels already bit-exact via the checker)
#   2. the whole GPU suite, smoke, the default bench
#   3. tools/bench_jpeg.py (encoder roofline line, decoder times) and the RCT_JDB=4 build of the encoder on all geometries
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zz_jpeg_decode_gpu.py -m gpu -q --tb=short -p no:cacheprovider > gpurun_out/pytest_jpeg_decode.log 2>&1
echo "decode pytest exit $?"; tail -3 gpurun_out/pytest_jpeg_decode.log | cut -c1-200
# the self-synchronising decoder: first through the stand-alone checker, then the gated test
python - <<'PY'
import io, numpy as np
from PIL import Image
rng = np.random.default_rng(1)
for name, (h, w) in {"540": (540, 960), "2160": (2160, 3840)}.items():
    yy, xx = np.mgrid[0:h, 0:w]
    a = np.clip(np.stack([(xx * 5 + yy) % 256, (yy * 3) % 256, (xx + yy * 2) % 256], -1) + rng.integers(-12, 12, (h, w, 3)), 0, 255).astype(np.uint8)
    Image.fromarray(a).save(f"/tmp/p{name}.jpg"); np.asarray(Image.open(f"/tmp/p{name}.jpg").convert("RGB")).tofile(f"/tmp/p{name}.ref")
PY
for n in 540 2160; do
  timeout 120 tools/jpeg_selftest --gpu-decode /tmp/p$n.jpg /tmp/p$n.raw par 2>&1 | tail -2; cmp /tmp/p$n.raw /tmp/p$n.ref && echo "parallel decode $n: bit-exact"
done 2>&1 | tee gpurun_out/jpeg_decode_par.log
RCT_TEST_UNVALIDATED=1 timeout 300 python -m pytest tests/test_zz_jpeg_decode_gpu.py -m gpu -q --tb=short -p no:cacheprovider -k self_synchronising > gpurun_out/pytest_jpeg_decode_par.log 2>&1
echo "parallel decode pytest exit $?"; tail -3 gpurun_out/pytest_jpeg_decode_par.log | cut -c1-200
timeout 1500 python -m pytest tests -m gpu -q --tb=short --timeout 600 -p no:cacheprovider > gpurun_out/pytest_gpu.log 2>&1
echo "pytest exit $?"; tail -4 gpurun_out/pytest_gpu.log | cut -c1-200
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke exit $?"
timeout 300 python tools/bench_jpeg.py --parallel > gpurun_out/bench_jpeg.json 2> gpurun_out/bench_jpeg.err; echo "bench_jpeg exit $?"; cut -c1-400 gpurun_out/bench_jpeg.json
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/bench.log 2> gpurun_out/bench.err; echo "bench exit $?"; cut -c1-200 gpurun_out/bench.log
