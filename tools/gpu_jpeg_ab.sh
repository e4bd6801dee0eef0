#!/bin/bash
# A/B of the JPEG encoder builds (tools/jpeg_selftest_v*: -DRCT_JDB variants) on 8 x 4K frames, then the GPU tests
# through the C ABI, the byte-exactness sweep and (time permitting) the ncu launch list.
mkdir -p gpurun_out
python - <<'PY'
import numpy as np
rng = np.random.default_rng(1)
h, w = 2160, 3840
yy, xx = np.mgrid[0:h, 0:w]
a = np.clip(np.stack([(xx * 5 + yy) % 256, (yy * 3) % 256, (xx + yy * 2) % 256], -1) + rng.integers(-12, 12, (h, w, 3)), 0, 255).astype(np.uint8)
a.tofile("/tmp/in4k.rgb")
PY
# builds from tools/build_jpeg_variants.sh; the reference bytes come from Pillow
python - <<'PY'
import numpy as np
from PIL import Image
a = np.fromfile("/tmp/in4k.rgb", np.uint8).reshape(2160, 3840, 3)
Image.fromarray(a).save("/tmp/ref4k.jpg")
PY
for exe in tools/jpeg_selftest_v*; do
  [ -x $exe ] || continue
  v=${exe#tools/jpeg_selftest_}
  echo -n "$v: "; timeout 20 $exe --gpu 2160 3840 /tmp/in4k.rgb /tmp/o_$v.jpg 75 8 2>&1 | tail -1
  cmp -s /tmp/o_$v.jpg /tmp/ref4k.jpg && echo "   same bytes as Pillow" || echo "   DIFFERS from Pillow"
done 2>&1 | tee gpurun_out/jpeg_ab.log
timeout 40 python -m pytest tests/test_jpeg_gpu.py -m gpu -x -q -p no:cacheprovider > gpurun_out/jpeg_pytest.log 2>&1; echo "pytest exit $?"; tail -2 gpurun_out/jpeg_pytest.log
echo "elapsed $SECONDS"
timeout 30 python tools/gpu_jpeg_check.py --gpu > gpurun_out/jpeg_gpu.log 2>&1; echo "sweep exit $?"; tail -1 gpurun_out/jpeg_gpu.log
echo "elapsed $SECONDS"
if [ $SECONDS -lt 26 ]; then
  timeout 20 ncu --metrics gpu__time_duration.sum --clock-control none -c 40 --csv --log-file gpurun_out/launches_jpeg.csv tools/jpeg_selftest --gpu 2160 3840 /tmp/in4k.rgb /tmp/o.jpg 75 8 > gpurun_out/ncu_jpeg.log 2>&1
  echo "ncu exit $?"
fi
