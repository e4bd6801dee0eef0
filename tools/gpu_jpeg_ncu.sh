#!/bin/bash
# launch list (ncu gpu__time_duration) of the JPEG encoder on 8 x 4K frames through tools/jpeg_selftest
mkdir -p gpurun_out
python - <<'PY'
import numpy as np
rng = np.random.default_rng(1)
h, w = 2160, 3840
yy, xx = np.mgrid[0:h, 0:w]
a = np.clip(np.stack([(xx * 5 + yy) % 256, (yy * 3) % 256, (xx + yy * 2) % 256], -1) + rng.integers(-12, 12, (h, w, 3)), 0, 255).astype(np.uint8)
a.tofile("/tmp/in4k.rgb")
PY
CMD="tools/jpeg_selftest --gpu 2160 3840 /tmp/in4k.rgb /tmp/o.jpg 75 8"
timeout 30 $CMD > gpurun_out/jpeg_plain.log 2>&1 || { echo plain failed; cat gpurun_out/jpeg_plain.log; exit 1; }
cat gpurun_out/jpeg_plain.log
timeout 60 ncu --metrics gpu__time_duration.sum --clock-control none -c 40 --csv --log-file gpurun_out/launches_jpeg.csv $CMD > gpurun_out/ncu_jpeg.log 2>&1
echo "ncu exit $?"
grep -c k_jpeg gpurun_out/launches_jpeg.csv
