#!/usr/bin/env python3
"""Decode one photo-like 4K JPEG (no restart markers) a few times through decode_jpeg: the workload for an ncu launch
list of the decoder kernels (`ncu --metrics gpu__time_duration.sum -k regex:k_j ...`)."""
import io
import os
import sys
import time

import numpy as np
import torch
from PIL import Image

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import applied_image_processing_b200 as aip  # noqa: E402

h, w = 2160, 3840
rng = np.random.default_rng(1)
yy, xx = np.mgrid[0:h, 0:w]
a = np.clip(np.stack([(xx * 5 + yy) % 256, (yy * 3) % 256, (xx + yy * 2) % 256], -1) + rng.integers(-12, 12, (h, w, 3)), 0, 255).astype(np.uint8)
buf = io.BytesIO()
Image.fromarray(a).save(buf, "JPEG")
data = buf.getvalue()
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 3
for _ in range(reps):
    t0 = time.perf_counter()
    px = aip.decode_jpeg(data, device=True)
    torch.cuda.synchronize()
    print("decode %.3f ms, %d bytes" % (1e3 * (time.perf_counter() - t0), len(data)))
