#!/usr/bin/env python3
"""Device time of the SURVEY 8f row f-3 kernels on one GPU through the C ABI: the JPEG encoder on a batch of frames
(CUDA events on the launching stream, frames resident, larger than L2) and the decoder on single files with and
without restart markers (host -> device copy of the scan inside the call).  Prints one JSON object.
    python tools/bench_jpeg.py [--frames 8] [--size 2160x3840] [--quality 75]"""
import argparse
import ctypes as C
import io
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import applied_image_processing_b200 as aip  # noqa: E402
from applied_image_processing_b200 import _cabi  # noqa: E402


def peak():
    try:
        return float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        return 6650.0


def frames_like_photos(n, H, W, dev):
    """smooth gradients + noise: about 0.2 bytes per pixel at quality 75, like a rendered frame"""
    g = torch.Generator(device=dev).manual_seed(0)
    yy = torch.arange(H, device=dev)[:, None]
    xx = torch.arange(W, device=dev)[None, :]
    base = torch.stack([(xx * 5 + yy) % 256, (yy * 3 + 0 * xx) % 256, (xx + yy * 2) % 256], -1)
    out = []
    for _ in range(n):
        noise = torch.randint(-12, 12, (H, W, 3), device=dev, generator=g)
        out.append((base + noise).clamp(0, 255).to(torch.uint8))
    return torch.stack(out).contiguous()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--frames", type=int, default=8)
    ap.add_argument("--size", default="2160x3840")
    ap.add_argument("--quality", type=int, default=75)
    ap.add_argument("--reps", type=int, default=20)
    ap.add_argument("--parallel", action="store_true", help="also time the self-synchronising decoder (decode_jpeg(parallel=True))")
    a = ap.parse_args()
    H, W = (int(v) for v in a.size.split("x"))
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    lib, P = _cabi.lib(), peak()
    out = {"peak_GBps": P, "encode": {}, "decode": {}}

    # ---- encoder: n frames per call, raw C ABI so that only the kernels are inside the events
    n = a.frames
    img = frames_like_photos(n, H, W, dev)
    cap = H * W + 4096
    ws = torch.empty(lib.rct_jpeg_workspace_bytes(H, W, n), dtype=torch.uint8, device=dev)
    dst = torch.empty((n, cap), dtype=torch.uint8, device=dev)
    nbytes = torch.empty(n, dtype=torch.int64, device=dev)
    st = torch.cuda.current_stream().cuda_stream

    def enc():
        _cabi.check(lib.rct_jpeg_encode_async(img.data_ptr(), 0, n, H, W, a.quality, ws.data_ptr(), dst.data_ptr(), cap, cap,
                                              nbytes.data_ptr(), st))

    for _ in range(3):
        enc()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.reps):
        enc()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / a.reps
    sizes = nbytes.cpu().tolist()
    px = n * H * W
    alg = 3 * px + sum(sizes)                      # pixels read once + file bytes written once
    out["encode"] = {"frames": n, "H": H, "W": W, "quality": a.quality, "ms": ms, "MP_per_s": px / 1e6 / (ms * 1e-3),
                     "file_bytes": sizes[0], "alg_bytes_per_pixel": alg / px, "GBps": alg / 1e9 / (ms * 1e-3),
                     "frac_of_hbm_peak": alg / 1e9 / (ms * 1e-3) / P, "launches_per_call": 8}
    # parity spot check of what was timed: frame 0 against Pillow
    from PIL import Image
    ref = io.BytesIO()
    Image.fromarray(img[0].cpu().numpy()).save(ref, "JPEG", quality=a.quality)
    out["encode"]["byte_exact_vs_pillow"] = bool(dst[0, :sizes[0]].cpu().numpy().tobytes() == ref.getvalue())

    # ---- decoder: one 1080p file per call, three restart layouts
    a1080 = frames_like_photos(1, 1080, 1920, dev)[0].cpu().numpy()
    for name, kw in (("no_restart_markers", {}), ("restart_per_mcu_row", {"restart_marker_rows": 1}),
                     ("restart_per_8_mcus", {"restart_marker_blocks": 8})):
        buf = io.BytesIO()
        Image.fromarray(a1080).save(buf, "JPEG", quality=a.quality, **kw)
        data = buf.getvalue()
        reps = 3
        aip.decode_jpeg(data)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(reps):
            got = aip.decode_jpeg(data, device=True)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / reps
        ok = bool(np.array_equal(got.cpu().numpy(), np.asarray(Image.open(io.BytesIO(data)).convert("RGB"))))
        out["decode"][name] = {"H": 1080, "W": 1920, "file_bytes": len(data), "ms": ms,
                               "MP_per_s": 1080 * 1920 / 1e6 / (ms * 1e-3), "bit_exact_vs_pillow": ok}
    if a.parallel:
        for name, (h, w) in (("1080p", (1080, 1920)), ("4k", (2160, 3840))):
            buf = io.BytesIO()
            Image.fromarray(frames_like_photos(1, h, w, dev)[0].cpu().numpy()).save(buf, "JPEG", quality=a.quality)
            data = buf.getvalue()
            aip.decode_jpeg(data, parallel=True)
            torch.cuda.synchronize()
            e0.record()
            for _ in range(5):
                got = aip.decode_jpeg(data, device=True, parallel=True)
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / 5
            ok = bool(np.array_equal(got.cpu().numpy(), np.asarray(Image.open(io.BytesIO(data)).convert("RGB"))))
            out["decode"]["self_synchronising_" + name] = {"H": h, "W": w, "file_bytes": len(data), "ms": ms,
                                                           "MP_per_s": h * w / 1e6 / (ms * 1e-3), "bit_exact_vs_pillow": ok}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
