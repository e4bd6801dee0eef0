#!/usr/bin/env python3
"""Device-time of the SURVEY 8f kernels (f-1, f-2, f-4) on one GPU: CUDA events on the launching
stream, inputs resident and larger than L2 (a ring of batches), algorithmic bytes / time against the
measured HBM peak.  Prints one JSON object."""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import applied_image_processing_b200 as aip  # noqa: E402
from applied_image_processing_b200 import video  # noqa: E402


def peak():
    try:
        return float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        return 6650.0


def timed(fn, reps=20, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def main():
    dev = torch.device("cuda", 0)
    g = torch.Generator(device=dev).manual_seed(0)
    P = peak()
    out = {"peak_GBps": P, "kernels": {}}
    # f-4: 32 frames of 1080p per launch (199 MB images + 531 MB flow: larger than L2)
    n, H, W = 32, 1080, 1920
    sty = torch.randint(0, 256, (n, H, W, 3), dtype=torch.uint8, device=dev, generator=g)
    prev = torch.randint(0, 256, (n, H, W, 3), dtype=torch.uint8, device=dev, generator=g)
    yy = torch.arange(H, device=dev, dtype=torch.float32)[:, None]
    xx = torch.arange(W, device=dev, dtype=torch.float32)[None, :]
    flow = torch.stack([torch.stack([6 * torch.sin(yy / 97 + k) + 0 * xx, 4 * torch.cos(xx / 131 - k) + 0 * yy])
                        for k in range(n)]).contiguous()
    flow += 0.25 * torch.randn(flow.shape, device=dev, generator=g)
    px = n * H * W
    for name, fn, bpp in (
        ("k_warp_blend<true>", lambda: video.warp_blend(sty, prev, flow, 0.7), 8 + 3 + 3 + 3),
        ("k_warp_blend<false>", lambda: video.warp_image(prev, flow), 8 + 3 + 3),
        ("k_blend", lambda: video.blend_images(sty, prev, 0.7), 3 + 3 + 3),
    ):
        ms = timed(fn)
        gbs = bpp * px / 1e9 / (ms * 1e-3)
        out["kernels"][name] = {"ms": ms, "MP_per_s": px / 1e6 / (ms * 1e-3), "alg_bytes_per_pixel": bpp,
                                "achieved_GBps": gbs, "frac": gbs / P, "workload": f"{n} x {W}x{H}"}
    del sty, prev, flow
    # f-1: AdaIN mask blend, 8 x 4K content, decoder output at 910x512
    n, H, W, oh, ow = 8, 2160, 3840, 512, 910
    content = torch.rand((n, 3, H, W), device=dev, generator=g)
    output = torch.rand((n, 3, oh, ow), device=dev, generator=g)
    mask = (torch.rand((H, W), device=dev, generator=g) < 0.5).float()
    ms = timed(lambda: aip.adain_mask_blend(content, output, mask))
    bpp = 12 + 12 + 4 + 12 * oh * ow / (H * W)
    gbs = bpp * n * H * W / 1e9 / (ms * 1e-3)
    out["kernels"]["k_mask_blend"] = {"ms": ms, "MP_per_s": n * H * W / 1e6 / (ms * 1e-3), "alg_bytes_per_pixel": bpp,
                                      "achieved_GBps": gbs, "frac": gbs / P, "workload": f"{n} x {W}x{H} <- {ow}x{oh}"}
    del content, output
    # f-2: 21-class logits at 4K
    logits = torch.randn((21, H, W), device=dev, generator=g)
    for name, fn, bpp in (("k_softmax_mask", lambda: aip.deeplab_background_mask(logits, 0.5, as_numpy=False), 85),
                          ("k_argmax_labels", lambda: aip.deeplab_label_map(logits), 85)):
        ms = timed(fn)
        gbs = bpp * H * W / 1e9 / (ms * 1e-3)
        out["kernels"][name] = {"ms": ms, "MP_per_s": H * W / 1e6 / (ms * 1e-3), "alg_bytes_per_pixel": bpp,
                                "achieved_GBps": gbs, "frac": gbs / P, "workload": f"21 x {W}x{H} logits (697 MB)"}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
