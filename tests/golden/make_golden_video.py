#!/usr/bin/env python3
"""Generate tests/golden/video_f4.npz from the UNMODIFIED reference functions of SURVEY.md 8f row
f-4 (run in the build container only, where /root/reference exists):

    python tests/golden/make_golden_video.py

video/utils.py cannot be imported (tensorflow, tensorflow_hub and a hub.load() download at module
level), so the three function definitions -- warp_image (video/utils.py:89-105), normalize_image
(:217-221) and blend_images (:223-229) -- are taken out of the reference file's syntax tree at
generation time and executed unmodified against the cv2 / numpy / torch of this image.  Nothing
of the reference is copied into the repository."""
import ast
import contextlib
import io
import json
import os

import cv2
import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference/video/utils.py"
WANT = ("warp_image", "normalize_image", "blend_images")

CASES = {  # name -> (H, W, flow amplitude in pixels, snap flow to 1/64 px (exact rounding ties), seed)
    "small_flow": (48, 64, 3.0, False, 0),
    "reflect": (40, 56, 60.0, False, 1),
    "ties": (33, 47, 5.0, True, 2),
    "one_row": (1, 9, 3.0, False, 3),
    "one_col": (9, 1, 3.0, False, 4),
}
ALPHAS = (0.7, 0.3, 0.0, 1.0, 0.123456789)


def reference_functions():
    tree = ast.parse(open(REF).read())
    ns = {"cv2": cv2, "np": np, "torch": torch}
    for node in tree.body:
        if isinstance(node, ast.FunctionDef) and node.name in WANT:
            exec(compile(ast.Module([node], []), REF, "exec"), ns)
    return ns


def make_case(H, W, amp, snap, seed):
    rng = np.random.default_rng(seed)
    img = rng.integers(0, 256, (H, W, 3), dtype=np.uint8)
    sty = rng.integers(0, 256, (H, W, 3), dtype=np.uint8)
    flow = (rng.standard_normal((2, H, W)) * amp).astype(np.float32)
    if snap:
        flow = (np.round(flow * 64) / 64).astype(np.float32)
    return img, sty, flow


def main():
    ref = reference_functions()
    out = {}
    for name, (H, W, amp, snap, seed) in CASES.items():
        img, sty, flow = make_case(H, W, amp, snap, seed)
        with contextlib.redirect_stdout(io.StringIO()):  # the reference prints progress
            warped = ref["warp_image"](img, torch.from_numpy(flow))
        out[f"{name}.image"], out[f"{name}.stylized"], out[f"{name}.flow"] = img, sty, flow
        out[f"{name}.warped"] = warped
        for a in ALPHAS:
            out[f"{name}.blend_{a}"] = ref["blend_images"](sty, warped, a)
    np.savez_compressed(os.path.join(HERE, "video_f4.npz"), **out)
    vpath = os.path.join(HERE, "versions.json")
    v = json.load(open(vpath))
    v["opencv-python (video_f4.npz)"] = cv2.__version__
    json.dump(v, open(vpath, "w"), indent=1, sort_keys=True)
    print("wrote video_f4.npz:", len(out), "arrays; cv2", cv2.__version__)


if __name__ == "__main__":
    main()
