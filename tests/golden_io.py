"""Loader for the committed reference outputs under tests/golden/ (see make_golden.py)."""
import json
import os

import numpy as np

from tests import synth

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")

PIPELINE_NAMES = [
    "pipeline_u_same", "pipeline_u_resample", "pipeline_n_bernoulli", "pipeline_fg_small",
    "pipeline_fg_large", "pipeline_equal_counts", "real_crop",
]


def _inputs_from_spec(spec):
    black = spec.get("black", 0.001)
    c = synth.synth_image(spec["H"], spec["W"], spec["ck"][1], spec["ck"][0], black)
    s = synth.synth_image(spec["sh"], spec["sw"], spec["sk"][1], spec["sk"][0], black)
    m = spec["mask"]
    if m[0] == "ellipse":
        mask = synth.ellipse_mask(spec["H"], spec["W"], m[1], m[2])
    elif m[0] == "bernoulli":
        mask = synth.bernoulli_mask(spec["H"], spec["W"], m[1], m[2])
    else:
        mask = np.zeros((spec["H"], spec["W"]), np.uint8)
        mask[:, spec["W"] // 2:] = 1
    return c, s, mask


def load_pipeline(name):
    """-> (content, stylised, background_mask, golden dict)."""
    g = dict(np.load(os.path.join(GOLDEN_DIR, name + ".npz"), allow_pickle=False))
    if name == "real_crop":
        c, s = g["content"], g["stylised"]
        mask = np.unpackbits(g["mask"])[: c.shape[0] * c.shape[1]].reshape(c.shape[:2])
    else:
        c, s, mask = _inputs_from_spec(json.loads(str(g["spec"])))
    return np.ascontiguousarray(c), np.ascontiguousarray(s), np.ascontiguousarray(mask), g


def load(name):
    return dict(np.load(os.path.join(GOLDEN_DIR, name + ".npz"), allow_pickle=False))
