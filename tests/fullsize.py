"""Full-size parity cases shared by tests/golden/make_golden_fullsize.py (which runs the live
reference on them) and the tests (which run the oracle / the CUDA path on them): the sizes the
metric is quoted on (BASELINE.json configs 2, 3, 5) and config 1 literally."""
import hashlib
import io
import json
import os

import numpy as np

from tests import synth

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")

# seeds of the 4K cases = frame 0 of bench.py's default batch (make_frames(a, n, 1000) on rank 0)
CASES = {
    "fs_1080p_U": dict(H=1080, W=1920, sh=1080, sw=1920, kind="U", cseed=3000, sseed=3001, mseed=3002),
    "fs_1080p_N_sty512": dict(H=1080, W=1920, sh=512, sw=910, kind="N", cseed=3010, sseed=3011, mseed=3012),
    "fs_4k_U": dict(H=2160, W=3840, sh=2160, sw=3840, kind="U", cseed=1000, sseed=1001, mseed=1000),
    "fs_4k_N": dict(H=2160, W=3840, sh=2160, sw=3840, kind="N", cseed=1000, sseed=1001, mseed=1000),
    "fs_1080p_labels8": dict(H=1080, W=1920, sh=1080, sw=1920, kind="N", cseed=3020, sseed=3021, mseed=3022, labels=8),
}


def inputs(spec):
    c = synth.synth_image(spec["H"], spec["W"], spec["cseed"], spec["kind"])
    s = synth.synth_image(spec["sh"], spec["sw"], spec["sseed"], spec["kind"])
    if spec.get("labels"):
        m = synth.voronoi_labels(spec["H"], spec["W"], spec["labels"], spec["mseed"])
    else:
        m = synth.ellipse_mask(spec["H"], spec["W"], 0.4, spec["mseed"])
    return c, s, m


def inputs_16k():
    """BASELINE config 4: the image bench.py --workload 16k runs (run_gpu_16k)."""
    H, W = 8640, 15360
    return synth.synth_image(H, W, 2000, "U"), synth.synth_image(H, W, 2001, "U"), synth.ellipse_mask(H, W, 0.4, 2002)


def golden():
    with open(os.path.join(GOLDEN_DIR, "fullsize.json")) as fh:
        return json.load(fh)


def sha256(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def unhex(v):
    return np.array([float.fromhex(x) for x in v])


def config1_cases(dec):
    """dec: decoded RGB arrays keyed like config1.npz.  Yields (tag, content, stylised, mask)."""
    from PIL import Image

    brad, brush, cat = dec["brad_pitt_jpg"], dec["brushstrokes_jpg"], dec["cat_jpg"]
    assert brad.shape == (512, 512, 3) and brush.shape == (700, 933, 3) and cat.shape == (485, 728, 3)
    # (i) brad_pitt 512x512 + brushstrokes at its own 933x700: LST:223-230 NEAREST-resizes it to the mask
    yield "brad", brad, brush, synth.ellipse_mask(512, 512, 0.42, 4)
    # (ii) cat 728x485 with a 768x512 stylised image (AdaIN's Resize(512) scale): 768->728, 512->485
    sty = np.array(Image.fromarray(brush).resize((768, 512), Image.NEAREST))
    yield "cat", cat, sty, synth.ellipse_mask(485, 728, 0.35, 5)


def load_config1():
    from PIL import Image

    g = dict(np.load(os.path.join(GOLDEN_DIR, "config1.npz"), allow_pickle=False))
    dec = {k: np.array(Image.open(io.BytesIO(g[k].tobytes())).convert("RGB"))
           for k in ("brad_pitt_jpg", "brushstrokes_jpg", "cat_jpg")}
    return [(tag, np.ascontiguousarray(c), np.ascontiguousarray(s), m, g[f"{tag}_combined"])
            for tag, c, s, m in config1_cases(dec)], g


# color_transfer_foreground with fg.shape != bg.shape: tag -> (fg H, fg W, bg H, bg W, fg kind, bg kind)
CTF_SHAPE_CASES = {
    "bg_smaller": (40, 50, 23, 31, "U", "N"),
    "bg_larger": (20, 30, 64, 48, "N", "U"),
    "same_count_other_shape": (24, 36, 36, 24, "U", "U"),
    "bg_much_larger": (17, 19, 200, 320, "N", "N"),
}


def ctf_inputs(fh, fw, bh, bw, fk, bk):
    return synth.synth_image(fh, fw, 41, fk, 0.05), synth.synth_image(bh, bw, 42, bk, 0.02)


def load_ctf_shapes():
    return dict(np.load(os.path.join(GOLDEN_DIR, "ctf_shapes.npz"), allow_pickle=False))
