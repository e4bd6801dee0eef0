#!/usr/bin/env python3
"""Generate tests/golden/*.npz from the UNMODIFIED reference (run in the build container only,
where /root/reference exists; the GPU box uses the committed files).

    python tests/golden/make_golden.py

Every fixture stores the exact inputs (or the seeds that regenerate them through
tests/synth.py) and the outputs of the reference's own functions, imported through
oracle/ref_shim.py:

  pipeline_*.npz   ref_pipeline(content, stylised, background_mask)  = LST:222-238 verbatim
                   + the intermediate values of color_transfer_foreground (LST:128-168):
                   fg/bg pixel counts, PCA mean_/components_, and -- for a strided sample of
                   foreground pixels -- lab, projection, matched value, float RGB before
                   quantisation.
  functions.npz    rgb_to_lab_pixels / lab_to_rgb_pixels / apply_pca / match_cdf on small
                   seeded arrays (all three length cases of match_cdf).
  resample.npz     Pillow Image.NEAREST source-index tables for a list of size pairs.
  real_crop.npz    a 256x256 crop of input/content/brad_pitt.jpg with a 192x256 crop of
                   input/style/brushstrokes.jpg as the stand-in stylised image (BASELINE.json
                   config 1 at reduced size; the DeepLab mask and the AdaIN output cannot be
                   produced offline).
Library versions used are recorded in versions.json.
"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from oracle import ref_shim  # noqa: E402
from tests import synth  # noqa: E402

# name -> (H, W, sh, sw, content kind/seed, stylised kind/seed, mask spec)
PIPELINE_CASES = {
    "pipeline_u_same": dict(H=96, W=128, sh=96, sw=128, ck=("U", 0), sk=("U", 1), mask=("ellipse", 0.4, 4)),
    "pipeline_u_resample": dict(H=135, W=240, sh=100, sw=150, ck=("U", 10), sk=("N", 11), mask=("ellipse", 0.4, 4)),
    "pipeline_n_bernoulli": dict(H=121, W=182, sh=128, sw=192, ck=("N", 2), sk=("N", 3), mask=("bernoulli", 0.5, 4)),
    "pipeline_fg_small": dict(H=90, W=160, sh=90, sw=160, ck=("N", 20), sk=("U", 21), mask=("ellipse", 0.1, 6)),
    "pipeline_fg_large": dict(H=90, W=160, sh=45, sw=80, ck=("U", 22), sk=("N", 23), mask=("ellipse", 0.8, 7)),
    "pipeline_equal_counts": dict(H=64, W=64, sh=64, sw=64, ck=("U", 30), sk=("U", 31), mask=("half",), black=0.0),
}


def build_case(spec):
    black = spec.get("black", 0.001)
    c = synth.synth_image(spec["H"], spec["W"], spec["ck"][1], spec["ck"][0], black)
    s = synth.synth_image(spec["sh"], spec["sw"], spec["sk"][1], spec["sk"][0], black)
    m = spec["mask"]
    if m[0] == "ellipse":
        mask = synth.ellipse_mask(spec["H"], spec["W"], m[1], m[2])
    elif m[0] == "bernoulli":
        mask = synth.bernoulli_mask(spec["H"], spec["W"], m[1], m[2])
    else:  # left half foreground, right half background: N_fg == N_bg when no black pixels
        mask = np.zeros((spec["H"], spec["W"]), np.uint8)
        mask[:, spec["W"] // 2:] = 1
    return c, s, mask


def trace_reference(lst, content, stylised, mask):
    """LST:222-238 + the intermediates of LST:128-168, all computed by the reference's own
    functions (nothing re-implemented here)."""
    from PIL import Image

    if stylised.shape[:2] != mask.shape:
        stylised = np.array(Image.fromarray(stylised).resize((mask.shape[1], mask.shape[0]), Image.NEAREST))
    fgm = 1 - mask
    fg = content * fgm[..., None]
    bg = stylised * mask[..., None]
    out = {"resized": stylised}
    fsel = fg.sum(-1) > 0
    bsel = bg.sum(-1) > 0
    out["n_fg"], out["n_bg"] = int(fsel.sum()), int(bsel.sum())
    if out["n_fg"] and out["n_bg"]:
        fg_lab = lst.rgb_to_lab_pixels(fg[fsel])
        bg_lab = lst.rgb_to_lab_pixels(bg[bsel])
        fp, fpca = lst.apply_pca(fg_lab)
        bp, bpca = lst.apply_pca(bg_lab)
        matched = lst.match_cdf(fp, bp)
        adj_lab = fpca.inverse_transform(matched)
        # float RGB before quantisation: LST:84-88 without the final cast
        log_lms = np.dot(adj_lab, lst.LAB_TO_LMS.T)
        rgbf = np.clip(np.dot(np.power(10.0, log_lms), lst.LMS_TO_RGB.T), 0, 1)
        idx = np.arange(0, out["n_fg"], max(1, out["n_fg"] // 512))
        flat = np.flatnonzero(fsel.ravel())[idx]
        out.update(
            fg_mean=fpca.mean_, fg_comp=fpca.components_[0], bg_mean=bpca.mean_, bg_comp=bpca.components_[0],
            sample_pixel=flat.astype(np.int64), sample_lab=fg_lab[idx], sample_proj=fp[idx, 0],
            sample_matched=matched[idx, 0], sample_rgbf=rgbf[idx],
        )
    out["combined"] = ref_shim.ref_pipeline(content, stylised, mask)
    return out


def main():
    lst = ref_shim.load_reference()
    for name, spec in PIPELINE_CASES.items():
        c, s, mask = build_case(spec)
        tr = trace_reference(lst, c, s, mask)
        np.savez_compressed(os.path.join(HERE, name + ".npz"), spec=json.dumps(spec), **tr)
        print(name, c.shape, s.shape, "n_fg", tr["n_fg"], "n_bg", tr["n_bg"])

    # ---- real image crop (config-1 stand-in) ------------------------------------------------
    from PIL import Image

    content = np.array(Image.open(os.path.join(ref_shim.REFERENCE_ROOT, "input/content/brad_pitt.jpg")).convert("RGB"))
    style = np.array(Image.open(os.path.join(ref_shim.REFERENCE_ROOT, "input/style/brushstrokes.jpg")).convert("RGB"))
    c = np.ascontiguousarray(content[128:384, 128:384])
    s = np.ascontiguousarray(style[200:392, 300:556])
    mask = synth.ellipse_mask(256, 256, 0.42, 4)
    tr = trace_reference(lst, c, s, mask)
    np.savez_compressed(os.path.join(HERE, "real_crop.npz"), content=c, stylised=s, mask=np.packbits(mask),
                        **tr)
    print("real_crop n_fg", tr["n_fg"], "n_bg", tr["n_bg"])

    # ---- stand-alone functions ----------------------------------------------------------------
    rng = np.random.default_rng(100)
    px = rng.integers(0, 256, (2048, 3), dtype=np.uint8)
    px[:4] = [[0, 0, 0], [255, 255, 255], [1, 0, 0], [0, 0, 1]]
    lab = lst.rgb_to_lab_pixels(px)
    back = lst.lab_to_rgb_pixels(lab)
    lab2 = lab + rng.normal(0, 0.05, lab.shape)
    back2 = lst.lab_to_rgb_pixels(lab2)
    proj, pca = lst.apply_pca(lab)
    fn = dict(px=px, lab=lab, back=back, lab2=lab2, back2=back2, proj=proj, pca_mean=pca.mean_,
              pca_comp=pca.components_[0], inv=pca.inverse_transform(proj))
    for tag, nt, ns in (("eq", 700, 700), ("tgt_longer", 900, 411), ("src_longer", 333, 1000), ("one_src", 50, 1),
                        ("one_tgt", 1, 40)):
        t = rng.normal(0, 1, (nt, 1))
        s_ = rng.normal(0.3, 2, (ns, 1))
        if tag == "eq":
            t[10] = t[20]  # a tie
        fn[f"cdf_{tag}_t"], fn[f"cdf_{tag}_s"], fn[f"cdf_{tag}_m"] = t, s_, lst.match_cdf(t, s_)
    np.savez_compressed(os.path.join(HERE, "functions.npz"), **fn)

    # ---- Pillow NEAREST tables ----------------------------------------------------------------
    pairs = [(512, 485), (768, 728), (100, 135), (150, 240), (933, 512), (700, 512), (45, 90), (80, 160),
             (3, 7), (7, 3), (1, 5), (5, 1), (1080, 2160), (1920, 3840), (511, 1023), (1000, 333)]
    tabs = {}
    for i, o in pairs:
        ramp = np.arange(i, dtype=np.int32).reshape(1, i)
        tabs[f"{i}_{o}"] = np.array(Image.fromarray(ramp, mode="I").resize((o, 1), Image.NEAREST))[0].astype(np.int32)
    np.savez_compressed(os.path.join(HERE, "resample.npz"), **tabs)

    import PIL
    import sklearn

    with open(os.path.join(HERE, "versions.json"), "w") as fh:
        json.dump({"numpy": np.__version__, "scikit-learn": sklearn.__version__, "Pillow": PIL.__version__,
                   "python": sys.version.split()[0]}, fh, indent=1)
    print("golden fixtures written")


if __name__ == "__main__":
    main()
