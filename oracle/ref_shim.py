"""TEST INFRASTRUCTURE ONLY -- import shim for the *unmodified* reference module.

Loads ``/root/reference/Style_3DGS/localized_style_transfer.py`` (abbrev. LST) in-process
so that the C oracle (``oracle/rct_oracle.c``) and the golden fixtures under
``tests/golden/`` can be pinned against the reference's own code.  Nothing is copied: the
reference file is read where it lies.  ``/root/reference`` does not exist on the GPU box,
so every caller must go through :func:`available` first.

Why a shim: ``import Style_3DGS.localized_style_transfer`` fails
because the package ``__init__`` pulls in the 3DGS trainer (needs the un-vendored
``diff_gaussian_rasterization``), LST:2 imports matplotlib (unused) and LST:5 imports
``Style_3DGS.AdaIN``.  We pre-register stub modules for exactly those three names.

Only ``tests/``, ``tests/golden/make_golden.py``, ``bench.py --impl reference`` (in this
container) may import this file.  The product package never does.
"""
from __future__ import annotations

import importlib
import os
import sys
import types

REFERENCE_ROOT = os.environ.get("RCT_REFERENCE_ROOT", "/root/reference")
_LST = None


def available() -> bool:
    return os.path.isfile(
        os.path.join(REFERENCE_ROOT, "Style_3DGS", "localized_style_transfer.py")
    )


def load_reference():
    """Return the reference LST module (unmodified), importing it on first use."""
    global _LST
    if _LST is not None:
        return _LST
    if not available():
        raise RuntimeError(f"reference not present under {REFERENCE_ROOT}")
    if "matplotlib" not in sys.modules:
        try:
            import matplotlib.pyplot  # noqa: F401
        except Exception:
            mpl = types.ModuleType("matplotlib")
            mpl.__path__ = []  # mark as package
            plt = types.ModuleType("matplotlib.pyplot")
            mpl.pyplot = plt
            sys.modules["matplotlib"] = mpl
            sys.modules["matplotlib.pyplot"] = plt
    pkg = types.ModuleType("Style_3DGS")
    pkg.__path__ = [os.path.join(REFERENCE_ROOT, "Style_3DGS")]
    sys.modules.setdefault("Style_3DGS", pkg)
    adain = types.ModuleType("Style_3DGS.AdaIN")

    def adain_inference(*a, **k):  # pragma: no cover - never called on the hot path
        raise RuntimeError("AdaIN is out of scope (weights absent); see SURVEY.md 8c")

    adain.adain_inference = adain_inference
    sys.modules.setdefault("Style_3DGS.AdaIN", adain)
    _LST = importlib.import_module("Style_3DGS.localized_style_transfer")
    return _LST


def ref_pipeline(content_np, stylized_np, background_mask):
    """LST:222-238 driven verbatim: resize-if-needed, mask algebra, transfer, combine.

    ``background_mask`` is the uint8 ``[H,W]`` array the reference holds after LST:218.
    Returns the ``combined_image`` uint8 ``[H,W,3]`` of LST:238.
    """
    import numpy as np
    from PIL import Image

    lst = load_reference()
    if (stylized_np.shape[0] != background_mask.shape[0]) or (
        stylized_np.shape[1] != background_mask.shape[1]
    ):
        stylized_np = np.array(
            Image.fromarray(stylized_np).resize(
                (background_mask.shape[1], background_mask.shape[0]), Image.NEAREST
            )
        )
    foreground_mask = 1 - background_mask
    foreground_np = content_np * foreground_mask[..., None]
    background_np = stylized_np * background_mask[..., None]
    adjusted_foreground = lst.color_transfer_foreground(foreground_np, background_np)
    combined_image = adjusted_foreground * foreground_mask[..., None] + background_np
    return combined_image.astype(np.uint8)
