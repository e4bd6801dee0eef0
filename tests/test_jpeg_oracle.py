"""SURVEY 8f row f-3 groundwork: the numpy baseline-JPEG encoder of oracle/jpeg_oracle.py writes the same
bytes as Pillow's `Image.fromarray(a).save(f, "JPEG")` -- the call at the end of the reference's pipeline
(localized_style_transfer.py:241-243).  Pillow is the reference's own dependency; it runs here on CPU."""
import io

import numpy as np
import pytest

PIL = pytest.importorskip("PIL.Image")
from oracle import jpeg_oracle as jo


def _pillow(a, **kw):
    buf = io.BytesIO()
    PIL.fromarray(a).save(buf, "JPEG", **kw)
    return buf.getvalue()


def _image(kind, h, w, seed):
    rng = np.random.default_rng(seed)
    if kind == "noise":
        return rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
    if kind == "flat":
        return np.broadcast_to(rng.integers(0, 256, 3, dtype=np.uint8), (h, w, 3)).copy()
    yy, xx = np.mgrid[0:h, 0:w]
    return np.stack([(xx * 5 + yy) % 256, (yy * 3) % 256, (xx + yy * 2) % 256], -1).astype(np.uint8)


# whole MCUs, ragged right / bottom edges (dummy luma blocks, replicated chroma rows), single pixel
SIZES = [(16, 16), (8, 8), (1, 1), (64, 48), (17, 23), (100, 75), (33, 65), (9, 40), (40, 9), (15, 17), (31, 16), (2, 3)]


@pytest.mark.parametrize("h,w", SIZES)
@pytest.mark.parametrize("kind", ["noise", "smooth", "flat"])
def test_encoder_is_byte_exact_against_pillow(h, w, kind):
    a = _image(kind, h, w, h * 1000 + w)
    assert jo.encode_jpeg(a) == _pillow(a)


@pytest.mark.parametrize("quality", [1, 10, 49, 50, 75, 90, 95, 100])
def test_quality_scaling(quality):
    a = _image("noise", 24, 40, quality)
    assert jo.encode_jpeg(a, quality) == _pillow(a, quality=quality)


def test_extreme_coefficients_and_stuffed_bytes():
    # checkerboards at full swing give the largest AC magnitudes; the noise image is long enough to hold 0xFF bytes
    yy, xx = np.mgrid[0:32, 0:32]
    a = np.where(((xx + yy) & 1)[..., None] == 1, 255, 0).astype(np.uint8).repeat(3, axis=2)
    assert jo.encode_jpeg(a, 100) == _pillow(a, quality=100)
    b = _image("noise", 96, 128, 7)
    enc = jo.encode_jpeg(b, 100)
    assert b"\xff\x00" in enc and enc == _pillow(b, quality=100)


def test_reference_geometry_frame():
    # one 480 x 270 frame (the reference's nerfstudio renders are 16:9), smooth content plus noise
    rng = np.random.default_rng(3)
    a = np.clip(_image("smooth", 270, 480, 0).astype(int) + rng.integers(-20, 20, (270, 480, 3)), 0, 255).astype(np.uint8)
    assert jo.encode_jpeg(a) == _pillow(a)
