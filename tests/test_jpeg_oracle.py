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


# ---- decoder half (Image.open of the content / stylised files, LST:200, 219) -----------------------------------
def _pillow_decode(data, mode="RGB"):
    return np.asarray(PIL.open(io.BytesIO(data)).convert(mode))


@pytest.mark.parametrize("h,w", [(16, 16), (1, 1), (2, 2), (3, 2), (2, 3), (5, 1), (1, 5), (17, 23), (33, 65), (9, 40),
                                 (40, 9), (64, 48), (100, 75)])
@pytest.mark.parametrize("subsampling", [2, 1, 0])      # 4:2:0 (Pillow's default), 4:2:2, 4:4:4
def test_decoder_is_bit_exact_against_pillow(h, w, subsampling):
    for kind, q in (("noise", 75), ("smooth", 95), ("noise", 30)):
        a = _image(kind, h, w, h * 31 + w + q)
        data = _pillow(a, quality=q, subsampling=subsampling)
        assert np.array_equal(jo.decode_jpeg(data), _pillow_decode(data)), (kind, q)


def test_decoder_grayscale_restart_markers_and_optimised_tables():
    rng = np.random.default_rng(5)
    g = rng.integers(0, 256, (37, 53), dtype=np.uint8)
    data = _pillow(g)
    assert np.array_equal(jo.decode_jpeg(data), _pillow_decode(data, "L"))
    a = _image("noise", 70, 90, 1)
    for kw in ({"restart_marker_blocks": 3}, {"restart_marker_rows": 1}, {"optimize": True}):
        data = _pillow(a, **kw)
        assert np.array_equal(jo.decode_jpeg(data), _pillow_decode(data)), kw


def test_encode_then_decode_through_the_oracle_equals_pillow_round_trip():
    a = _image("smooth", 45, 70, 2)
    assert np.array_equal(jo.decode_jpeg(jo.encode_jpeg(a)), _pillow_decode(_pillow(a)))


def test_decoder_rejects_progressive_files():
    data = _pillow(_image("noise", 16, 16, 0), progressive=True)
    with pytest.raises(NotImplementedError):
        jo.decode_jpeg(data)
