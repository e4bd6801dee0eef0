"""f-3, PNG decode (host side, csrc/rct_png.h) against Pillow's `Image.open(...).convert("RGB")` (LST:200, 219)."""
import io

import numpy as np
import pytest
from PIL import Image

import applied_image_processing_b200 as aip


def _png(img, **kw):
    buf = io.BytesIO()
    img.save(buf, "PNG", **kw)
    return buf.getvalue()


def _want(data):
    return np.array(Image.open(io.BytesIO(data)).convert("RGB"))


def _noise(h, w, c, seed):
    rng = np.random.default_rng(seed)
    yy, xx = np.mgrid[0:h, 0:w]
    base = ((xx * 3 + yy * 2)[..., None] + np.arange(c) * 40) % 256
    return np.clip(base + rng.integers(-20, 20, (h, w, c)), 0, 255).astype(np.uint8)


@pytest.mark.parametrize("mode,h,w", [("RGB", 37, 53), ("RGB", 300, 300), ("RGBA", 64, 48), ("L", 33, 70), ("LA", 20, 31),
                                      ("P", 45, 67), ("1", 30, 41), ("RGB", 1, 1), ("RGB", 1, 257), ("L", 255, 1)])
def test_png_matches_pillow(mode, h, w):
    if mode == "P":
        img = Image.fromarray(_noise(h, w, 3, 1)).quantize(37)
    elif mode == "1":
        img = Image.fromarray(_noise(h, w, 1, 2)[..., 0]).convert("1")
    else:
        c = {"RGB": 3, "RGBA": 4, "L": 1, "LA": 2}[mode]
        a = _noise(h, w, c, 3)
        img = Image.fromarray(a[..., 0] if c == 1 else a, mode)
    for kw in ({}, {"optimize": True}, {"compress_level": 0}):
        data = _png(img, **kw)
        got = aip.decode_png(data)
        assert got.dtype == np.uint8 and np.array_equal(got, _want(data)), (mode, kw)


def test_png_low_bit_depths_and_palette_bits():
    rng = np.random.default_rng(5)
    for bits in (1, 2, 4):
        img = Image.fromarray(rng.integers(0, 1 << bits, (23, 37), dtype=np.uint8), "P")
        img.putpalette([int(v) for v in rng.integers(0, 256, 3 * (1 << bits))])
        data = _png(img, bits=bits)
        assert np.array_equal(aip.decode_png(data), _want(data)), bits


def test_png_refuses_what_it_does_not_decode():
    a = _noise(40, 40, 3, 7)
    with pytest.raises(ValueError, match="interlaced"):
        # Pillow cannot write Adam7: flip the interlace byte of the IHDR (the CRC is not checked before the refusal)
        d = bytearray(_png(Image.fromarray(a)))
        d[28] = 1
        aip.decode_png(bytes(d))
    with pytest.raises(ValueError, match="16-bit"):
        aip.decode_png(_png(Image.fromarray((a[..., 0].astype(np.uint16) << 8), "I;16")))
    with pytest.raises(ValueError, match="not a PNG"):
        aip.decode_png(b"\xff\xd8\xff\xe0" + b"\0" * 64)
    with pytest.raises(ValueError):
        aip.decode_png(_png(Image.fromarray(a))[:200])


def test_reference_png_inputs():
    import os

    p = "/root/reference/input/style/sketch.png"
    if not os.path.exists(p):
        pytest.skip("reference inputs are not on this machine")
    data = open(p, "rb").read()
    assert np.array_equal(aip.decode_png(data), _want(data))
    assert np.array_equal(aip.decode_image(data), _want(data))
