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


def _chunks(f):
    import struct

    pos, out = 8, []
    while pos < len(f):
        n = struct.unpack(">I", f[pos:pos + 4])[0]
        out.append((f[pos + 4:pos + 8], f[pos + 8:pos + 8 + n], f[pos + 8 + n:pos + 12 + n]))
        pos += 12 + n
    return out


def _build(chunks, fix_crc):
    import struct
    import zlib

    o = b"\x89PNG\r\n\x1a\n"
    for t, d, c in chunks:
        o += struct.pack(">I", len(d)) + t + d + (struct.pack(">I", zlib.crc32(t + d)) if fix_crc else c)
    return o


def _damaged_files():
    """name -> file bytes: the ways a PNG can be damaged that Pillow has an opinion on."""
    import zlib

    a = _noise(30, 40, 3, 11)
    ch = _chunks(_png(Image.fromarray(a)))
    raw = zlib.decompress(b"".join(d for t, d, _ in ch if t == b"IDAT"))
    zero = b"\0\0\0\0"
    text = (b"tEXt", b"k\0v", zero)

    def idat(z):
        return [(t, (z if t == b"IDAT" else d), c) for t, d, c in ch]

    z = b"".join(d for t, d, _ in ch if t == b"IDAT")
    return {
        "bad IDAT crc": _build([(t, d, zero if t == b"IDAT" else c) for t, d, c in ch], False),
        "bad IHDR crc": _build([(t, d, zero if t == b"IHDR" else c) for t, d, c in ch], False),
        "bad IEND crc": _build([(t, d, zero if t == b"IEND" else c) for t, d, c in ch], False),
        "bad tEXt crc before IDAT": _build([ch[0], text] + ch[1:], False),
        "bad tEXt crc after IDAT": _build(ch[:-1] + [text] + ch[-1:], False),
        "adler cut off": _build(idat(z[:-4]), True),
        "wrong adler": _build(idat(z[:-4] + b"\1\2\3\4"), True),
        "rows after the last": _build(idat(zlib.compress(raw + b"\0" * 50)), True),
        "stream ends early": _build(idat(zlib.compress(raw[:-10])), True),
        "no IEND": _build(ch[:-1], True),
    }


def test_damaged_png_files_get_pillows_verdict():
    """csrc/rct_png.h accepts exactly the damaged files Pillow's Image.open().convert("RGB") accepts (same pixels) and
    refuses the ones Pillow raises on: checksums are enforced in front of the first IDAT only, a cut-off trailer is
    tolerated once every row is there, a wrong Adler-32 or a short stream is not."""
    verdicts = {}
    for name, data in _damaged_files().items():
        try:
            want = _want(data)
        except Exception:
            want = None
        verdicts[name] = want is not None
        if want is None:
            with pytest.raises(ValueError):
                aip.decode_png(data)
        else:
            assert np.array_equal(aip.decode_png(data), want), name
    # the probe itself must have seen both kinds, or the test proves nothing
    assert verdicts["bad IHDR crc"] is False and verdicts["wrong adler"] is False and verdicts["stream ends early"] is False
    assert verdicts["bad IDAT crc"] is True and verdicts["adler cut off"] is True


def test_png_inflate_in_steps(tmp_path):
    """zlib's avail_in / avail_out are 32-bit, a 65535 x 65535 PNG is not: rct_png.h feeds zlib in steps of
    RCT_PNG_ZCHUNK bytes.  Built here with a step of 7 bytes so that a small file crosses many step boundaries."""
    import os
    import shutil
    import subprocess

    if shutil.which("g++") is None:
        pytest.skip("no g++")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    exe = tmp_path / "png_chunk"
    subprocess.check_call(["g++", "-O1", "-DRCT_PNG_ZCHUNK=7u", "-I", os.path.join(root, "applied_image_processing_b200", "csrc"),
                           "-o", str(exe), os.path.join(root, "tests", "helpers", "png_chunk_main.cpp"), "-lz"])
    for mode, h, w in (("RGB", 61, 47), ("L", 33, 70), ("RGBA", 20, 31)):
        c = {"RGB": 3, "L": 1, "RGBA": 4}[mode]
        a = _noise(h, w, c, 13)
        data = _png(Image.fromarray(a[..., 0] if c == 1 else a, mode))
        (tmp_path / "t.png").write_bytes(data)
        out = subprocess.run([str(exe), str(tmp_path / "t.png")], capture_output=True, check=True).stdout
        assert np.array_equal(np.frombuffer(out, np.uint8).reshape(h, w, 3), _want(data)), mode
    for name in ("wrong adler", "stream ends early", "bad IHDR crc"):
        (tmp_path / "d.png").write_bytes(_damaged_files()[name])
        r = subprocess.run([str(exe), str(tmp_path / "d.png")], capture_output=True)
        assert r.returncode == 1 and r.stdout.startswith(b"ERR"), name
