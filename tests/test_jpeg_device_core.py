"""The __host__ __device__ bodies of csrc/rct_jpeg.cuh (sample preparation, islow DCT, quantiser, Huffman coder, bit
sink, header) run serially on the CPU by tools/jpeg_selftest --host and compared with Pillow and the oracle: pins the
kernels' arithmetic without a GPU.  The kernels themselves are checked by tests/test_jpeg_gpu.py."""
import io
import os
import subprocess

import numpy as np
import pytest

PIL = pytest.importorskip("PIL.Image")
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "tools", "jpeg_selftest")


@pytest.fixture(scope="module")
def selftest():
    if not os.path.exists(EXE):
        import __graft_entry__

        __graft_entry__.build()
    return EXE


def _pillow(a, q):
    buf = io.BytesIO()
    PIL.fromarray(a).save(buf, "JPEG", quality=q)
    return buf.getvalue()


@pytest.mark.parametrize("h,w,q", [(16, 16, 75), (1, 1, 75), (17, 23, 75), (33, 65, 50), (9, 40, 100), (40, 9, 5),
                                   (100, 75, 90), (270, 480, 75), (31, 16, 75), (2, 3, 75)])
def test_device_core_on_host_is_byte_exact(selftest, tmp_path, h, w, q):
    from oracle import jpeg_oracle as jo

    rng = np.random.default_rng(h * 977 + w)
    a = rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
    if (h + w) & 1:  # smooth content on half of the cases: long zero runs, ZRL codes, small DC differences
        yy, xx = np.mgrid[0:h, 0:w]
        a = np.stack([(xx * 5 + yy) % 256, (yy * 3) % 256, (xx + yy * 2) % 256], -1).astype(np.uint8)
    a.tofile(tmp_path / "in.rgb")
    subprocess.run([selftest, "--host", str(h), str(w), str(tmp_path / "in.rgb"), str(tmp_path / "o.jpg"), str(q)], check=True)
    got = (tmp_path / "o.jpg").read_bytes()
    assert got == _pillow(a, q)
    if h * w <= 100 * 75:
        assert got == jo.encode_jpeg(a, q)


def test_cabi_exports_the_encoder_and_sizes_are_consistent():
    from applied_image_processing_b200 import _cabi

    lib = _cabi.lib()
    assert lib.rct_jpeg_workspace_bytes(0, 8, 1) == 0 and lib.rct_jpeg_max_bytes(70000, 8) == 0
    one, two = lib.rct_jpeg_workspace_bytes(2160, 3840, 1), lib.rct_jpeg_workspace_bytes(2160, 3840, 2)
    assert 0 < one < two <= 2 * one
    assert lib.rct_jpeg_max_bytes(16, 16) >= 640 + 2 * 6 * 208


# ---- decoder half ------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("h,w,kw", [(16, 16, {}), (1, 1, {}), (2, 2, {}), (3, 2, {}), (5, 1, {}), (17, 23, {"quality": 95}),
                                    (33, 65, {"subsampling": 1}), (40, 9, {"subsampling": 0}), (70, 90, {"restart_marker_blocks": 3}),
                                    (100, 75, {"restart_marker_rows": 1}), (64, 48, {"optimize": True}), (270, 480, {"quality": 30})])
def test_decoder_core_on_host_is_bit_exact(selftest, tmp_path, h, w, kw):
    rng = np.random.default_rng(h * 131 + w)
    a = rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
    if (h + w) & 1:
        yy, xx = np.mgrid[0:h, 0:w]
        a = np.stack([(xx * 5 + yy) % 256, (yy * 3) % 256, (xx + yy * 2) % 256], -1).astype(np.uint8)
    buf = io.BytesIO()
    PIL.fromarray(a).save(buf, "JPEG", **kw)
    (tmp_path / "i.jpg").write_bytes(buf.getvalue())
    p = subprocess.run([selftest, "--host-decode", str(tmp_path / "i.jpg"), str(tmp_path / "o.raw")], check=True, capture_output=True, text=True)
    assert p.stdout.split() == [str(h), str(w), "3"]
    got = np.fromfile(tmp_path / "o.raw", np.uint8).reshape(h, w, 3)
    assert np.array_equal(got, np.asarray(PIL.open(io.BytesIO(buf.getvalue())).convert("RGB")))


def test_decode_info_parses_headers_on_the_host_and_rejects_what_it_cannot_decode():
    import applied_image_processing_b200 as aip

    buf = io.BytesIO()
    PIL.fromarray(np.zeros((37, 53, 3), np.uint8)).save(buf, "JPEG")
    assert aip.jpeg_info(buf.getvalue()) == (37, 53, 3)
    buf = io.BytesIO()
    PIL.fromarray(np.zeros((9, 5), np.uint8)).save(buf, "JPEG")
    assert aip.jpeg_info(buf.getvalue()) == (9, 5, 1)
    buf = io.BytesIO()
    PIL.fromarray(np.zeros((16, 16, 3), np.uint8)).save(buf, "JPEG", progressive=True)
    with pytest.raises(ValueError, match="baseline"):
        aip.jpeg_info(buf.getvalue())
    with pytest.raises(ValueError):
        aip.jpeg_info(b"\xff\xd8\xff")


@pytest.mark.skipif(__import__("torch").cuda.is_available(), reason="checks the behaviour without a GPU")
def test_jpeg_entry_points_fail_loudly_without_a_gpu():
    import applied_image_processing_b200 as aip

    buf = io.BytesIO()
    PIL.fromarray(np.zeros((16, 16, 3), np.uint8)).save(buf, "JPEG")
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        aip.encode_jpeg(np.zeros((8, 8, 3), np.uint8))
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        aip.decode_jpeg(buf.getvalue())


@pytest.mark.parametrize("h,w,kw,noise", [(270, 480, {}, False), (270, 480, {"quality": 95}, True), (100, 75, {"optimize": True}, True),
                                          (96, 128, {"subsampling": 0}, False), (64, 200, {"subsampling": 1}, True), (37, 53, {}, None)])
def test_self_synchronising_decoder_core_on_host_is_bit_exact(selftest, tmp_path, h, w, kw, noise):
    # the parallel Huffman decoder's bodies (speculative pass, rounds to a fixed point, count, write, DC prefix) run serially
    rng = np.random.default_rng(h + w)
    yy, xx = np.mgrid[0:h, 0:w]
    a = np.stack([(xx * 5 + yy) % 256, (yy * 3) % 256, (xx + yy * 2) % 256], -1) + rng.integers(-12, 12, (h, w, 3))
    if noise:
        a = rng.integers(0, 256, (h, w, 3))
    img = PIL.fromarray(np.clip(a, 0, 255).astype(np.uint8))
    if noise is None:
        img = img.convert("L")
    buf = io.BytesIO()
    img.save(buf, "JPEG", **kw)
    (tmp_path / "i.jpg").write_bytes(buf.getvalue())
    p = subprocess.run([selftest, "--host-decode", str(tmp_path / "i.jpg"), str(tmp_path / "o.raw"), "par"], check=True,
                       capture_output=True, text=True)
    assert "converged" in p.stderr and "NOT" not in p.stderr
    ref = np.asarray(PIL.open(io.BytesIO(buf.getvalue())))
    got = np.fromfile(tmp_path / "o.raw", np.uint8).reshape(ref.shape)
    assert np.array_equal(got, ref)
