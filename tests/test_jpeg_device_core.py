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
