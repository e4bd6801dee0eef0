"""SURVEY 8f row f-3, decode half, on the GPU through the C ABI: rct_jpeg_decode_async gives the pixels Pillow decodes
(`Image.open`, localized_style_transfer.py:200, 219), bit for bit.  (Sorted last: the newest entry point.)"""
import io

import numpy as np
import pytest

PIL = pytest.importorskip("PIL.Image")
pytestmark = pytest.mark.gpu


def _file(h, w, seed, mode="RGB", **kw):
    rng = np.random.default_rng(seed)
    yy, xx = np.mgrid[0:h, 0:w]
    a = np.clip(np.stack([(xx * 5 + yy) % 256, (yy * 3) % 256, (xx + yy * 2) % 256], -1) + rng.integers(-20, 20, (h, w, 3)),
                0, 255).astype(np.uint8)
    buf = io.BytesIO()
    PIL.fromarray(a).convert(mode).save(buf, "JPEG", **kw)
    return buf.getvalue()


def _pillow(data, mode="RGB"):
    return np.array(PIL.open(io.BytesIO(data)).convert(mode))


@pytest.mark.parametrize("h,w", [(16, 16), (1, 1), (2, 2), (2, 3), (5, 1), (17, 23), (33, 65), (40, 9), (100, 75)])
@pytest.mark.parametrize("subsampling", [2, 1, 0])
def test_decoder_matches_pillow(h, w, subsampling):
    import applied_image_processing_b200 as aip

    data = _file(h, w, 7 * h + w, subsampling=subsampling, quality=85)
    assert aip.jpeg_info(data) == (h, w, 3)
    assert np.array_equal(aip.decode_jpeg(data), _pillow(data))


def test_restart_markers_gray_files_paths_and_device_output(tmp_path):
    import torch

    import applied_image_processing_b200 as aip

    for kw in ({"restart_marker_rows": 1}, {"restart_marker_blocks": 5}, {"optimize": True}):
        data = _file(270, 480, 3, **kw)
        assert np.array_equal(aip.decode_jpeg(data), _pillow(data)), kw
    gray = _file(37, 53, 4, mode="L")
    assert aip.jpeg_info(gray) == (37, 53, 1)
    assert np.array_equal(aip.decode_jpeg(gray), _pillow(gray, "L"))
    path = tmp_path / "content.jpg"
    path.write_bytes(_file(64, 48, 5))
    t = aip.decode_jpeg(path, device=True)
    assert isinstance(t, torch.Tensor) and t.is_cuda and np.array_equal(t.cpu().numpy(), _pillow(path.read_bytes()))


def test_encode_decode_round_trip_on_the_device_equals_pillows():
    import applied_image_processing_b200 as aip

    a = _pillow(_file(96, 128, 6))
    ours = aip.decode_jpeg(aip.encode_jpeg(a))
    buf = io.BytesIO()
    PIL.fromarray(a).save(buf, "JPEG")
    assert np.array_equal(ours, _pillow(buf.getvalue()))


def test_unsupported_files_raise():
    import applied_image_processing_b200 as aip

    with pytest.raises(ValueError, match="baseline"):
        aip.decode_jpeg(_file(16, 16, 0, progressive=True))
    with pytest.raises(ValueError):
        aip.decode_jpeg(b"not a jpeg at all")


def test_pipeline_tail_writes_the_file_the_reference_would(tmp_path):
    # LST:191-245 with the mask and the stylised image injected: open -> transfer -> combine -> save JPEG
    import applied_image_processing_b200 as aip
    from applied_image_processing_b200.localized_style_transfer import run_localized_style_transfer
    from oracle import oracle as orc
    from tests.synth import ellipse_mask, synth_image

    H, W = 270, 480
    content, sty, mask = synth_image(H, W, 0, "N"), synth_image(200, 300, 1, "N"), ellipse_mask(H, W, 0.4, 4)
    PIL.fromarray(content).save(tmp_path / "content.png")
    path = run_localized_style_transfer(str(tmp_path / "content.png"), None, output_path=str(tmp_path),
                                        background_mask=mask[None], stylized_image=sty)
    assert path.endswith("localized_style_transfer_result.jpg")
    ref, status, _ = orc.transfer(content, sty, mask)
    assert status == 0
    buf = io.BytesIO()
    PIL.fromarray(ref).save(buf, "JPEG")
    assert open(path, "rb").read() == buf.getvalue()
    assert np.array_equal(aip.decode_jpeg(path), _pillow(buf.getvalue()))


@pytest.mark.parametrize("h,w,kw", [(270, 480, {}), (540, 960, {"quality": 90}), (1080, 1920, {}), (96, 128, {"subsampling": 0}),
                                    (37, 53, {"optimize": True})])
def test_self_synchronising_decoder_matches_pillow(h, w, kw):
    import applied_image_processing_b200 as aip

    data = _file(h, w, h + w, **kw)
    assert np.array_equal(aip.decode_jpeg(data, parallel=True), _pillow(data))
    assert np.array_equal(aip.decode_jpeg(data, parallel=False), _pillow(data))
    assert np.array_equal(aip.decode_jpeg(data), _pillow(data))
