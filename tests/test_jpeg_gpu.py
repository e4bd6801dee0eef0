"""SURVEY 8f row f-3 on the GPU: rct_jpeg_encode_async (through applied_image_processing_b200.jpeg) writes the same
bytes as Pillow's save -- the reference's own call (localized_style_transfer.py:241-243) -- and as the oracle."""
import io

import numpy as np
import pytest
import torch

PIL = pytest.importorskip("PIL.Image")
pytestmark = pytest.mark.gpu


def _pillow(a, q=75):
    buf = io.BytesIO()
    PIL.fromarray(a).save(buf, "JPEG", quality=q)
    return buf.getvalue()


def _image(kind, h, w, seed):
    rng = np.random.default_rng(seed)
    if kind == "noise":
        return rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
    yy, xx = np.mgrid[0:h, 0:w]
    base = np.stack([(xx * 5 + yy) % 256, (yy * 3) % 256, (xx + yy * 2) % 256], -1)
    return np.clip(base + rng.integers(-12, 12, (h, w, 3)), 0, 255).astype(np.uint8)


@pytest.mark.parametrize("h,w", [(16, 16), (1, 1), (8, 8), (17, 23), (33, 65), (9, 40), (40, 9), (100, 75), (270, 480)])
@pytest.mark.parametrize("kind", ["noise", "smooth"])
def test_encoder_matches_pillow_and_oracle(h, w, kind):
    import applied_image_processing_b200 as aip
    from oracle import jpeg_oracle as jo

    a = _image(kind, h, w, 100 * h + w)
    got = aip.encode_jpeg(a)
    assert got == _pillow(a)
    if h * w <= 100 * 75:
        assert got == jo.encode_jpeg(a)


@pytest.mark.parametrize("quality", [1, 30, 75, 95, 100])
def test_quality_levels_and_capacity_retry(quality):
    # noise at quality 100 needs about 2 bytes per pixel: the first-try buffer overflows and the call re-encodes
    import applied_image_processing_b200 as aip

    a = _image("noise", 360, 640, quality)
    assert aip.encode_jpeg(a, quality) == _pillow(a, quality)


def test_batch_of_frames_cuda_tensor_input_and_save(tmp_path):
    import applied_image_processing_b200 as aip

    frames = np.stack([_image("smooth", 1080, 1920, k) for k in range(3)])
    files = aip.encode_jpeg(torch.from_numpy(frames).cuda())
    assert [len(f) for f in files] == [len(_pillow(f)) for f in frames]
    for f, a in zip(files, frames):
        assert f == _pillow(a)
    aip.save_jpeg(frames[1], tmp_path / "r.jpg")
    assert (tmp_path / "r.jpg").read_bytes() == _pillow(frames[1])
    back = np.asarray(PIL.open(tmp_path / "r.jpg").convert("RGB"))
    assert back.shape == frames[1].shape and np.abs(back.astype(int) - frames[1]).mean() < 12


def test_transfer_then_encode_equals_the_reference_pipeline_tail():
    # LST:232-243: combined_image -> Image.fromarray(...).save(jpg), with both steps on the device
    import applied_image_processing_b200 as aip
    from tests.synth import ellipse_mask, synth_image

    H, W = 270, 480
    content, sty, mask = synth_image(H, W, 0, "N"), synth_image(200, 300, 1, "N"), ellipse_mask(H, W, 0.4, 4)
    combined = aip.regional_color_transfer(content, sty, mask)
    assert aip.encode_jpeg(combined) == _pillow(combined)


def test_bad_arguments_raise():
    import applied_image_processing_b200 as aip

    with pytest.raises(ValueError):
        aip.encode_jpeg(np.zeros((8, 8, 3), np.uint8), quality=0)
    with pytest.raises(ValueError):
        aip.encode_jpeg(np.zeros((8, 8), np.uint8))
    with pytest.raises(TypeError):
        aip.encode_jpeg(np.zeros((8, 8, 3), np.float32))
    assert aip.encode_jpeg(np.zeros((0, 8, 8, 3), np.uint8)) == []
