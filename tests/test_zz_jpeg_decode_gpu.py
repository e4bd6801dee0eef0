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


def test_pipeline_tail_from_jpeg_files_stays_on_the_device(tmp_path):
    """content.jpg + stylised.jpg (LST:200, 219) -> device decode -> transfer -> device encode: the file equals
    Pillow-decode -> oracle -> Pillow-encode byte for byte; a progressive file goes through Pillow's reader."""
    from applied_image_processing_b200.localized_style_transfer import run_localized_style_transfer
    from oracle import oracle as orc
    from tests.synth import ellipse_mask, synth_image

    H, W = 360, 640
    PIL.fromarray(synth_image(H, W, 3, "N")).save(tmp_path / "content.jpg", quality=90)
    PIL.fromarray(synth_image(256, 455, 4, "N")).save(tmp_path / "sty.jpg")
    PIL.fromarray(synth_image(256, 455, 4, "N")).save(tmp_path / "sty_prog.jpg", progressive=True)
    mask = ellipse_mask(H, W, 0.45, 6)
    for sty_name in ("sty.jpg", "sty_prog.jpg"):
        c = np.array(PIL.open(tmp_path / "content.jpg").convert("RGB"))
        s = np.array(PIL.open(tmp_path / sty_name).convert("RGB"))
        ref, status, _ = orc.transfer(c, s, mask)
        buf = io.BytesIO()
        PIL.fromarray(ref).save(buf, "JPEG")
        path = run_localized_style_transfer(str(tmp_path / "content.jpg"), None, output_path=str(tmp_path),
                                            background_mask=mask, stylized_image=str(tmp_path / sty_name))
        assert status == 0 and open(path, "rb").read() == buf.getvalue(), sty_name


@pytest.mark.parametrize("cfmt,sfmt,H,W,sh,sw", [("JPEG", "JPEG", 360, 640, 256, 455), ("PNG", "JPEG", 270, 480, 270, 480),
                                                 ("JPEG", "PNG", 97, 131, 64, 80), ("GRAY", "JPEG", 128, 192, 128, 192),
                                                 # a grey stylised file of more than 3*H*W pixels: its grey plane once
                                                 # overran the output buffer it borrowed
                                                 ("JPEG", "GRAY", 64, 96, 256, 455), ("GRAY", "GRAY", 40, 56, 300, 300)])
def test_files_entry_point_matches_pillow_plus_oracle(cfmt, sfmt, H, W, sh, sw):
    """rct_transfer_files_host: file bytes in (JPEG on the device, PNG on the host, a grey JPEG replicated to RGB as
    convert("RGB") does), mask, JPEG bytes out == Pillow decode -> oracle -> Pillow encode, byte for byte."""
    import applied_image_processing_b200 as aip
    from oracle import oracle as orc
    from tests.synth import ellipse_mask, synth_image

    def enc(a, fmt):
        buf = io.BytesIO()
        if fmt == "GRAY":
            PIL.fromarray(a[..., 0]).save(buf, "JPEG", quality=85)
        else:
            PIL.fromarray(a).save(buf, fmt, **({"quality": 90} if fmt == "JPEG" else {}))
        return buf.getvalue()

    cf, sf = enc(synth_image(H, W, 3, "N"), cfmt), enc(synth_image(sh, sw, 4, "N"), sfmt)
    mask = ellipse_mask(H, W, 0.45, 6)
    c = np.array(PIL.open(io.BytesIO(cf)).convert("RGB"))
    s = np.array(PIL.open(io.BytesIO(sf)).convert("RGB"))
    ref, status, _ = orc.transfer(c, s, mask)
    for q in (75, 92):
        buf = io.BytesIO()
        PIL.fromarray(ref).save(buf, "JPEG", quality=q)
        got = aip.regional_color_transfer_files(cf, sf, mask, quality=q)
        assert status == 0 and got == buf.getvalue(), (cfmt, sfmt, q)
    assert aip.image_info(cf) == (H, W) and aip.image_info(sf) == (sh, sw)
    with pytest.raises(ValueError, match="mask is"):
        aip.regional_color_transfer_files(cf, sf, mask[:-1])


@pytest.mark.parametrize("h,w,kw", [(270, 480, {}), (540, 960, {"quality": 90}), (1080, 1920, {}), (96, 128, {"subsampling": 0}),
                                    (37, 53, {"optimize": True})])
def test_self_synchronising_decoder_matches_pillow(h, w, kw):
    import applied_image_processing_b200 as aip

    data = _file(h, w, h + w, **kw)
    assert np.array_equal(aip.decode_jpeg(data, parallel=True), _pillow(data))
    assert np.array_equal(aip.decode_jpeg(data, parallel=False), _pillow(data))
    assert np.array_equal(aip.decode_jpeg(data), _pillow(data))
