"""SURVEY.md 8f rows f-1 (AdaIN mask-blend tail) and f-2 (DeepLab post-processing).

The reference computes both with torch (F.interpolate, torch.softmax), which is present on this
machine and on the GPU box, so torch's own CPU result is the pin: the numpy oracle is checked
against it without a GPU, the CUDA kernels against both on the GPU."""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

from oracle import adjacent_oracle as aorc

TOL_F32 = 2e-5  # float32: the source coordinate carries ~1e-5 of rounding at x ~ 100-700, times a tap difference <= 1.3; 1 LSB of uint8 is 4e-3

SIZES = [  # (H, W, oh, ow, mh, mw)
    (64, 96, 64, 96, 64, 96),      # nothing to resize
    (97, 131, 64, 86, 97, 131),    # AdaIN output smaller than the content (512-short-side rule)
    (120, 90, 200, 150, 60, 45),   # down-sampling the output, up-sampling the mask
    (485, 728, 512, 768, 485, 728),
    (33, 57, 7, 5, 3, 2),
]


def _inputs(H, W, oh, ow, mh, mw, n=2, seed=0):
    rng = np.random.default_rng(seed)
    content = rng.random((n, 3, H, W), dtype=np.float32)
    output = rng.random((n, 3, oh, ow), dtype=np.float32) * 1.2 - 0.1  # decoder output leaves [0,1] a little
    mask = (rng.random((mh, mw)) < 0.5).astype(np.float32)
    return content, output, mask


def _torch_reference(content, output, mask):
    """AdaIN/test.py:222-236 verbatim, on the CPU."""
    c = torch.from_numpy(content)
    o = torch.from_numpy(output)
    m = torch.from_numpy(mask)[None, None]
    m = F.interpolate(m, size=c.shape[-2:], mode="nearest")
    o = F.interpolate(o, size=c.shape[-2:], mode="bilinear", align_corners=False)
    return (c * (1.0 - m) + o * m).numpy()


@pytest.mark.parametrize("geom", SIZES)
def test_oracle_blend_matches_torch_cpu(geom):
    content, output, mask = _inputs(*geom)
    ref = _torch_reference(content, output, mask)
    sty, rgb8 = aorc.adain_mask_blend(content, output, mask)
    assert np.abs(sty - ref).max() <= TOL_F32
    # nearest rule is an index rule: pixels where the mask is 0 must be the content, bit for bit
    m = torch.from_numpy(mask)[None, None]
    m = F.interpolate(m, size=content.shape[-2:], mode="nearest").numpy()[0, 0]
    assert np.array_equal(sty[:, :, m == 0], content[:, :, m == 0])
    want8 = np.clip(ref * 255 + 0.5, 0, 255).astype(np.uint8).transpose(0, 2, 3, 1)
    assert np.abs(rgb8.astype(int) - want8.astype(int)).max() <= 1


def test_oracle_nearest_index_rule_matches_torch():
    for in_size, out_size in [(3, 7), (7, 3), (512, 485), (485, 512), (600, 800), (1, 9), (1080, 2160), (2160, 1080)]:
        src = torch.arange(in_size, dtype=torch.float32)[None, None, None]
        got = F.interpolate(src, size=(1, out_size), mode="nearest")[0, 0, 0].numpy().astype(np.int64)
        assert np.array_equal(got, aorc.nearest_idx(out_size, in_size)), (in_size, out_size)


def test_oracle_softmax_mask_and_argmax_match_torch():
    rng = np.random.default_rng(3)
    logits = (rng.normal(0, 3, (21, 70, 90))).astype(np.float32)
    logits[0] += 1.0
    probs = torch.softmax(torch.from_numpy(logits), dim=0)
    want = (probs[0] > 0.5).float().unsqueeze(0).numpy().astype(np.uint8)  # LST:183-186
    got, p = aorc.softmax_mask(logits, 0.5)
    decidable = np.abs(p - 0.5) > 1e-6
    assert got.shape == want.shape == (1, 70, 90) and got.dtype == np.uint8
    assert np.array_equal(got[0][decidable], want[0][decidable]) and decidable.mean() > 0.999
    assert np.array_equal(aorc.argmax_labels(logits), torch.from_numpy(logits).argmax(0).numpy().astype(np.uint8))


# ---- CUDA ----------------------------------------------------------------------------------------
@pytest.mark.gpu
@pytest.mark.parametrize("geom", SIZES)
def test_cuda_blend_matches_oracle_and_torch(geom):
    from applied_image_processing_b200 import adjacent

    content, output, mask = _inputs(*geom, seed=1)
    sty_o, rgb8_o = aorc.adain_mask_blend(content, output, mask)
    sty, rgb8 = adjacent.adain_mask_blend(torch.from_numpy(content).cuda(), torch.from_numpy(output).cuda(),
                                          mask[None], return_uint8=True)
    sty, rgb8 = sty.cpu().numpy(), rgb8.cpu().numpy()
    assert np.abs(sty - sty_o).max() <= TOL_F32
    assert np.abs(sty - _torch_reference(content, output, mask)).max() <= TOL_F32
    assert np.abs(rgb8.astype(int) - rgb8_o.astype(int)).max() <= 1 and (rgb8 != rgb8_o).mean() < 1e-3
    # same thing through torch on the GPU (what the reference runs when CUDA is available)
    c, o = torch.from_numpy(content).cuda(), torch.from_numpy(output).cuda()
    m = F.interpolate(torch.from_numpy(mask).cuda()[None, None], size=c.shape[-2:], mode="nearest")
    ref_gpu = c * (1.0 - m) + F.interpolate(o, size=c.shape[-2:], mode="bilinear", align_corners=False) * m
    assert (ref_gpu.cpu().numpy() - sty).__abs__().max() <= TOL_F32


@pytest.mark.gpu
def test_cuda_deeplab_postprocessing():
    from applied_image_processing_b200 import adjacent

    rng = np.random.default_rng(5)
    logits = rng.normal(0, 3, (21, 213, 317)).astype(np.float32)
    logits[0] += 1.0
    want, p = aorc.softmax_mask(logits, 0.5)
    got = adjacent.deeplab_background_mask(torch.from_numpy(logits).cuda(), 0.5)
    assert got.shape == (1, 213, 317) and got.dtype == np.uint8
    decidable = np.abs(p - 0.5) > 1e-9
    assert np.array_equal(got[0][decidable], want[0][decidable])
    probs = torch.softmax(torch.from_numpy(logits), dim=0)
    ref = (probs[0] > 0.5).numpy().astype(np.uint8)
    band = np.abs(p - 0.5) > 1e-6
    assert np.array_equal(got[0][band], ref[band])
    labels = adjacent.deeplab_label_map(torch.from_numpy(logits).cuda(), as_numpy=True)
    assert np.array_equal(labels, aorc.argmax_labels(logits))


@pytest.mark.gpu
@pytest.mark.parametrize("thr", [0.5, 0.3, 0.9])
def test_cuda_softmax_mask_fast_path_and_guard_band(thr):
    """Vectorised float32 fast path (npix % 4 == 0) with the float64 re-evaluation inside the guard band:
    a third of the pixels are constructed to sit within ~1e-6 of the threshold."""
    from applied_image_processing_b200 import adjacent

    rng = np.random.default_rng(9)
    H, W = 212, 316
    logits = rng.normal(0, 3, (21, H, W)).astype(np.float32)
    # rows 0..69: 20 equal competitors a, class 0 at a + ln(20 thr / (1 - thr)) + tiny -> p0 = thr +- ~1e-7
    a = rng.normal(0, 2, (70, W)).astype(np.float32)
    logits[1:, :70] = a
    logits[0, :70] = a + np.float32(np.log(20 * thr / (1 - thr))) + rng.normal(0, 2e-6, (70, W)).astype(np.float32)
    want, p = aorc.softmax_mask(logits, thr)
    got = adjacent.deeplab_background_mask(torch.from_numpy(logits).cuda(), thr)
    decidable = np.abs(p - np.float64(np.float32(thr))) > 1e-12
    assert (np.abs(p[:70] - thr) < 1e-4).mean() > 0.9          # the band really is populated
    assert np.array_equal(got[0][decidable], want[0][decidable])
    assert 0 < got[0, :70].mean() < 1


@pytest.mark.gpu
def test_blend_feeds_the_colour_transfer_without_a_file_round_trip():
    """f-1 -> hot path on the device: the uint8 image the blend emits is what save_image would have
    written (before JPEG), and goes straight into regional_color_transfer as the stylised input."""
    import applied_image_processing_b200 as aip
    from applied_image_processing_b200 import adjacent
    from oracle import oracle as orc
    from tests import synth

    H, W = 96, 128
    content8 = synth.synth_image(H, W, 3, "N")
    content = torch.from_numpy(content8).cuda().permute(2, 0, 1).float().div(255)[None]
    output = torch.rand((1, 3, 64, 86), device="cuda")
    bg = synth.ellipse_mask(H, W, 0.4, 4)
    _, rgb8 = adjacent.adain_mask_blend(content, output, bg[None].astype(np.float32), return_uint8=True)
    out = aip.regional_color_transfer(torch.from_numpy(content8).cuda(), rgb8[0], torch.from_numpy(bg).cuda())
    ref, st, _ = orc.transfer(content8, rgb8[0].cpu().numpy(), bg)
    assert st == 0 and np.array_equal(out.cpu().numpy(), ref)
