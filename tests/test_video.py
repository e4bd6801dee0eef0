"""warp_image / blend_images of the reference's video driver
(video/utils.py:89-105, 217-229).

Pin: tests/golden/video_f4.npz holds outputs of the reference's own, unmodified functions
(tests/golden/make_golden_video.py executes their definitions from /root/reference against the
cv2 of the build image).  The numpy oracle is checked against them (and against the live cv2 when
it is importable) without a GPU; the CUDA kernels against both on the GPU.  Everything here is
integer / float32-with-a-written-order arithmetic, so the bar is bit-exact."""
import os

import numpy as np
import pytest

from oracle import adjacent_oracle as aorc

GOLD = os.path.join(os.path.dirname(__file__), "golden", "video_f4.npz")
CASES = ("small_flow", "reflect", "ties", "one_row", "one_col")
ALPHAS = (0.7, 0.3, 0.0, 1.0, 0.123456789)


@pytest.fixture(scope="module")
def gold():
    return np.load(GOLD)


@pytest.mark.parametrize("case", CASES)
def test_oracle_warp_matches_reference_golden(gold, case):
    got = aorc.warp_image(gold[f"{case}.image"], gold[f"{case}.flow"])
    assert np.array_equal(got, gold[f"{case}.warped"])


@pytest.mark.parametrize("case", CASES)
def test_oracle_blend_matches_reference_golden(gold, case):
    for a in ALPHAS:
        got = aorc.blend_images(gold[f"{case}.stylized"], gold[f"{case}.warped"], a)
        assert np.array_equal(got, gold[f"{case}.blend_{a}"]), a


def test_oracle_warp_matches_live_cv2():
    cv2 = pytest.importorskip("cv2")
    rng = np.random.default_rng(7)
    for H, W, amp in [(37, 53, 3.0), (120, 200, 30.0), (33, 47, 400.0), (2, 2, 5.0)]:
        img = rng.integers(0, 256, (H, W, 3), dtype=np.uint8)
        flow = (rng.standard_normal((2, H, W)) * amp).astype(np.float32)
        mx, my = aorc.flow_maps(flow)
        ref = cv2.remap(img, mx, my, interpolation=cv2.INTER_LINEAR, borderMode=cv2.BORDER_REFLECT)
        assert np.array_equal(aorc.warp_image(img, flow), ref), (H, W, amp)


def test_oracle_reflect_rule():
    # fedcba|abcdefgh|hgfedcb
    assert aorc.reflect_index(np.array([-1, -2, -8, -9, 8, 9, 15, 16, 3]), 8).tolist() == [0, 1, 7, 7, 7, 6, 0, 0, 3]
    assert aorc.reflect_index(np.array([-5, 0, 7]), 1).tolist() == [0, 0, 0]


def test_blend_identity_properties():
    rng = np.random.default_rng(3)
    s = rng.integers(0, 256, (31, 17, 3), dtype=np.uint8)
    w = rng.integers(0, 256, (31, 17, 3), dtype=np.uint8)
    # alpha = 1 keeps the stylised frame, alpha = 0 the warped one, up to the float32 x/255*255 round trip
    for a, keep in ((1.0, s), (0.0, w)):
        out = aorc.blend_images(s, w, a)
        assert np.abs(out.astype(int) - keep.astype(int)).max() <= 1


# ---- GPU -----------------------------------------------------------------------------------------
@pytest.mark.gpu
@pytest.mark.parametrize("case", CASES)
def test_gpu_warp_and_blend_match_golden(gold, case):
    import applied_image_processing_b200 as aip

    img, sty, flow = gold[f"{case}.image"], gold[f"{case}.stylized"], gold[f"{case}.flow"]
    warped = aip.warp_image(img, flow)
    assert warped.dtype == np.uint8 and np.array_equal(warped, gold[f"{case}.warped"])
    for a in ALPHAS:
        want = gold[f"{case}.blend_{a}"]
        assert np.array_equal(aip.blend_images(sty, warped, a), want), a
        assert np.array_equal(aip.warp_blend(sty, img, flow, a), want), a


@pytest.mark.gpu
def test_gpu_warp_blend_batch_1080p_vs_oracle():
    import torch

    import applied_image_processing_b200 as aip

    rng = np.random.default_rng(11)
    n, H, W = 2, 1080, 1920
    prev = rng.integers(0, 256, (n, H, W, 3), dtype=np.uint8)
    sty = rng.integers(0, 256, (n, H, W, 3), dtype=np.uint8)
    # smooth flow of a few pixels plus a region that leaves the frame (reflection)
    yy, xx = np.mgrid[0:H, 0:W].astype(np.float32)
    flow = np.stack([np.stack([6 * np.sin(yy / 97 + k) + 0.25 * rng.standard_normal((H, W)),
                               4 * np.cos(xx / 131 - k) + 0.25 * rng.standard_normal((H, W))]) for k in range(n)])
    flow[:, :, :40, :60] += 90.0
    flow = flow.astype(np.float32)
    dev = torch.device("cuda", 0)
    t_prev, t_sty, t_flow = (torch.from_numpy(x).to(dev) for x in (prev, sty, flow))
    warped = aip.warp_image(t_prev, t_flow)
    out = aip.warp_blend(t_sty, t_prev, t_flow, 0.7)
    assert warped.is_cuda and out.is_cuda
    for k in range(n):
        want_w = aorc.warp_image(prev[k], flow[k])
        assert np.array_equal(warped[k].cpu().numpy(), want_w)
        assert np.array_equal(out[k].cpu().numpy(), aorc.blend_images(sty[k], want_w, 0.7))
    # flat blend on a length that is not a multiple of 16, and on an unaligned view
    a = t_sty.reshape(-1)[: 1000003].clone()
    b = t_prev.reshape(-1)[: 1000003].clone()
    got = aip.blend_images(a, b, 0.3).cpu().numpy()
    assert np.array_equal(got, aorc.blend_images(a.cpu().numpy(), b.cpu().numpy(), 0.3))


@pytest.mark.gpu
def test_gpu_blend_sequence_is_the_reference_frame_loop():
    import applied_image_processing_b200.video as video

    rng = np.random.default_rng(21)
    n, H, W = 5, 90, 160
    sty = rng.integers(0, 256, (n, H, W, 3), dtype=np.uint8)
    flows = (rng.standard_normal((n - 1, 2, H, W)) * 4).astype(np.float32)
    got = video.blend_sequence(sty, flows, 0.7)
    prev = sty[0]
    assert np.array_equal(got[0], prev)
    for i in range(1, n):  # video/utils.py:355-371: prev_stylized is the *blended* previous frame
        prev = aorc.blend_images(sty[i], aorc.warp_image(prev, flows[i - 1]), 0.7)
        assert np.array_equal(got[i], prev), i


@pytest.mark.gpu
def test_gpu_warp_grey_and_rgba():
    import applied_image_processing_b200 as aip

    rng = np.random.default_rng(5)
    H, W = 61, 45
    flow = (rng.standard_normal((2, H, W)) * 8).astype(np.float32)
    for ch in (1, 4):
        img = rng.integers(0, 256, (H, W, ch), dtype=np.uint8)
        assert np.array_equal(aip.warp_image(img, flow), aorc.warp_image(img, flow))
    grey = rng.integers(0, 256, (H, W), dtype=np.uint8)
    assert np.array_equal(aip.warp_image(grey, flow), aorc.warp_image(grey, flow))


@pytest.mark.gpu
def test_gpu_video_argument_errors():
    import applied_image_processing_b200 as aip

    img = np.zeros((8, 8, 3), np.uint8)
    with pytest.raises(ValueError):
        aip.warp_image(img, np.zeros((2, 7, 8), np.float32))
    with pytest.raises(TypeError):
        aip.blend_images(img.astype(np.float32), img, 0.5)
    with pytest.raises(ValueError):
        aip.blend_images(img, img[:4], 0.5)
