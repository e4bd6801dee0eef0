"""Zero-variance target region (one flat colour): a DEFINED DEVIATION from the reference
(VERDICT r1 weak #2).

The reference's covariance of identical rows is pure round-off (~1e-15), so sklearn's
`components_` is whatever OpenBLAS's summation order leaves: it changes with the NUMBER of
(identical) pixels.  No fixed-order restatement can reproduce it, and the output colour of the
whole region depends on it.  This repo's axis comes from exact integer moments and a fixed-order
Jacobi (C5/C6): deterministic, independent of pixel order, sharding and GPU count.  The
near-degenerate neighbours (two colours, collinear colours, flat SOURCE) do match the reference."""
import numpy as np
import pytest

from oracle import oracle as orc
from oracle import ref_shim
from tests import synth


def _flat_case(colour=(200, 31, 7), H=40, W=50):
    c = np.empty((H, W, 3), np.uint8)
    c[...] = colour
    s = synth.synth_image(H, W, 1, "N")
    mask = synth.ellipse_mask(H, W, 0.5, 3)
    return c, s, mask


@pytest.mark.skipif(not ref_shim.available(), reason="needs the live reference (/root/reference)")
def test_reference_axis_of_a_flat_region_is_round_off_noise():
    lst = ref_shim.load_reference()
    c, s, mask = _flat_case()
    fg = c * (1 - mask)[..., None]
    lab = lst.rgb_to_lab_pixels(fg[fg.sum(-1) > 0])
    axes = []
    for n in (100, 333, 800, lab.shape[0]):
        proj, pca = lst.apply_pca(lab[:n])
        assert pca.explained_variance_[0] < 1e-12          # the "variance" is round-off
        assert np.abs(proj).max() < 1e-12                   # every projection is (numerically) zero
        axes.append(pca.components_[0])
    dots = [abs(float(np.dot(axes[0], a))) for a in axes[1:]]
    assert min(dots) < 0.99, "the reference's axis does not depend on n here: revisit the deviation note"


def test_oracle_flat_region_is_deterministic_and_order_independent():
    c, s, mask = _flat_case()
    out, st, d = orc.transfer(c, s, mask)
    assert st == 0
    fsel = mask == 0
    assert len(np.unique(out[fsel].reshape(-1, 3), axis=0)) == 1    # one input colour -> one output colour
    assert np.array_equal(out[~fsel], s[~fsel])
    # permuting the rows (another pixel order, another region size split) leaves the axis unchanged
    out2, _, d2 = orc.transfer(c[::-1].copy(), s[::-1].copy(), mask[::-1].copy())
    assert np.array_equal(d["t_comp"], d2["t_comp"]) and np.array_equal(out2[::-1], out)
    assert abs(np.linalg.norm(d["t_comp"]) - 1) < 1e-12


@pytest.mark.skipif(not ref_shim.available(), reason="needs the live reference (/root/reference)")
def test_near_degenerate_neighbours_match_the_reference():
    H, W = 40, 50
    s = synth.synth_image(H, W, 1, "N")
    mask = synth.ellipse_mask(H, W, 0.5, 3)
    two = np.empty((H, W, 3), np.uint8)
    two[...] = (200, 31, 7)
    two[::3, ::2] = (200, 31, 8)                                     # two adjacent colours
    line = np.empty((H, W, 3), np.uint8)
    line[...] = np.stack([np.linspace(10, 240, W).astype(np.uint8)] * 3, -1)[None]   # collinear greys
    flat_src = np.empty((H, W, 3), np.uint8)
    flat_src[...] = (9, 140, 77)
    for c, sty in ((two, s), (line, s), (synth.synth_image(H, W, 5, "U"), flat_src)):
        ref = ref_shim.ref_pipeline(c, sty, mask)
        out, st, _ = orc.transfer(c, sty, mask)
        assert st == 0 and np.abs(out.astype(int) - ref.astype(int)).max() <= 1


@pytest.mark.gpu
@pytest.mark.parametrize("colour", [(200, 31, 7), (255, 255, 255), (1, 0, 0)])
def test_cuda_flat_region_equals_the_oracle(colour):
    import torch

    import applied_image_processing_b200 as aip
    from applied_image_processing_b200 import _cabi
    from applied_image_processing_b200 import localized_style_transfer as lst

    c, s, mask = _flat_case(colour, 300, 400)       # > 65536 target pixels of one colour: split bucket
    ref, st, d = orc.transfer(c, s, mask)
    out = aip.regional_color_transfer(c, s, mask)
    assert st == 0 and np.array_equal(out, ref)
    plan = lst.Plan(torch.device("cuda", 0), 300, 400, 300, 400, slots=1, flags=_cabi.FLAG_DEBUG)
    dc, ds, dm = (torch.from_numpy(x).cuda() for x in (c, s, mask))
    o = torch.empty_like(dc)
    _cabi.check(_cabi.lib().rct_transfer_async(plan.handle, dc.data_ptr(), c.size, ds.data_ptr(), s.size, dm.data_ptr(),
                                               mask.size, o.data_ptr(), c.size, 1, 0, 0, torch.cuda.current_stream().cuda_stream))
    info = plan.info()
    assert np.array_equal(info["t_comp"], d["t_comp"]) and np.array_equal(info["t_mean"], d["t_mean"])
