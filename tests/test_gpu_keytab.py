"""The key table in front of the colour LUT (csrc/rct_keytab.cuh): bit-identical output with and without it, every
distinct target colour of a frame re-checked against the LUT, the measured fp32 key error against its bound."""
import numpy as np
import pytest
import torch

from applied_image_processing_b200 import _cabi
from applied_image_processing_b200 import localized_style_transfer as lst
from oracle import oracle as orc
from tests import fullsize, synth

pytestmark = pytest.mark.gpu
EPS_KEY = 2.0 ** -16


def _run(plan, c, s, mask):
    dc, ds, dm = (torch.from_numpy(x).cuda() for x in (c, s, mask))
    out = torch.zeros_like(dc)
    _cabi.check(_cabi.lib().rct_transfer_async(plan.handle, dc.data_ptr(), c.size, ds.data_ptr(), s.size, dm.data_ptr(), mask.size,
                                               out.data_ptr(), c.size, 1, _cabi.MODE_MASK, 0, torch.cuda.current_stream().cuda_stream))
    torch.cuda.synchronize()
    return out.cpu().numpy()


@pytest.mark.parametrize("H,W,kind,fg", [(272, 480, "U", 0.4), (272, 480, "N", 0.4), (544, 960, "U", 0.8), (64, 96, "N", 0.5),
                                         (1080, 1920, "U", 0.4), (1080, 1920, "N", 0.1)])
def test_keytab_is_exact_and_used(H, W, kind, fg):
    c, s = synth.synth_image(H, W, 21, kind), synth.synth_image(H, W, 22, kind)
    mask = synth.ellipse_mask(H, W, fg, 5)
    ref, st, _ = orc.transfer(c, s, mask)
    assert st == 0
    outs = {}
    for kt in (1, 0):
        plan = lst.Plan(torch.device("cuda", 0), H, W, H, W, slots=1)
        plan.set_option(_cabi.OPT_KEYTAB, kt)
        outs[kt] = _run(plan, c, s, mask)
        assert np.array_equal(outs[kt], ref), (kt, int((outs[kt] != ref).sum()))
        chk = plan.debug_keytab()
        if kt:
            assert chk["valid"] == 1 and chk["runs"] > 0 and chk["cells"] > 60000
            assert chk["colours_wrong"] == 0 and chk["pixels_wrong"] == 0, chk
            assert chk["max_key_err"] < EPS_KEY / 2, chk            # the bound the builder widens every cell by
            assert chk["pixels_flagged"] < 0.25 * (chk["pixels_pure"] + chk["pixels_flagged"]), chk
        else:
            assert chk["valid"] == 0


def test_keytab_4k_bench_frame():
    spec, rec = fullsize.CASES["fs_4k_U"], fullsize.golden()["fs_4k_U"]
    c, s, mask = fullsize.inputs(spec)
    plan = lst.Plan(torch.device("cuda", 0), spec["H"], spec["W"], spec["sh"], spec["sw"], slots=1)
    out = _run(plan, c, s, mask)
    assert fullsize.sha256(out) == rec["sha256"]
    chk = plan.debug_keytab()
    assert chk["valid"] == 1 and chk["colours_wrong"] == 0 and chk["max_key_err"] < EPS_KEY / 2, chk
    assert chk["pixels_flagged"] < 0.2 * rec["n_fg"], chk


def test_keytab_degenerate_and_falls_back():
    """Flat target (one run), two colours, a frame with an empty region in a batch, non-interleaved match."""
    H, W = 304, 400
    s = synth.synth_image(H, W, 1, "N")
    mask = synth.ellipse_mask(H, W, 0.5, 3)
    flat = np.empty((H, W, 3), np.uint8)
    flat[...] = (200, 31, 7)
    two = flat.copy()
    two[::3, ::2] = (10, 200, 90)
    for c in (flat, two):
        ref, st, _ = orc.transfer(c, s, mask)
        plan = lst.Plan(torch.device("cuda", 0), H, W, H, W, slots=1)
        assert st == 0 and np.array_equal(_run(plan, c, s, mask), ref)
        assert plan.debug_keytab()["colours_wrong"] == 0
    c = synth.synth_image(H, W, 9, "U")
    ref, _, _ = orc.transfer(c, s, mask)
    plan = lst.Plan(torch.device("cuda", 0), H, W, H, W, slots=1)
    plan.set_option(_cabi.OPT_MATCH, 0)                  # contiguous traversal records no runs: LUT only
    assert np.array_equal(_run(plan, c, s, mask), ref) and plan.debug_keytab()["valid"] == 0
