"""CPU: the C oracle against the reference's outputs at the benchmarked sizes (tests/golden/
fullsize.json, config1.npz; generator tests/golden/make_golden_fullsize.py runs the unmodified
reference).  sha256 of the uint8 result must be equal: the oracle differed from the reference in 0
bytes on every case when the fixtures were made, so equality pins it byte for byte."""
import numpy as np
import pytest

from oracle import oracle as orc
from tests import fullsize

TOL_PCA = 1e-9
G = fullsize.golden()


def _check_pca(d, rec):
    for k_o, k_g in (("t_mean", "fg_mean"), ("t_comp", "fg_comp"), ("s_mean", "bg_mean"), ("s_comp", "bg_comp")):
        np.testing.assert_allclose(d[k_o], fullsize.unhex(rec[k_g]), atol=TOL_PCA, rtol=0)


@pytest.mark.parametrize("name", [n for n in fullsize.CASES if not fullsize.CASES[n].get("labels")])
def test_oracle_equals_reference_hash(name):
    rec = G[name]
    assert rec["oracle_bytes_differing"] == 0
    c, s, m = fullsize.inputs(fullsize.CASES[name])
    out, st, d = orc.transfer(c, s, m)
    assert st == 0 and d["n_t"] == rec["n_fg"] and d["n_s"] == rec["n_bg"]
    _check_pca(d, rec)
    assert fullsize.sha256(out) == rec["sha256"], f"{name}: oracle output differs from the reference's"


def test_oracle_equals_reference_hash_labels():
    name = "fs_1080p_labels8"
    rec = G[name]
    c, s, m = fullsize.inputs(fullsize.CASES[name])
    out = c.copy()
    for lab in range(8):
        _, st, d = orc.transfer(c, s, m, mode=orc.MODE_LABEL, label=lab, out=out)
        per = rec["per_label"][lab]
        assert st == 0 and d["n_t"] == per["n_fg"] and d["n_s"] == per["n_bg"]
        _check_pca(d, per)
    assert fullsize.sha256(out) == rec["sha256"]


def test_oracle_config1_literal():
    cases, g = fullsize.load_config1()
    for tag, c, s, mask, combined in cases:
        out, st, _ = orc.transfer(c, s, mask)
        assert st == 0
        assert np.array_equal(out, combined), f"config 1 ({tag}): {(out != combined).sum()} bytes differ from the reference"


@pytest.mark.parametrize("tag", list(fullsize.CTF_SHAPE_CASES))
def test_color_transfer_foreground_with_different_shapes(tag):
    """LST:128-168 called with fg.shape != bg.shape: all non-black pixels of each image, no resize."""
    fg, bg = fullsize.ctf_inputs(*fullsize.CTF_SHAPE_CASES[tag])
    want = fullsize.load_ctf_shapes()[tag]
    out, st, _ = orc.transfer(fg, bg, None, mode=orc.MODE_ALL)
    assert st == 0 and out.shape == fg.shape and np.array_equal(out, want)
