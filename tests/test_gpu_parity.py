"""Parity of the CUDA path (through the C ABI / the drop-in Python surface) against the C
oracle and the committed reference outputs.  Needs a B200: every test is marked gpu.

Bars (BASELINE.json north_star): histograms, region counts, CDF indices, resample tables and
the uint8 output BIT-EXACT vs the oracle; float lab / PCA / RGB within 1e-4 (far tighter in
practice) and <= 1 LSB vs the reference's own outputs (tests/golden)."""
import numpy as np
import pytest
import torch

import applied_image_processing_b200 as aip
from applied_image_processing_b200 import _cabi
from applied_image_processing_b200 import localized_style_transfer as lst
from oracle import oracle as orc
from tests import golden_io, synth

pytestmark = pytest.mark.gpu

TOL_RGBF = 1e-4  # max-abs in [0,1] RGB (north_star)
TOL_PCA = 1e-9


def _dev():
    return torch.device("cuda", 0)


def _debug_plan(H, W, sh, sw):
    return lst.Plan(_dev(), H, W, sh, sw, slots=1, flags=_cabi.FLAG_DEBUG)


def _run_plan(plan, c, s, mask, mode=_cabi.MODE_MASK, label=0):
    L = _cabi.lib()
    dc, ds = torch.from_numpy(c).cuda(), torch.from_numpy(s).cuda()
    dm = torch.from_numpy(mask).cuda() if mask is not None else None
    out = torch.zeros_like(dc)
    _cabi.check(L.rct_transfer_async(plan.handle, dc.data_ptr(), c.size, ds.data_ptr(), s.size,
                                     dm.data_ptr() if dm is not None else None, mask.size if mask is not None else 0,
                                     out.data_ptr(), c.size, 1, mode, label, torch.cuda.current_stream().cuda_stream))
    torch.cuda.synchronize()
    return out.cpu().numpy()


# ---- the fused pipeline -------------------------------------------------------------------------
@pytest.mark.parametrize("name", golden_io.PIPELINE_NAMES)
def test_pipeline_vs_reference_golden_and_oracle(name):
    c, s, mask, g = golden_io.load_pipeline(name)
    out = aip.regional_color_transfer(c, s, mask)
    assert out.dtype == np.uint8 and out.shape == c.shape
    diff = np.abs(out.astype(np.int16) - g["combined"].astype(np.int16))
    assert diff.max() <= 1, f"{name}: {diff.max()} LSB from the reference"
    ref, st, _ = orc.transfer(c, s, mask)
    assert st == 0 and np.array_equal(out, ref), f"{name}: {(out != ref).sum()} bytes differ from the oracle"


@pytest.mark.parametrize("H,W,sh,sw,ck,sk,mk", [
    (270, 480, 270, 480, "U", "U", "ellipse"),
    (270, 480, 200, 300, "U", "N", "ellipse"),
    (243, 364, 256, 384, "N", "N", "bernoulli"),
    (97, 131, 97, 131, "N", "U", "ellipse"),      # odd sizes: unaligned rows, ragged tail group
    (540, 960, 540, 960, "U", "U", "bernoulli"),
])
def test_integer_artefacts_bit_exact(H, W, sh, sw, ck, sk, mk):
    c, s = synth.synth_image(H, W, 60, ck), synth.synth_image(sh, sw, 61, sk)
    mask = synth.ellipse_mask(H, W, 0.4, 4) if mk == "ellipse" else synth.bernoulli_mask(H, W, 0.5, 4)
    plan = _debug_plan(H, W, sh, sw)
    out = _run_plan(plan, c, s, mask)
    ref, st, d = orc.transfer(c, s, mask, debug=True)
    info = plan.info()
    assert info["status"] == st == 0
    assert info["n_target"] == d["n_t"] and info["n_source"] == d["n_s"]       # region counts
    ht, hs = plan.debug_hist()
    assert np.array_equal(ht, d["hist_t"]) and np.array_equal(hs, d["hist_s"])  # colour histograms
    assert info["d_target"] == int((ht > 0).sum()) and info["d_source"] == int((hs > 0).sum())
    for k in ("t_mean", "t_comp", "s_mean", "s_comp"):                         # PCA: same canonical arithmetic
        assert np.array_equal(info[k], d[k]), (k, info[k], d[k])
    # per distinct target colour: key, CDF index, matched value -- against the per-pixel oracle
    t = plan.debug_target()
    assert np.all(np.diff(t["key"]) >= 0)
    colour = (c[..., 0].astype(np.uint32) | (c[..., 1].astype(np.uint32) << 8) | (c[..., 2].astype(np.uint32) << 16))
    sel = d["j"] >= 0
    order = np.argsort(t["colour"])
    pos = order[np.searchsorted(t["colour"][order], colour[sel])]
    assert np.array_equal(t["colour"][pos], colour[sel])
    assert np.array_equal(t["key"][pos], d["key"][sel])
    assert np.array_equal(t["j"][pos], d["j"][sel]), "CDF indices differ from the oracle"
    assert np.array_equal(t["matched"][pos], d["matched"][sel])
    assert int(t["cum"][-1]) == d["n_t"]
    assert np.array_equal(out, ref), f"{(out != ref).sum()} output bytes differ"


def test_empty_regions_and_black_pixels():
    H, W = 64, 96
    c, s = synth.synth_image(H, W, 1, "U", 0.05), synth.synth_image(H, W, 2, "U", 0.05)
    for mask, want_status in ((np.ones((H, W), np.uint8), 1), (np.zeros((H, W), np.uint8), 2)):
        out = aip.regional_color_transfer(c, s, mask)
        ref, st, _ = orc.transfer(c, s, mask)
        assert st == want_status and np.array_equal(out, ref)
    # black foreground pixels stay black, black background pixels are passed through
    mask = synth.ellipse_mask(H, W, 0.5, 3)
    out = aip.regional_color_transfer(c, s, mask)
    blk = (c.sum(-1) == 0) & (mask == 0)
    assert blk.any() and np.all(out[blk] == 0)
    assert np.array_equal(out[mask == 1], s[mask == 1])


@pytest.mark.parametrize("H,W", [(1, 1), (1, 7), (3, 5), (16, 16), (17, 33)])
def test_tiny_images(H, W):
    c, s = synth.synth_image(H, W, 5, "U", 0), synth.synth_image(H, W, 6, "U", 0)
    mask = synth.bernoulli_mask(H, W, 0.5, 7)
    out = aip.regional_color_transfer(c, s, mask)
    ref, _, _ = orc.transfer(c, s, mask)
    assert np.array_equal(out, ref)


def test_color_transfer_foreground_signature_and_result():
    c, s, mask, g = golden_io.load_pipeline("pipeline_u_same")
    fg = c * (1 - mask)[..., None]
    bg = s * mask[..., None]
    adj = aip.color_transfer_foreground(fg, bg)
    ref, st, _ = orc.transfer(fg, bg, None, mode=orc.MODE_ALL)
    assert st == 0 and np.array_equal(adj, ref)
    combined = adj * (1 - mask)[..., None] + bg  # LST:238
    assert np.abs(combined.astype(int) - g["combined"].astype(int)).max() <= 1
    assert np.array_equal(fg, c * (1 - mask)[..., None])  # inputs are never mutated


def test_batch_tensor_path_and_slots():
    n, H, W = 5, 120, 160
    cs = np.stack([synth.synth_image(H, W, 100 + i, "U" if i % 2 else "N") for i in range(n)])
    ss = np.stack([synth.synth_image(H, W, 200 + i, "N" if i % 2 else "U") for i in range(n)])
    ms = np.stack([synth.ellipse_mask(H, W, 0.2 + 0.1 * i, i) for i in range(n)])
    ms[3] = 1  # one frame without foreground
    refs = np.stack([orc.transfer(cs[i], ss[i], ms[i])[0] for i in range(n)])
    out_np = aip.regional_color_transfer(cs, ss, ms, slots=2)  # 3 launch groups, ragged last
    assert np.array_equal(out_np, refs)
    out_t = aip.regional_color_transfer(torch.from_numpy(cs).cuda(), torch.from_numpy(ss).cuda(),
                                        torch.from_numpy(ms).cuda(), slots=4)
    assert out_t.is_cuda and np.array_equal(out_t.cpu().numpy(), refs)


def test_label_mode_eight_regions():
    H, W, nl = 135, 240, 8
    c, s = synth.synth_image(H, W, 70, "N"), synth.synth_image(H, W, 71, "U")
    labels = synth.voronoi_labels(H, W, nl, 5)
    out = aip.regional_color_transfer_labels(c, s, labels, nl)
    ref = c.copy()
    for lab in range(nl):
        orc.transfer(c, s, labels, mode=orc.MODE_LABEL, label=lab, out=ref)
    assert np.array_equal(out, ref)


def test_fused_label_batch_equals_per_label_passes_and_oracle():
    """rct_transfer_labels_async (all labels of a frame group side by side) against one pass per label and
    against the oracle; a label that is absent from a frame, labels >= n_labels, shared stylised image."""
    n, H, W, nl = 3, 150, 208, 5
    cs = np.stack([synth.synth_image(H, W, 700 + i, "UN"[i % 2]) for i in range(n)])
    ss = np.stack([synth.synth_image(H, W, 800 + i, "NU"[i % 2]) for i in range(n)])
    ls = np.stack([synth.voronoi_labels(H, W, 7, 40 + i) for i in range(n)])  # labels 0..6: 5 and 6 stay content
    ls[1][ls[1] == 2] = 3                                                      # frame 1 has no label 2
    refs = cs.copy()
    for i in range(n):
        for lab in range(nl):
            orc.transfer(cs[i], ss[i], ls[i], mode=orc.MODE_LABEL, label=lab, out=refs[i])
    tc, ts, tl = (torch.from_numpy(x).cuda() for x in (cs, ss, ls))
    fused = aip.regional_color_transfer_labels(tc, ts, tl, nl, slots=2)        # 2 frames per group: a ragged last group
    seq = aip.regional_color_transfer_labels(tc, ts, tl, nl, fused=False)
    assert np.array_equal(fused.cpu().numpy(), refs), int((fused.cpu().numpy() != refs).sum())
    assert np.array_equal(seq.cpu().numpy(), refs)
    # numpy in -> numpy out, one shared stylised image
    refs1 = cs.copy()
    for i in range(n):
        for lab in range(nl):
            orc.transfer(cs[i], ss[0], ls[i], mode=orc.MODE_LABEL, label=lab, out=refs1[i])
    got = aip.regional_color_transfer_labels(cs, ss[0], ls, nl)
    assert isinstance(got, np.ndarray) and np.array_equal(got, refs1)


def test_row_sharded_phases_equal_single_pass():
    """per-shard histograms summed (what the NCCL all-reduce does) -> identical
    LUT -> per-shard apply == the unsharded result, for any split."""
    H, W = 150, 200
    c, s = synth.synth_image(H, W, 80, "U"), synth.synth_image(120, 170, 81, "N")
    mask = synth.ellipse_mask(H, W, 0.45, 2)
    L = _cabi.lib()
    st = torch.cuda.current_stream().cuda_stream
    dc, ds, dm = (torch.from_numpy(a).cuda() for a in (c, s, mask))
    plans = [lst.Plan(_dev(), H, W, 120, 170) for _ in range(3)]
    bounds = [0, 37, 101, H]
    hists = []
    for r, p in enumerate(plans):
        _cabi.check(L.rct_shard_hist_async(p.handle, dc.data_ptr(), ds.data_ptr(), dm.data_ptr(), 0, 0,
                                           bounds[r], bounds[r + 1] - bounds[r], st))
        n = L.rct_plan_hist_bytes(p.handle) // 4
        hists.append(_as_tensor(L.rct_plan_hist_ptr(p.handle), n))
    total = hists[0] + hists[1] + hists[2]
    out = torch.zeros_like(dc)
    for r, p in enumerate(plans):
        hists[r].copy_(total)
        _cabi.check(L.rct_shard_solve_async(p.handle, st))
        _cabi.check(L.rct_shard_apply_async(p.handle, dc.data_ptr(), ds.data_ptr(), dm.data_ptr(), out.data_ptr(),
                                            0, 0, bounds[r], bounds[r + 1] - bounds[r], st))
    torch.cuda.synchronize()
    ref, _, _ = orc.transfer(c, s, mask)
    assert np.array_equal(out.cpu().numpy(), ref)


def _as_tensor(ptr, n_int32):
    """View library-owned device memory as an int32 tensor (no copy) via __cuda_array_interface__."""
    class _Buf:
        __cuda_array_interface__ = {"shape": (n_int32,), "typestr": "<i4", "data": (ptr, False), "version": 3}
    return torch.as_tensor(_Buf(), device="cuda")


# ---- the reference's stand-alone functions ---------------------------------------------------------
def test_rgb_lab_functions_vs_golden_and_oracle():
    g = golden_io.load("functions")
    lab = aip.rgb_to_lab_pixels(g["px"])
    assert lab.dtype == np.float64 and lab.shape == g["px"].shape
    assert np.array_equal(lab, orc.rgb_to_lab_pixels(g["px"]))            # canonical arithmetic: bit-exact
    assert np.abs(lab - g["lab"]).max() <= 1e-9                           # vs numpy reference
    back, rgbf = aip.lab_to_rgb_pixels(g["lab2"], return_float=True)
    oback, orgbf = orc.lab_to_rgb_pixels(g["lab2"], want_float=True)
    assert np.array_equal(back, oback) and np.array_equal(rgbf, orgbf)
    assert np.abs(back.astype(int) - g["back2"].astype(int)).max() <= 1
    img = g["px"].reshape(32, 64, 3)
    assert np.array_equal(aip.rgb_to_lab_image(img), lab.reshape(32, 64, 3))
    assert np.array_equal(aip.lab_to_rgb_image(g["lab2"].reshape(32, 64, 3)), back.reshape(32, 64, 3))


def test_apply_pca_vs_golden_and_oracle():
    g = golden_io.load("functions")
    proj, pca = aip.apply_pca(g["lab"])
    assert proj.shape == (g["lab"].shape[0], 1) and pca.components_.shape == (1, 3) and pca.mean_.shape == (3,)
    np.testing.assert_allclose(pca.mean_, g["pca_mean"], atol=TOL_PCA, rtol=0)
    np.testing.assert_allclose(pca.components_[0], g["pca_comp"], atol=TOL_PCA, rtol=0)
    np.testing.assert_allclose(proj, g["proj"], atol=1e-9, rtol=0)
    oproj, omean, ocomp = orc.apply_pca(g["lab"])
    assert np.array_equal(proj, oproj) and np.array_equal(pca.mean_, omean) and np.array_equal(pca.components_[0], ocomp)
    inv = pca.inverse_transform(proj)
    np.testing.assert_allclose(inv, g["inv"], atol=1e-9, rtol=0)


@pytest.mark.parametrize("tag", ["eq", "tgt_longer", "src_longer", "one_src", "one_tgt"])
def test_match_cdf_bit_exact_vs_numpy_golden(tag):
    g = golden_io.load("functions")
    t, s = g[f"cdf_{tag}_t"], g[f"cdf_{tag}_s"]
    m, j = aip.match_cdf(t, s, return_index=True)
    assert m.shape == (t.shape[0], 1)
    assert np.array_equal(m, g[f"cdf_{tag}_m"]), np.abs(m - g[f"cdf_{tag}_m"]).max()
    om, oj = orc.match_cdf(t, s, want_index=True)
    assert np.array_equal(j, oj)


def test_match_cdf_many_ties_and_large():
    rng = np.random.default_rng(7)
    t = np.round(rng.normal(0, 1, 200_000), 2)  # heavy ties
    s = rng.normal(1, 3, 150_001)
    m, j = aip.match_cdf(t, s, return_index=True)
    om, oj = orc.match_cdf(t, s, want_index=True)
    assert np.array_equal(j, oj) and np.array_equal(m, om)


# ---- full-size, size-independent properties ---------------------------------------------------------
def test_4k_properties():
    """BASELINE config 2 size: the oracle would take minutes, so check properties the domain
    offers: histogram mass == region counts, background pass-through, LUT consistency (equal
    input colours -> equal output), idempotence of a second run, monotone key->matched map."""
    H, W = 2160, 3840
    c, s = synth.synth_image(H, W, 0, "U"), synth.synth_image(H, W, 1, "U")
    mask = synth.ellipse_mask(H, W, 0.4, 4)
    plan = _debug_plan(H, W, H, W)
    out = _run_plan(plan, c, s, mask)
    info = plan.info()
    fsel = (mask == 0) & (c.sum(-1) > 0)
    bsel = (mask == 1) & (s.sum(-1) > 0)
    assert info["n_target"] == int(fsel.sum()) and info["n_source"] == int(bsel.sum())
    ht, hs = plan.debug_hist()
    assert int(ht.sum(dtype=np.int64)) == info["n_target"] and int(hs.sum(dtype=np.int64)) == info["n_source"]
    assert np.array_equal(out[mask == 1], s[mask == 1])
    assert np.all(out[(mask == 0) & ~fsel] == 0)
    lut = plan.debug_lut()
    col = c[..., 0].astype(np.uint32) | (c[..., 1].astype(np.uint32) << 8) | (c[..., 2].astype(np.uint32) << 16)
    o32 = out[..., 0].astype(np.uint32) | (out[..., 1].astype(np.uint32) << 8) | (out[..., 2].astype(np.uint32) << 16)
    assert np.array_equal(o32[fsel], lut[col[fsel]])
    t = plan.debug_target()
    assert np.all(np.diff(t["key"]) > 0) and np.all(np.diff(t["matched"]) >= 0) and np.all(np.diff(t["j"]) > 0)
    assert int(t["cum"][-1]) == info["n_target"]
    out2 = _run_plan(plan, c, s, mask)
    assert np.array_equal(out, out2)
    # a 1/64 sub-sample of rows run through the oracle has (nearly) the same statistics: the
    # PCA axes agree to sampling error -- guards against a grossly wrong full-size reduction
    _, _, d = orc.transfer(c[::8, ::8].copy(), s[::8, ::8].copy(), mask[::8, ::8].copy())
    assert np.abs(info["t_mean"] - d["t_mean"]).max() < 0.02
    assert abs(abs(np.dot(info["t_comp"], d["t_comp"])) - 1) < 1e-3


# ---- the row-sharded driver (sharded.py) with two ranks on one GPU ---------------------------------
def _shard_worker(rank, world, port, H, W, sh, sw, q):
    import os

    import torch.distributed as dist

    from applied_image_processing_b200 import sharded

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(0)
    dist.init_process_group("gloo", rank=rank, world_size=world)  # NCCL refuses two ranks on one GPU
    try:
        c, s = synth.synth_image(H, W, 90, "U"), synth.synth_image(sh, sw, 91, "N")
        mask = synth.ellipse_mask(H, W, 0.45, 2)
        out, (r0, n) = sharded.regional_color_transfer_sharded(c, s, mask)
        q.put((rank, r0, n, out))
    finally:
        dist.destroy_process_group()


def test_sharded_driver_two_ranks_one_gpu():
    import socket

    import torch.multiprocessing as mp

    H, W, sh, sw, world = 151, 203, 120, 170, 2
    with socket.socket() as sk:
        sk.bind(("127.0.0.1", 0))
        port = sk.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_shard_worker, args=(r, world, port, H, W, sh, sw, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = [q.get(timeout=240) for _ in range(world)]
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    c, s = synth.synth_image(H, W, 90, "U"), synth.synth_image(sh, sw, 91, "N")
    ref, st, _ = orc.transfer(c, s, synth.ellipse_mask(H, W, 0.45, 2))
    assert st == 0
    full = np.zeros_like(ref)
    for rank, r0, n, out in got:
        full[r0:r0 + n] = out
    assert np.array_equal(full, ref)


def test_sharded_driver_without_a_mask_in_mode_all():
    """MODE_ALL has no mask (LST:128-168 on whole images): the row-sharded drivers pass NULL like the unsharded paths,
    and a MASK call without one is refused by the library instead of dereferencing None."""
    from applied_image_processing_b200 import sharded

    H, W, sh, sw = 90, 120, 90, 120
    c, s = synth.synth_image(H, W, 92, "N"), synth.synth_image(sh, sw, 93, "N")
    dc, ds = torch.from_numpy(c).cuda(), torch.from_numpy(s).cuda()
    drv = sharded.ShardedTransfer(H, W, sh, sw)
    out = drv(dc, ds, None, mode=_cabi.MODE_ALL)
    outb = drv.binned(dc, ds, None, bins=1024, mode=_cabi.MODE_ALL)
    torch.cuda.synchronize()
    ref, st, _ = orc.transfer(c, s, None, mode=orc.MODE_ALL)
    refb, stb, _ = orc.transfer_binned(c, s, None, 1024, mode=orc.MODE_ALL)
    assert st == 0 and np.array_equal(out.cpu().numpy(), ref)
    assert stb == 0 and np.array_equal(outb.cpu().numpy(), refb)
    with pytest.raises(_cabi.RctError, match="mask required"):
        drv(dc, ds, None, mode=_cabi.MODE_MASK)


@pytest.mark.parametrize("slots", [1, 2])
def test_host_entry_point_pageable_pinned_and_mixed_buffers(slots):
    """rct_transfer_host stages pageable arrays (what np.array(Image.open(...)) gives, LST:200-201) through a pinned ring
    with host threads and takes pinned arrays directly: every combination gives the oracle's bytes.  7 frames in groups
    of `slots` frames = more groups than staging sets, so every pinned set is reused, and an odd tail group."""
    n, H, W, sh, sw = 7, 96, 136, 64, 90
    cs = np.stack([synth.synth_image(H, W, 300 + i, "U") for i in range(n)])
    ss = np.stack([synth.synth_image(sh, sw, 320 + i, "N") for i in range(n)])
    ms = np.stack([synth.ellipse_mask(H, W, 0.4, 340 + i) for i in range(n)])
    ref = np.stack([orc.transfer(cs[i], ss[i], ms[i])[0] for i in range(n)])

    def pinned(a):
        return torch.from_numpy(a).pin_memory().numpy()

    pc, ps, pm, po = pinned(cs), pinned(ss), pinned(ms), pinned(np.zeros_like(cs))
    assert np.array_equal(aip.regional_color_transfer(cs, ss, ms, slots=slots), ref)                   # all pageable
    assert np.array_equal(aip.regional_color_transfer(pc, ps, pm, slots=slots, out=po), ref)           # all pinned
    assert np.array_equal(aip.regional_color_transfer(pc, ss, pm, slots=slots), ref)                   # mixed in, pageable out
    po[:] = 0
    assert np.array_equal(aip.regional_color_transfer(cs, ps, ms, slots=slots, out=po), ref)           # pageable in, pinned out
    # LABEL mode reads the output array as an input too (untouched pixels keep the caller's values)
    labels = np.stack([synth.voronoi_labels(H, W, 3, 360 + i) for i in range(n)])
    want = cs.copy()
    for i in range(n):
        orc.transfer(cs[i], ss[i], labels[i], mode=orc.MODE_LABEL, label=1, out=want[i])
    for out in (cs.copy(), pinned(cs.copy())):
        got = lst._transfer(cs, ss, labels, _cabi.MODE_LABEL, label=1, out=out, slots=slots, quiet=True)
        assert np.array_equal(got, want)


# ---- exact sort: fallback paths ---------------------------------------------------------------------
@pytest.mark.parametrize("cap", [1, 64, 700])
def test_sort_global_fallback_is_exact(cap):
    """Lower the shared-memory part capacity so that parts take the global-memory bitonic
    fallback of the exact sort; results must not change."""
    H, W = 180, 320
    c, s = synth.synth_image(H, W, 40, "U"), synth.synth_image(H, W, 41, "N")
    mask = synth.ellipse_mask(H, W, 0.55, 9)
    plan = _debug_plan(H, W, H, W)
    plan.set_option(_cabi.OPT_PART_CAP, cap)
    out = _run_plan(plan, c, s, mask)
    ref, st, d = orc.transfer(c, s, mask, debug=True)
    assert st == 0 and np.array_equal(out, ref)
    t = plan.debug_target()
    assert np.all(np.diff(t["key"]) >= 0) and int(t["cum"][-1]) == d["n_t"]


def test_match_cdf_degenerate_key_sets():
    """All keys equal / two clusters / huge dynamic range: the equalised partition degenerates
    (one fine bin holds everything) and the bitonic fallbacks take over."""
    rng = np.random.default_rng(3)
    cases = [
        (np.full(9000, 0.25), rng.normal(0, 1, 5000)),
        (rng.normal(0, 1, 7000), np.full(12000, -3.0)),
        (np.concatenate([np.full(6000, 1.0), np.full(6000, 1.0 + 1e-12)]), rng.uniform(0, 1, 12000)),
        (np.concatenate([rng.normal(0, 1e-9, 20000), [1e6, -1e6]]), rng.normal(0, 1, 30000)),
        (rng.integers(0, 3, 50000).astype(np.float64), rng.integers(0, 5, 40000).astype(np.float64)),
    ]
    for t, s in cases:
        m, j = aip.match_cdf(t, s, return_index=True)
        om, oj = orc.match_cdf(t, s, want_index=True)
        assert np.array_equal(j, oj) and np.array_equal(m, om)


# ---- front ends (colour partition vs colour cube) and apply kernels (bulk-async vs register) ----------
def _flat_pair(H, W):
    """One colour per region on > 65536 pixels: every item lands in one bucket, which is split
    over several CTAs that merge through the cube window (rct_front.cuh)."""
    c = np.empty((H, W, 3), np.uint8)
    c[...] = (200, 31, 7)
    s = np.empty((H, W, 3), np.uint8)
    s[...] = (9, 140, 77)
    c[::7, ::5] = (200, 31, 8)     # a second colour in the same bucket
    s[::3, ::11] = (250, 1, 1)     # and one in a different bucket
    return c, s


@pytest.mark.parametrize("case", ["U", "N", "flat", "resample", "ragged"])
def test_front_ends_and_apply_kernels_agree(case):
    if case == "flat":
        H, W = 400, 512
        c, s = _flat_pair(H, W)
        sh, sw = H, W
    elif case == "resample":
        H, W, sh, sw = 300, 416, 211, 307
        c, s = synth.synth_image(H, W, 11, "N"), synth.synth_image(sh, sw, 12, "U")
    elif case == "ragged":
        H, W, sh, sw = 251, 333, 251, 333  # npix % 16 != 0: bulk-async kernel + scalar tail
        c, s = synth.synth_image(H, W, 13, "U"), synth.synth_image(sh, sw, 14, "N")
    else:
        H, W, sh, sw = 360, 640, 360, 640
        c, s = synth.synth_image(H, W, 15, case), synth.synth_image(sh, sw, 16, case)
    mask = synth.ellipse_mask(H, W, 0.4, 4)
    ref, st, d = orc.transfer(c, s, mask, debug=True)
    assert st == 0
    for front in (1, 0):
        for apply_tma in (1, 7, 6, 2, 5, 4, 3, 0):
            plan = _debug_plan(H, W, sh, sw)
            plan.set_option(_cabi.OPT_FRONTEND, front)
            plan.set_option(_cabi.OPT_APPLY, apply_tma)
            out = _run_plan(plan, c, s, mask)
            out2 = _run_plan(plan, c, s, mask)  # second run: the cube windows were left zero
            info = plan.info()
            assert info["n_target"] == d["n_t"] and info["n_source"] == d["n_s"], (front, apply_tma)
            ht, hs = plan.debug_hist()
            assert np.array_equal(ht, d["hist_t"]) and np.array_equal(hs, d["hist_s"]), (front, apply_tma)
            assert np.array_equal(out, ref), (front, apply_tma, int((out != ref).sum()))
            assert np.array_equal(out2, ref), (front, apply_tma, "second run")


@pytest.mark.parametrize("H,W,sh,sw", [
    (300, 416, 208, 288),    # up-sampling 1.44x (stylised at a lower scale: the reference's usual case)
    (300, 512, 360, 624),    # down-sampling 1.22x: wide spans, W % 256 == 0
    (512, 1024, 128, 256),   # up-sampling 4x, W % 256 == 0
    (200, 300, 128, 192),    # W % 16 != 0: tiles cross rows at arbitrary x
    (201, 301, 128, 192),    # ragged tail of 5 pixels through the register-staged kernel
    (64, 256, 64, 128),      # W == 256: one tile per row
])
def test_tma_staged_resample_apply(H, W, sh, sw):
    """k_apply_warp_rs (bulk-async staged NEAREST resample, opt-in RCT_OPT_APPLY=8, W >= 256) and the
    register-staged resample kernel (the default) against the oracle, single frames and a strided batch."""
    n = 3
    cs = np.stack([synth.synth_image(H, W, 500 + i, "UN"[i % 2]) for i in range(n)])
    ss = np.stack([synth.synth_image(sh, sw, 600 + i, "NU"[i % 2]) for i in range(n)])
    ms = np.stack([synth.ellipse_mask(H, W, 0.3 + 0.15 * i, i) if i else synth.bernoulli_mask(H, W, 0.5, 9)
                   for i in range(n)])
    refs = np.stack([orc.transfer(cs[i], ss[i], ms[i])[0] for i in range(n)])
    for apply_opt in (8, 1, 0):
        plan = _debug_plan(H, W, sh, sw)
        plan.set_option(_cabi.OPT_APPLY, apply_opt)
        for i in range(n):
            out = _run_plan(plan, cs[i], ss[i], ms[i])
            assert np.array_equal(out, refs[i]), (apply_opt, i, int((out != refs[i]).sum()))
    out = aip.regional_color_transfer(torch.from_numpy(cs).cuda(), torch.from_numpy(ss).cuda(),
                                      torch.from_numpy(ms).cuda(), slots=2)
    assert np.array_equal(out.cpu().numpy(), refs)


@pytest.mark.parametrize("case", ["U", "N", "resample", "tiny_target"])
def test_lut_layouts_agree(case):
    """RCT_OPT_LUT: uint16 run indices + palette (opt-in) against the uint32 RGB table (default) and the
    oracle, through the bulk-async and the register-staged apply kernels."""
    H, W, sh, sw = 360, 640, 360, 640
    fg = 0.4
    if case == "resample":
        sh, sw = 256, 448
    if case == "tiny_target":
        fg = 0.01
    kind = "N" if case == "N" else "U"
    c, s = synth.synth_image(H, W, 31, kind), synth.synth_image(sh, sw, 32, "N" if case == "resample" else kind)
    mask = synth.ellipse_mask(H, W, fg, 4)
    ref, st, _ = orc.transfer(c, s, mask)
    assert st == 0
    luts = []
    for lut_mode in (1, 0):
        for apply_opt in (1, 0):
            plan = _debug_plan(H, W, sh, sw)
            plan.set_option(_cabi.OPT_LUT, lut_mode)
            plan.set_option(_cabi.OPT_APPLY, apply_opt)
            out = _run_plan(plan, c, s, mask)
            assert np.array_equal(out, ref), (lut_mode, apply_opt, int((out != ref).sum()))
        luts.append(plan.debug_lut())
    col = c[..., 0].astype(np.uint32) | (c[..., 1].astype(np.uint32) << 8) | (c[..., 2].astype(np.uint32) << 16)
    sel = (mask == 0) & (col > 0)
    assert np.array_equal(luts[0][col[sel]], luts[1][col[sel]])
    assert len(np.unique(luts[0][col[sel]])) <= 3 * 3 * 256  # a-7: the outputs lie on one curve


def test_partition_front_end_all_and_label_modes():
    """Two-pass item generation (a pixel feeds both regions) of the partition front end."""
    H, W = 200, 304
    c, s = synth.synth_image(H, W, 21, "N"), synth.synth_image(H, W, 22, "U")
    labels = synth.voronoi_labels(H, W, 4, 6)
    for mode, mask, label in ((_cabi.MODE_ALL, None, 0), (_cabi.MODE_LABEL, labels, 2)):
        outs = []
        for front in (1, 0):
            plan = _debug_plan(H, W, H, W)
            plan.set_option(_cabi.OPT_FRONTEND, front)
            L = _cabi.lib()
            dc, ds = torch.from_numpy(c).cuda(), torch.from_numpy(s).cuda()
            dm = torch.from_numpy(mask).cuda() if mask is not None else None
            out = dc.clone()
            _cabi.check(L.rct_transfer_async(plan.handle, dc.data_ptr(), c.size, ds.data_ptr(), s.size,
                                             dm.data_ptr() if dm is not None else None, H * W, out.data_ptr(), c.size,
                                             1, mode, label, torch.cuda.current_stream().cuda_stream))
            torch.cuda.synchronize()
            outs.append((out.cpu().numpy(), plan.debug_hist()))
        ref = c.copy()
        if mode == _cabi.MODE_ALL:
            ref, st, _ = orc.transfer(c, s, None, mode=orc.MODE_ALL)
        else:
            orc.transfer(c, s, labels, mode=orc.MODE_LABEL, label=label, out=ref)
        for out, _ in outs:
            assert np.array_equal(out, ref)
        assert np.array_equal(outs[0][1][0], outs[1][1][0]) and np.array_equal(outs[0][1][1], outs[1][1][1])


def test_batch_apply_kernels_agree_on_aligned_and_unaligned_strides():
    for (H, W) in ((120, 160), (97, 131)):  # frame stride % 16 == 0 / != 0 (falls back per call)
        n = 4
        cs = np.stack([synth.synth_image(H, W, 300 + i, "U") for i in range(n)])
        ss = np.stack([synth.synth_image(H, W, 400 + i, "N") for i in range(n)])
        ms = np.stack([synth.ellipse_mask(H, W, 0.3 + 0.1 * i, i) for i in range(n)])
        refs = np.stack([orc.transfer(cs[i], ss[i], ms[i])[0] for i in range(n)])
        out = aip.regional_color_transfer(torch.from_numpy(cs).cuda(), torch.from_numpy(ss).cuda(),
                                          torch.from_numpy(ms).cuda(), slots=3)
        assert np.array_equal(out.cpu().numpy(), refs)


@pytest.mark.parametrize("fg", [0.25, 0.5, 0.75])
def test_match_traversals_agree(fg):
    """RCT_OPT_MATCH: lane-interleaved vs one contiguous run per thread; target smaller / equal /
    larger than the source (the three branches of match_cdf, LST:108-121)."""
    H, W = 300, 400
    c, s = synth.synth_image(H, W, 31, "U"), synth.synth_image(H, W, 32, "N")
    mask = synth.ellipse_mask(H, W, fg, 8)
    if fg == 0.5:  # force N_fg == N_bg exactly: equal counts of non-black pixels
        mask = (np.arange(H * W).reshape(H, W) % 2).astype(np.uint8)
        c, s = synth.synth_image(H, W, 31, "U", 0), synth.synth_image(H, W, 32, "N", 0)
        c[c.sum(-1) == 0] = 1
        s[s.sum(-1) == 0] = 1
    ref, st, d = orc.transfer(c, s, mask, debug=True)
    assert st == 0
    colour = (c[..., 0].astype(np.uint32) | (c[..., 1].astype(np.uint32) << 8) | (c[..., 2].astype(np.uint32) << 16))
    sel = d["j"] >= 0
    for order in (1, 0):
        plan = _debug_plan(H, W, H, W)
        plan.set_option(_cabi.OPT_MATCH, order)
        out = _run_plan(plan, c, s, mask)
        assert np.array_equal(out, ref), (order, int((out != ref).sum()))
        t = plan.debug_target()
        o = np.argsort(t["colour"])
        pos = o[np.searchsorted(t["colour"][o], colour[sel])]
        assert np.array_equal(t["j"][pos], d["j"][sel]), order
        assert np.array_equal(t["matched"][pos], d["matched"][sel]), order


def test_randomised_geometries_against_the_oracle():
    """Seeded sweep over sizes (including < 16 pixels, odd rows, W around the 256-pixel tile), stylised
    sizes (same / up- / down-sampled), mask kinds and image families: uint8 output bit-exact vs the
    oracle, statuses equal (empty regions leave the content / stylised pixels alone)."""
    rng = np.random.default_rng(2026)
    kinds = ("U", "N")
    for case in range(40):
        H = int(rng.integers(1, 40)) if case % 5 == 0 else int(rng.integers(40, 320))
        W = int(rng.choice([int(rng.integers(1, 64)), 255, 256, 257, int(rng.integers(64, 400))]))
        if rng.random() < 0.5:
            sh, sw = H, W
        else:
            sh, sw = max(1, int(H * rng.uniform(0.4, 1.6))), max(1, int(W * rng.uniform(0.4, 1.6)))
        c = synth.synth_image(H, W, 9000 + case, kinds[case % 2], black_frac=float(rng.choice([0.0, 0.001, 0.2])))
        s = synth.synth_image(sh, sw, 9500 + case, kinds[(case // 2) % 2], black_frac=float(rng.choice([0.0, 0.01])))
        mk = case % 6
        if mk == 0:
            mask = synth.ellipse_mask(H, W, float(rng.uniform(0.05, 0.95)), case)
        elif mk == 1:
            mask = synth.bernoulli_mask(H, W, float(rng.uniform(0.1, 0.9)), case)
        elif mk == 2:
            mask = np.zeros((H, W), np.uint8)          # everything foreground: no source pixels
        elif mk == 3:
            mask = np.ones((H, W), np.uint8)           # everything background: no target pixels
        elif mk == 4:
            mask = (np.add.outer(np.arange(H) // 7, np.arange(W) // 5) % 2).astype(np.uint8)
        else:
            mask = np.ones((H, W), np.uint8)
            mask[rng.integers(0, H), rng.integers(0, W)] = 0   # a single target pixel
        ref, st, _ = orc.transfer(c, s, mask)
        plan = lst.get_plan(_dev(), H, W, sh, sw, 1)
        out = _run_plan(plan, c, s, mask)
        status = np.zeros(1, np.int32)
        _cabi.check(_cabi.lib().rct_plan_status(plan.handle, status.ctypes.data, 1, torch.cuda.current_stream().cuda_stream))
        assert int(status[0]) == st, (case, H, W, sh, sw, mk, int(status[0]), st)
        assert np.array_equal(out, ref), (case, H, W, sh, sw, mk, int((out != ref).sum()))
    lst.clear_plans()
