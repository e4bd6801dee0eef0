"""BINNED mode (BASELINE.json configs[4] "1024-bin histograms"; no counterpart in the reference: an approximation
with its own definition).  CPU: the C oracle's orc_match_binned against an independent numpy
statement of the same definition, its deviation from the exact path, degenerate inputs.  GPU: the CUDA path against
that oracle bit for bit (histograms, key ranges, CDF index pairs, matched values, uint8 output), the row-sharded
phases against the unsharded call, full size."""
import numpy as np
import pytest

from oracle import oracle as orc
from tests import synth


def _numpy_binned(tkey, skey, B):
    """The definition of rct_binned.cuh / orc_match_binned, vectorised in numpy (same fp64 operations)."""
    lo_t, hi_t, lo_s, hi_s = tkey.min(), tkey.max(), skey.min(), skey.max()
    sc_t = B / (hi_t - lo_t) if hi_t > lo_t else 0.0
    sc_s = B / (hi_s - lo_s) if hi_s > lo_s else 0.0
    bt = np.minimum(((tkey - lo_t) * sc_t).astype(np.int64), B - 1)
    bs = np.minimum(((skey - lo_s) * sc_s).astype(np.int64), B - 1)
    ht, hs = np.bincount(bt, minlength=B), np.bincount(bs, minlength=B)
    ct, cs = np.concatenate([[0], np.cumsum(ht)]), np.concatenate([[0], np.cumsum(hs)])
    pos = (tkey - lo_t) * sc_t
    f = pos - bt
    q = ct[bt] + f * ht[bt]
    qs = (q / float(len(tkey))) * float(len(skey))
    b2 = np.minimum(np.searchsorted(cs[1:].astype(np.float64), qs, side="left"), B - 1)
    with np.errstate(divide="ignore", invalid="ignore"):
        f2 = np.where(hs[b2] > 0, (qs - cs[b2]) / hs[b2], 0.0)
    f2 = np.clip(f2, 0.0, 1.0)
    matched = lo_s + (b2 + f2) * ((hi_s - lo_s) / B)
    return matched, bt * 65536 + b2, ht.astype(np.uint32), hs.astype(np.uint32)


def _keys_of(c, s, mask):
    """Per-pixel keys of both regions in row-major order, from the exact oracle's debug output."""
    _, st, d = orc.transfer(c, s, mask, debug=True)
    assert st == 0
    tkey = d["key"][~np.isnan(d["key"])]
    return tkey, d


@pytest.mark.parametrize("kind,B", [("U", 256), ("U", 1024), ("N", 1024), ("N", 4096), ("U", 2)])
def test_oracle_binned_equals_numpy_statement(kind, B):
    H, W = 120, 200
    c, s = synth.synth_image(H, W, 3, kind), synth.synth_image(H, W, 4, kind)
    mask = synth.ellipse_mask(H, W, 0.4, 5)
    out, st, d = orc.transfer_binned(c, s, mask, B, debug=True)
    assert st == 0
    assert int(d["hist_t"].sum()) == d["n_t"] and int(d["hist_s"].sum()) == d["n_s"]
    sel = ~np.isnan(d["key"])
    tkey = d["key"][sel]
    # source keys: the oracle's exact run on the swapped problem exposes them as ITS target keys
    sw_mask = 1 - mask
    _, st2, d2 = orc.transfer(s, c, sw_mask, debug=True)
    assert st2 == 0 and np.array_equal(d2["t_comp"], d["s_comp"])
    skey = d2["key"][~np.isnan(d2["key"])]
    m, code, ht, hs = _numpy_binned(tkey, skey, B)
    assert np.array_equal(ht, d["hist_t"]) and np.array_equal(hs, d["hist_s"])
    assert np.array_equal(code, d["code"][sel])
    assert np.array_equal(m, d["matched"][sel])
    assert np.allclose(d["range"], [tkey.min(), tkey.max(), skey.min(), skey.max()], rtol=0, atol=0)
    # monotone: a larger key never gets a smaller matched value
    o = np.argsort(tkey, kind="stable")
    assert np.all(np.diff(m[o]) >= 0)


@pytest.mark.parametrize("kind", ["U", "N"])
def test_binned_is_close_to_the_exact_path(kind):
    H, W = 270, 480
    c, s = synth.synth_image(H, W, 11, kind), synth.synth_image(200, 300, 12, kind)
    mask = synth.ellipse_mask(H, W, 0.4, 4)
    exact, _, _ = orc.transfer(c, s, mask)
    for B, max_frac in ((256, 0.06), (1024, 0.03)):
        out, st, _ = orc.transfer_binned(c, s, mask, B)
        diff = np.abs(out.astype(int) - exact.astype(int))
        # <= 1 LSB except at isolated pixels in the extreme tails of the key range (measured max 3)
        assert st == 0 and diff.max() <= 3 and (diff.max(-1) > 1).mean() <= 1e-4, (B, diff.max())
        assert (diff.max(-1) > 0).mean() <= max_frac, (B, (diff.max(-1) > 0).mean())
    assert np.array_equal(out[mask == 1], exact[mask == 1])


def test_binned_degenerate_regions():
    H, W = 40, 50
    s = synth.synth_image(H, W, 1, "N")
    mask = synth.ellipse_mask(H, W, 0.5, 3)
    flat = np.empty((H, W, 3), np.uint8)
    flat[...] = (200, 31, 7)
    out, st, d = orc.transfer_binned(flat, s, mask, 256)                   # flat target: hi == lo
    assert st == 0 and d["hist_t"][0] == d["n_t"] and len(np.unique(out[mask == 0].reshape(-1, 3), axis=0)) == 1
    out, st, d = orc.transfer_binned(synth.synth_image(H, W, 5, "U"), flat, mask, 256)   # flat source
    assert st == 0 and d["hist_s"][0] == d["n_s"]
    for m_, want in ((np.ones((H, W), np.uint8), 1), (np.zeros((H, W), np.uint8), 2)):   # empty regions
        assert orc.transfer_binned(flat, s, m_, 64)[1] == want


# ---- GPU -------------------------------------------------------------------------------------------------------
def _run_binned_debug(c, s, mask, B, mode=0, label=0):
    import torch

    from applied_image_processing_b200 import _cabi
    from applied_image_processing_b200 import localized_style_transfer as lst

    H, W = c.shape[:2]
    plan = lst.Plan(torch.device("cuda", 0), H, W, s.shape[0], s.shape[1], slots=1, flags=_cabi.FLAG_DEBUG)
    dc, ds = torch.from_numpy(c).cuda(), torch.from_numpy(s).cuda()
    dm = torch.from_numpy(mask).cuda() if mask is not None else None
    out = torch.zeros_like(dc) if mode != 1 else dc.clone()
    _cabi.check(_cabi.lib().rct_transfer_binned_async(
        plan.handle, dc.data_ptr(), c.size, ds.data_ptr(), s.size, dm.data_ptr() if dm is not None else None,
        mask.size if mask is not None else 0, out.data_ptr(), c.size, 1, mode, label, B, torch.cuda.current_stream().cuda_stream))
    torch.cuda.synchronize()
    return out.cpu().numpy(), plan


@pytest.mark.gpu
@pytest.mark.parametrize("H,W,sh,sw,kind,B", [
    (270, 480, 270, 480, "U", 1024), (270, 480, 200, 300, "N", 256), (97, 131, 97, 131, "U", 4096),
    (540, 960, 540, 960, "N", 1024), (33, 47, 40, 60, "U", 2), (1080, 1920, 512, 910, "N", 1024),
])
def test_cuda_binned_bit_exact_vs_its_oracle(H, W, sh, sw, kind, B):
    c, s = synth.synth_image(H, W, 60, kind), synth.synth_image(sh, sw, 61, kind)
    mask = synth.ellipse_mask(H, W, 0.4, 4)
    ref, st, d = orc.transfer_binned(c, s, mask, B, debug=True)
    out, plan = _run_binned_debug(c, s, mask, B)
    info = plan.info()
    assert info["status"] == st == 0 and info["n_target"] == d["n_t"] and info["n_source"] == d["n_s"]
    for k in ("t_mean", "t_comp", "s_mean", "s_comp"):
        assert np.array_equal(info[k], d[k]), k
    g = plan.debug_binned(B)
    assert np.array_equal(g["hist_t"], d["hist_t"]) and np.array_equal(g["hist_s"], d["hist_s"])   # histograms
    assert np.array_equal(g["range"], d["range"])                                                   # key ranges
    col = c[..., 0].astype(np.uint32) | (c[..., 1].astype(np.uint32) << 8) | (c[..., 2].astype(np.uint32) << 16)
    sel = ~np.isnan(d["key"])
    order = np.argsort(g["colour"])
    pos = order[np.searchsorted(g["colour"][order], col[sel])]
    assert np.array_equal(g["colour"][pos], col[sel])
    assert np.array_equal(g["code"][pos], d["code"][sel]), "CDF index pairs differ from the oracle"
    assert np.array_equal(g["matched"][pos], d["matched"][sel])
    assert np.array_equal(out, ref), f"{(out != ref).sum()} bytes differ from the binned oracle"


@pytest.mark.gpu
def test_cuda_binned_public_api_batches_modes_and_empty_regions():
    import torch

    import applied_image_processing_b200 as aip
    from applied_image_processing_b200 import _cabi

    n, H, W, B = 5, 120, 160, 1024
    cs = np.stack([synth.synth_image(H, W, 100 + i, "UN"[i % 2]) for i in range(n)])
    ss = np.stack([synth.synth_image(H, W, 200 + i, "NU"[i % 2]) for i in range(n)])
    ms = np.stack([synth.ellipse_mask(H, W, 0.2 + 0.1 * i, i) for i in range(n)])
    ms[3] = 1                                                                # a frame without target pixels
    refs = np.stack([orc.transfer_binned(cs[i], ss[i], ms[i], B)[0] for i in range(n)])
    assert np.array_equal(aip.regional_color_transfer_binned(cs, ss, ms, B, slots=2), refs)
    t = aip.regional_color_transfer_binned(torch.from_numpy(cs).cuda(), torch.from_numpy(ss).cuda(), torch.from_numpy(ms).cuda(), B)
    assert t.is_cuda and np.array_equal(t.cpu().numpy(), refs)
    labels = synth.voronoi_labels(H, W, 4, 6)                                # LABEL and ALL modes
    ref = cs[0].copy()
    orc.transfer_binned(cs[0], ss[0], labels, 256, mode=orc.MODE_LABEL, label=2, out=ref)
    got = aip.regional_color_transfer_binned(cs[0], ss[0], labels, 256, mode=_cabi.MODE_LABEL, label=2)
    assert np.array_equal(got, ref)
    ref, _, _ = orc.transfer_binned(cs[1], ss[1], None, 256, mode=orc.MODE_ALL)
    assert np.array_equal(aip.regional_color_transfer_binned(cs[1], ss[1], None, 256, mode=_cabi.MODE_ALL), ref)


@pytest.mark.gpu
@pytest.mark.parametrize("H,W,sh,sw", [(150, 200, 120, 170), (1080, 1920, 1080, 1920)])
def test_cuda_binned_row_sharded_phases_equal_the_unsharded_call(H, W, sh, sw):
    """Three uneven bands on one GPU; the exchange buffers are combined as the all-reduces would (SUM / MAX /
    SUM).  Every band's output rows must equal the unsharded binned result, which equals the oracle."""
    import torch

    from applied_image_processing_b200 import _cabi
    from applied_image_processing_b200 import localized_style_transfer as lst

    B = 1024
    c, s = synth.synth_image(H, W, 80, "U"), synth.synth_image(sh, sw, 81, "N")
    mask = synth.ellipse_mask(H, W, 0.45, 2)
    ref, st0, _ = orc.transfer_binned(c, s, mask, B)
    assert st0 == 0
    L = _cabi.lib()
    st = torch.cuda.current_stream().cuda_stream
    dc, ds, dm = (torch.from_numpy(a).cuda() for a in (c, s, mask))
    out = torch.zeros_like(dc)
    plans = [lst.Plan(torch.device("cuda", 0), H, W, sh, sw) for _ in range(3)]
    bounds = [0, H // 4 + 1, H // 4 + 1 + H // 2, H]

    def view(p, which):
        n, ts = {0: (56, "<i8"), 1: (4, "<f8"), 2: (2 * B, "<i4")}[which]

        class _Buf:
            __cuda_array_interface__ = {"shape": (n,), "typestr": ts, "data": (L.rct_plan_xchg_ptr(p.handle), False), "version": 3}
        return torch.as_tensor(_Buf(), device="cuda")

    for phase in range(4):
        for r, p in enumerate(plans):
            _cabi.check(L.rct_shard_binned_phase_async(p.handle, phase, dc.data_ptr(), ds.data_ptr(), dm.data_ptr(), out.data_ptr(),
                                                       0, 0, bounds[r], bounds[r + 1] - bounds[r], B, st))
        if phase < 3:
            bufs = [view(p, phase) for p in plans]
            tot = torch.stack(bufs).max(0).values if phase == 1 else torch.stack(bufs).sum(0)
            for b in bufs:
                b.copy_(tot)
    torch.cuda.synchronize()
    assert np.array_equal(out.cpu().numpy(), ref), int((out.cpu().numpy() != ref).sum())


def _binned_shard_worker(rank, world, port, H, W, q):
    import os

    import torch
    import torch.distributed as dist

    from applied_image_processing_b200 import sharded

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(0)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        c, s = synth.synth_image(H, W, 90, "U"), synth.synth_image(H, W, 91, "N")
        mask = synth.ellipse_mask(H, W, 0.45, 2)
        out, (r0, n) = sharded.regional_color_transfer_sharded(c, s, mask, bins=256)
        q.put((rank, r0, n, out))
    finally:
        dist.destroy_process_group()


@pytest.mark.gpu
def test_cuda_binned_sharded_driver_two_ranks_one_gpu():
    import socket

    import torch.multiprocessing as mp

    H, W, world = 151, 203, 2
    with socket.socket() as sk:
        sk.bind(("127.0.0.1", 0))
        port = sk.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_binned_shard_worker, args=(r, world, port, H, W, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = [q.get(timeout=240) for _ in range(world)]
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    c, s = synth.synth_image(H, W, 90, "U"), synth.synth_image(H, W, 91, "N")
    ref, st, _ = orc.transfer_binned(c, s, synth.ellipse_mask(H, W, 0.45, 2), 256)
    full = np.zeros_like(ref)
    for rank, r0, n, out in got:
        full[r0:r0 + n] = out
    assert st == 0 and np.array_equal(full, ref)


@pytest.mark.gpu
def test_cuda_binned_4k_vs_oracle_and_distance_from_exact():
    import applied_image_processing_b200 as aip
    from tests import fullsize

    c, s, mask = fullsize.inputs(fullsize.CASES["fs_4k_U"])
    ref, st, _ = orc.transfer_binned(c, s, mask, 1024)
    out = aip.regional_color_transfer_binned(c, s, mask, 1024)
    assert st == 0 and np.array_equal(out, ref), int((out != ref).sum())
    exact = aip.regional_color_transfer(c, s, mask)
    diff = np.abs(out.astype(int) - exact.astype(int))
    assert diff.max() <= 3 and (diff.max(-1) > 1).mean() <= 1e-4 and (diff.max(-1) > 0).mean() < 0.03
