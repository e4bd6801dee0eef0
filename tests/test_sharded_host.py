"""Host-side logic of the row-sharded path (applied_image_processing_b200/sharded.py) on CPU:
band partition, the stylised rows a band needs, and -- with two ``gloo`` ranks -- that the
all-reduced per-band colour histograms equal the full-image histograms (the property the
sharded protocol rests on).  The histograms here come from the oracle."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from applied_image_processing_b200 import sharded
from oracle import oracle as orc
from tests import synth


@pytest.mark.parametrize("H,world", [(10, 1), (10, 3), (7, 8), (8640, 8), (2160, 4), (5, 2)])
def test_row_band_is_a_partition(H, world):
    nxt = 0
    sizes = []
    for r in range(world):
        r0, n = sharded.row_band(H, r, world)
        assert r0 == nxt and n >= 0
        nxt += n
        sizes.append(n)
    assert nxt == H and max(sizes) - min(sizes) <= 1


@pytest.mark.parametrize("sh,H,world", [(512, 485, 2), (200, 270, 3), (1080, 2160, 4), (300, 300, 2), (900, 128, 8)])
def test_stylised_rows_cover_the_band(sh, H, world):
    tab = sharded.resample_table(sh, H)
    assert np.array_equal(tab, orc.resample_table(sh, H))
    for r in range(world):
        r0, n = sharded.row_band(H, r, world)
        s0, s1 = sharded.stylised_rows_needed(sh, H, r0, n)
        if n == 0:
            assert (s0, s1) == (0, 0)
            continue
        band = tab[r0:r0 + n] if sh != H else np.arange(r0, r0 + n)
        assert s0 == band.min() and s1 == band.max() + 1 and 0 <= s0 < s1 <= sh


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, H, W, sh, sw, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        c, s = synth.synth_image(H, W, 11, "U"), synth.synth_image(sh, sw, 12, "N")
        mask = synth.ellipse_mask(H, W, 0.45, 3)
        r0, n = sharded.row_band(H, rank, world)
        s0, s1 = sharded.stylised_rows_needed(sh, H, r0, n)
        xt, yt = sharded.resample_table(sw, W), sharded.resample_table(sh, H)
        # the band sees only stylised rows [s0, s1): index them relative to s0
        sty_band = s[s0:s1][yt[r0:r0 + n] - s0][:, xt]
        _, _, d = orc.transfer(c[r0:r0 + n], sty_band, mask[r0:r0 + n], debug=True)
        h = torch.from_numpy(np.concatenate([d["hist_t"], d["hist_s"]]).view(np.int32).copy())
        dist.all_reduce(h, op=dist.ReduceOp.SUM)
        if rank == 0:
            _, _, full = orc.transfer(c, s, mask, debug=True)
            want = np.concatenate([full["hist_t"], full["hist_s"]])
            q.put(bool(np.array_equal(h.numpy().view(np.uint32), want)) and int(want.sum()) > 0)
    finally:
        dist.destroy_process_group()


def test_gloo_allreduce_of_band_histograms_equals_full_histogram():
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, 61, 83, 40, 57, q)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    assert q.get(timeout=5) is True
