"""The row-sharded drivers over REAL NCCL (two GPUs, one process each): exact mode (all-reduce of the colour cubes) and
binned mode (three tiny all-reduces) against the oracles.  Skipped on boxes with fewer than two GPUs (the gloo tests in
test_gpu_parity.py / test_binned.py cover the same drivers with two ranks on one GPU through a host copy)."""
import socket

import numpy as np
import pytest
import torch

from oracle import oracle as orc
from tests import synth

pytestmark = pytest.mark.gpu


def _worker(rank, world, port, H, W, sh, sw, bins, q):
    import os

    import torch.distributed as dist

    from applied_image_processing_b200 import sharded

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        c, s = synth.synth_image(H, W, 90, "U"), synth.synth_image(sh, sw, 91, "N")
        mask = synth.ellipse_mask(H, W, 0.45, 2)
        out, (r0, n) = sharded.regional_color_transfer_sharded(c, s, mask, bins=bins)
        q.put((rank, r0, n, out))
    finally:
        dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs for NCCL")
@pytest.mark.parametrize("bins", [0, 1024])
@pytest.mark.parametrize("H,W,sh,sw", [(151, 203, 120, 170), (1080, 1920, 1080, 1920)])
def test_sharded_drivers_over_nccl(H, W, sh, sw, bins):
    import torch.multiprocessing as mp

    world = 2
    with socket.socket() as sk:
        sk.bind(("127.0.0.1", 0))
        port = sk.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, H, W, sh, sw, bins, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = [q.get(timeout=300) for _ in range(world)]
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    c, s = synth.synth_image(H, W, 90, "U"), synth.synth_image(sh, sw, 91, "N")
    mask = synth.ellipse_mask(H, W, 0.45, 2)
    ref, st, _ = orc.transfer_binned(c, s, mask, bins) if bins else orc.transfer(c, s, mask)
    assert st == 0
    full = np.zeros_like(ref)
    for rank, r0, n, out in got:
        full[r0:r0 + n] = out
    assert np.array_equal(full, ref), int((full != ref).sum())
