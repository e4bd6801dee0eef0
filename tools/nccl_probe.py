#!/usr/bin/env python3
"""Where does an in-pipeline all-reduce lose its time?  (The row-sharded 16K workload measured 1.1 ms per all-reduce at 8
ranks inside the step, whatever the message size, against 0.41 ms for 128 MiB back to back.)  Under torchrun: tiny and
large all-reduces (a) back to back, (b) each behind ~1 ms of compute on the same stream, (c) as (b) with a host
synchronisation per iteration (what a per-step phase timer does).  Rank 0 prints one JSON line: ms per iteration minus
the compute alone."""
import json
import os
import time

import torch
import torch.distributed as dist

world, rank, local = int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
a = torch.randn(4096, 4096, device=dev)


def compute():
    for _ in range(2):
        torch.mm(a, a)


def timed(fn, iters=30):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(iters):
        fn()
    torch.cuda.synchronize()
    t = (time.perf_counter() - t0) / iters * 1e3
    tt = torch.tensor([t], dtype=torch.float64, device=dev)
    dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    return float(tt.item())


out = {"ranks": world}
t_compute = timed(compute)
out["compute_ms"] = round(t_compute, 4)
for name, n in (("tiny_56xi64", 56), ("hist_2x1024xi32", 2048), ("cube_128MiB", 32 << 20)):
    buf = torch.zeros(n, dtype=torch.int64 if n == 56 else torch.int32, device=dev)
    r = {}
    r["back_to_back_ms"] = round(timed(lambda: dist.all_reduce(buf)), 4)

    def behind():
        compute()
        dist.all_reduce(buf)

    def behind_sync():
        compute()
        dist.all_reduce(buf)
        torch.cuda.synchronize()

    r["behind_compute_ms"] = round(timed(behind) - t_compute, 4)
    r["behind_compute_host_sync_ms"] = round(timed(behind_sync) - t_compute, 4)
    out[name] = r
if rank == 0:
    print(json.dumps(out), flush=True)
dist.destroy_process_group()
