#!/usr/bin/env python3
"""Raw pinned-memory copy ceiling of the box for the e2e leg of bench.py: every rank copies the 4K workload's
bytes (464 MB host->device + 199 MB device->host per step) on two streams at once, all ranks concurrently.

    python tools/pcie_bw.py                                   # one GPU
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P tools/pcie_bw.py

Rank 0 prints one JSON line: per-step time (max over ranks), aggregate GB/s, per-rank GB/s, for H2D alone, D2H alone
and both directions busy; `--affinity` pins each rank's host thread to the cores of its GPU's NUMA node (from
/sys/bus/pci/devices/<bdf>/local_cpulist) before allocating the pinned buffers (first touch), `--wc` allocates the
input buffer write-combined (cudaHostAllocWriteCombined)."""
import argparse
import json
import os
import time

import torch
import torch.distributed as dist

H2D_BYTES, D2H_BYTES = 464486400, 199065600


def gpu_cpulist(index):
    try:
        import pynvml

        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(index)
        bdf = pynvml.nvmlDeviceGetPciInfo(h).busId
        bdf = (bdf.decode() if isinstance(bdf, bytes) else bdf).lower()
        if len(bdf.split(":")[0]) == 8:
            bdf = bdf[4:]
        with open(f"/sys/bus/pci/devices/{bdf}/local_cpulist") as fh:
            txt = fh.read().strip()
        with open(f"/sys/bus/pci/devices/{bdf}/numa_node") as fh:
            node = int(fh.read().strip())
        cpus = set()
        for part in txt.split(","):
            a, _, b = part.partition("-")
            cpus.update(range(int(a), int(b or a) + 1))
        return sorted(cpus), node, bdf
    except Exception as e:  # noqa: BLE001
        return None, None, str(e)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--reps", type=int, default=10)
    ap.add_argument("--affinity", action="store_true")
    ap.add_argument("--wc", action="store_true")
    a = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    cpus, node, bdf = gpu_cpulist(local)
    if a.affinity and cpus:
        os.sched_setaffinity(0, cpus)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    if a.wc:
        import ctypes

        rt = ctypes.CDLL("libcudart.so")
        p = ctypes.c_void_p()
        assert rt.cudaHostAlloc(ctypes.byref(p), ctypes.c_size_t(H2D_BYTES), ctypes.c_uint(4)) == 0  # WriteCombined
        buf = (ctypes.c_ubyte * H2D_BYTES).from_address(p.value)
        h_in = torch.frombuffer(buf, dtype=torch.uint8)
    else:
        h_in = torch.empty(H2D_BYTES, dtype=torch.uint8).pin_memory()
    h_in.fill_(1)  # first touch on this rank's node
    h_out = torch.empty(D2H_BYTES, dtype=torch.uint8).pin_memory()
    h_out.fill_(0)
    d_in = torch.empty(H2D_BYTES, dtype=torch.uint8, device="cuda")
    d_out = torch.empty(D2H_BYTES, dtype=torch.uint8, device="cuda")
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()

    def h2d():
        with torch.cuda.stream(s1):
            d_in.copy_(h_in, non_blocking=True)

    def d2h():
        with torch.cuda.stream(s2):
            h_out.copy_(d_out, non_blocking=True)

    def both():
        h2d()
        d2h()

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    out = {"ranks": world, "affinity": bool(a.affinity and cpus), "write_combined": a.wc, "gpu0_numa_node": node}
    for name, fn, nb in (("h2d", h2d, H2D_BYTES), ("d2h", d2h, D2H_BYTES), ("both", both, H2D_BYTES + D2H_BYTES)):
        fn()
        barrier()
        t0 = time.perf_counter()
        for _ in range(a.reps):
            fn()
        torch.cuda.synchronize()
        t = (time.perf_counter() - t0) / a.reps
        if world > 1:
            tt = torch.tensor([t], dtype=torch.float64, device="cuda")
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            t = float(tt.item())
        barrier()
        out[name] = {"ms_per_step": round(1e3 * t, 3), "aggregate_GBps": round(world * nb / t / 1e9, 1),
                     "per_rank_GBps": round(nb / t / 1e9, 1)}
    if rank == 0:
        # the 4K workload moves 66.36 MP per rank per step
        out["e2e_ceiling_MPps"] = round(world * 66.3552 / (out["both"]["ms_per_step"] * 1e-3), 0)
        print(json.dumps(out), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
