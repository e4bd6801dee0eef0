#!/usr/bin/env python3
"""bench.py -- megapixels/s of the regional colour transfer (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

One "step" = one pass of the hot path (LST:222-238 + color_transfer_foreground, LST:128-168)
over a batch of `--frames` distinct synthetic 3840x2160 content/stylised pairs with binary
fg/bg masks (BASELINE.json configs[1]); the batch is larger than L2, so no flush is needed
between steps.  N>1: every rank processes its own batch (frames are independent units, no
data-path collective) -> weak scaling; `value` = all ranks' pixels / max-over-ranks time.

Printed keys (one JSON line, rank 0): see the task contract; `roofline` describes the
dominant per-pixel kernel, `cpu_baseline` the C port of the reference (oracle/) timed on this
box's host cores on a bounded sample, `e2e` the same metric through the public numpy API
(pinned host buffers in, host buffer out, copies inside the timed region).
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

METRIC = "megapixels/sec region colour transfer"
UNIT = "MP/s"
H4K, W4K = 2160, 3840
ALG_BYTES_PER_PIXEL = {"hist": 7.0, "apply": 10.0, "pipeline": 10.0}  # SURVEY.md 8(d), DESIGN.md
# kernels that stream the pixels and their algorithmic bytes per pixel: content 3 + stylised 3 + mask 1
# read (+ combined 3 written by the apply pass)
PIXEL_KERNELS = {"k_front_count": 7.0, "k_front_part": 7.0, "k_hist": 7.0, "k_apply": 10.0}  # "k_apply" = the timing slot of the apply pass


TRAFFIC_FILE = "profiles/r02_traffic.json"


def load_traffic():
    p = os.path.join(ROOT, TRAFFIC_FILE)
    if not os.path.exists(p):
        p = os.path.join(ROOT, "profiles", "r01_traffic.json")
    try:
        return json.load(open(p))
    except Exception:
        return {}


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="4k", choices=["4k", "1080p", "labels", "16k"],
                    help="4k: BASELINE configs[1] (default, the quoted metric); 1080p: configs[2] frame batch; "
                         "labels: configs[4] 8-label 1080p batch; 16k: configs[3] one image row-sharded over the ranks")
    ap.add_argument("--frames", type=int, default=0, help="frames per step and per rank (0 = workload default)")
    ap.add_argument("--slots", type=int, default=0, help="frames per launch group (0 = min(frames, 8))")
    ap.add_argument("--kind", default="U", choices=["U", "N"], help="synthetic image family (SURVEY 8d)")
    ap.add_argument("--height", type=int, default=0)
    ap.add_argument("--width", type=int, default=0)
    ap.add_argument("--cpu-frames", type=int, default=0, help="frames in the CPU sample (0 = auto)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--labels-fused", type=int, default=1, choices=[0, 1],
                    help="labels workload: 1 all labels of a frame group in one launch group (rct_transfer_labels_async), "
                         "0 one pass per label (A/B runs)")
    ap.add_argument("--lanes", type=int, default=0, choices=[0, 1, 2, 3, 4],
                    help="launch groups in flight at once: the step's frames are split over this many plans, each on its own "
                         "stream (the solve kernels are latency-bound: concurrent groups overlap their stalls)")
    ap.add_argument("--keytab", type=int, default=1, choices=[0, 1],
                    help="RCT_OPT_KEYTAB: 1 key table in front of the colour LUT (default), 0 colour LUT only (A/B runs)")
    ap.add_argument("--lut", type=int, default=0, choices=[0, 1],
                    help="RCT_OPT_LUT: 0 uint32 RGB per colour (default), 1 uint16 run index + palette (A/B runs)")
    ap.add_argument("--sty-short", type=int, default=0,
                    help="stylised frames at this short-side size (512 = what AdaIN's Resize(512) hands the reference, "
                         "LST:223-230), NEAREST-resampled inside the kernels; 0 = same size as the content")
    ap.add_argument("--files", type=int, default=1, choices=[0, 1],
                    help="4k / 1080p workloads on one GPU: also report e2e_files, the same transfer through the reference's real "
                         "boundary (JPEG files in, JPEG file out; rct_transfer_files_host) beside PIL decode + C port + PIL encode")
    ap.add_argument("--bins", type=int, default=0,
                    help="0 (default): the reference's exact CDF matching (the parity path); > 0: the BINNED approximation with "
                         "this many histogram bins (BASELINE.json configs[4] '1024-bin histograms'; reported separately, never "
                         "the headline).  Workloads 4k / 1080p / 16k")
    ap.add_argument("--e2e-slots", type=int, default=0, help="frames per copy/compute group of the host entry point (0 = its default)")
    ap.add_argument("--stage-sets", type=int, default=0, help="RCT_OPT_STAGE_SETS of the host entry point (0 = default)")
    ap.add_argument("--front", type=int, default=1, choices=[0, 1],
                    help="1: colour-partition front end (default); 0: colour-cube histograms (A/B runs)")
    ap.add_argument("--match", type=int, default=1, choices=[0, 1], help="k_match traversal (RCT_OPT_MATCH), A/B runs")
    ap.add_argument("--apply", type=int, default=1, choices=[0, 1, 2, 3, 4, 5, 6, 7, 8],
                    help="apply kernel (RCT_OPT_APPLY): 1 per-warp bulk-async pipeline, 256-px tiles (default), 2 same with "
                         "cp.async, 4 per-warp bulk-async 512-px tiles, 3 CTA-wide bulk-async ring, 6 three-stage variant of 1, 7 = 1 with one bulk store per tile, 0 register-staged k_apply")
    a = ap.parse_args()
    dflt = {"4k": (H4K, W4K, 16), "1080p": (1080, 1920, 32), "labels": (1080, 1920, 16), "16k": (8640, 15360, 1)}[a.workload]
    a.height = a.height or dflt[0]
    a.width = a.width or dflt[1]
    a.frames = a.frames or dflt[2]
    a.slots = a.slots or min(a.frames, 8)
    if a.lanes == 0:
        a.lanes = 2 if (a.frames >= 2 * a.slots and a.workload in ("4k", "1080p") and not a.bins) else 1
    return a


def workload_name(a):
    if a.bins:
        a_bins, a.bins = a.bins, 0
        name = workload_name(a)
        a.bins = a_bins
        return name + f"; BINNED APPROXIMATION, {a_bins}-bin key histograms + CDF inverse lookup instead of the exact match_cdf"
    if a.workload == "16k":
        return (f"one {a.width}x{a.height} synthetic content/stylised pair, binary fg/bg ellipse mask (40% fg), image "
                f"family {a.kind}, row-sharded over the ranks with an all-reduce of the colour histograms "
                f"(BASELINE.json configs[3])")
    if a.workload == "labels":
        return (f"{a.width}x{a.height} synthetic frames with 8-label Voronoi masks, per-label exact transfer, image "
                f"family {a.kind} (BASELINE.json configs[4], exact-rank mode); {a.frames} frames per step per GPU")
    cfg = "configs[1]" if a.workload == "4k" else "configs[2]"
    rs = ""
    if a.sty_short:
        sh, sw = sty_size(a)
        rs = f", stylised at {sw}x{sh} (short side {a.sty_short}) NEAREST-resampled in-kernel"
    return (f"{a.width}x{a.height} synthetic content/stylised pair{rs}, binary fg/bg ellipse mask (40% fg), "
            f"image family {a.kind} (BASELINE.json {cfg}); {a.frames} distinct pairs per step per GPU")


def config_dict(a, world):
    """`config` of the JSON line: a function of the arguments only, so that both arms (ours / reference)
    print the same dict for the same command line."""
    if a.workload == "16k":
        rows = (a.height + world - 1) // world
        read_mb = rows * a.width * 7 / 1e6
        return {"workload": workload_name(a), "rows_per_rank": rows,
                "l2": f"inputs larger than L2: {read_mb:.0f} MB read per rank per step",
                "parallelism": f"rows sharded over {world} GPU(s); exchange of the per-region statistics over NCCL"}
    sh, sw = sty_size(a)
    read_mb = a.frames * (a.height * a.width * 4 + sh * sw * 3) / 1e6
    return {"workload": workload_name(a), "frames_per_step_per_gpu": a.frames, "frames_per_launch_group": a.slots,
            "concurrent_launch_groups": a.lanes,
            "l2": f"inputs larger than L2: {read_mb:.0f} MB read per step",
            "parallelism": f"frames sharded over {world} GPU(s), no collective"}


def sty_size(a):
    """(sh, sw) of the stylised frames: torchvision Resize(short) keeps the aspect ratio (test.py:16-24)."""
    if not a.sty_short or a.workload in ("16k", "labels"):
        return a.height, a.width
    short = min(a.height, a.width)
    return (a.sty_short, int(a.sty_short * a.width / short)) if a.height <= a.width else (
        int(a.sty_short * a.height / short), a.sty_short)


def make_frames(a, n, seed0):
    from tests.synth import ellipse_mask, synth_image

    sh, sw = sty_size(a)
    cs = np.empty((n, a.height, a.width, 3), np.uint8)
    ss = np.empty((n, sh, sw, 3), np.uint8)
    ms = np.empty((n, a.height, a.width), np.uint8)
    for i in range(n):
        cs[i] = synth_image(a.height, a.width, seed0 + 2 * i, a.kind)
        ss[i] = synth_image(sh, sw, seed0 + 2 * i + 1, a.kind)
        if a.workload == "labels":
            from tests.synth import voronoi_labels

            ms[i] = voronoi_labels(a.height, a.width, 8, seed0 + i)
        else:
            ms[i] = ellipse_mask(a.height, a.width, 0.4, seed0 + i)
    return cs, ss, ms


# ---- CPU leg: the C port of the reference (oracle/), all host threads ------------------------------
def host_threads():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def cpu_pass(frames, threads, labels=False, keep=False, bins=0):
    """Run every frame through the oracle, `threads` at a time (ctypes releases the GIL).
    keep=True: returns (seconds, outputs) so the GPU arm can assert parity on the same frames."""
    from concurrent.futures import ThreadPoolExecutor

    from oracle import oracle as orc

    orc.lib()
    cs, ss, ms = frames

    def one(i):
        if labels:
            out = cs[i].copy()
            for lab in range(8):
                orc.transfer(cs[i], ss[i], ms[i], mode=orc.MODE_LABEL, label=lab, out=out)
            return out if keep else None
        out = (orc.transfer_binned(cs[i], ss[i], ms[i], bins) if bins else orc.transfer(cs[i], ss[i], ms[i]))[0]
        return out if keep else None

    t0 = time.perf_counter()
    with ThreadPoolExecutor(max_workers=threads) as ex:
        outs = list(ex.map(one, range(cs.shape[0])))
    t = time.perf_counter() - t0
    return (t, outs) if keep else t


def cpu_sample_size(a):
    """(frames, threads) of the CPU arm: one frame per thread on EVERY core this process may use (~6.7 s per 4K frame per
    core), fewer only where the host's free memory would not hold one oracle working set (~160 bytes per pixel) per thread."""
    threads = max(1, host_threads())
    try:
        avail = next(int(l.split()[1]) * 1024 for l in open("/proc/meminfo") if l.startswith("MemAvailable:"))
        threads = max(1, min(threads, int(0.5 * avail / (160 * a.height * a.width))))
    except Exception:  # noqa: BLE001
        pass
    n = a.cpu_frames or threads
    return n, min(threads, n)


def run_reference_arm(a):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return  # rank 0 alone runs the CPU arm
    n, threads = cpu_sample_size(a)
    full_h = a.height
    if a.workload == "16k":  # one 16K image takes minutes on one core: time a 1/16 band of it per thread
        a.height = a.height // 16
    frames = make_frames(a, n, 1000)
    labels = a.workload == "labels"
    for _ in range(a.warmup):
        cpu_pass(frames, threads, labels, bins=a.bins)
    times = [cpu_pass(frames, threads, labels, bins=a.bins) for _ in range(a.steps)]
    mp = n * a.height * a.width / 1e6
    band_note = f" (each a {a.height}-row band of the {full_h}-row image)" if a.workload == "16k" else ""
    rows_used = a.height
    a.height = full_h
    t = sum(times) / len(times)
    val = mp / t
    sample = f"{n} frame(s) of the workload per step{band_note}, {threads} thread(s), one frame per thread"
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": a.gpus, "steps": a.steps,
        "warmup": a.warmup, "ms_per_step": 1e3 * t, "higher_is_better": True,
        "scaling": "strong" if a.workload == "16k" else "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": config_dict(a, a.gpus),
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": threads, "host_cores_available": host_threads(), "kind": "port",
                         "sample": sample},
        # `config` is the repo arm's (the workload both arms are quoted on); what THIS arm ran per step is here
        "reference_sample": {"frames_per_step": n, "threads": threads, "rows_per_frame": rows_used,
                             "note": "MP/s is per pixel processed, so the sample size does not enter the ratio"},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def files_leg(a, aip, cs, ss, ms, H, W):
    """The step's frames as JPEG files (Pillow defaults) through regional_color_transfer_files (device decode ->
    transfer -> device encode; file bytes and the mask cross PCIe, not pixels), and -- on a sample, one frame per host
    thread -- the same thing on the CPU: PIL decode + C port + PIL encode.  Lossy inputs: parity here is against the
    CPU chain on the SAME files, asserted byte for byte on the sample."""
    import io
    from concurrent.futures import ThreadPoolExecutor

    from PIL import Image

    from oracle import oracle as orc

    def enc(arr):
        buf = io.BytesIO()
        Image.fromarray(arr).save(buf, "JPEG")
        return buf.getvalue()

    n = min(cs.shape[0], 8)
    cf, sf = [enc(cs[i]) for i in range(n)], [enc(ss[i]) for i in range(n)]
    outs = [aip.regional_color_transfer_files(cf[i], sf[i], ms[i]) for i in range(min(2, n))]  # warm-up
    t0 = time.perf_counter()
    reps = 2
    for _ in range(reps):
        outs = [aip.regional_color_transfer_files(cf[i], sf[i], ms[i]) for i in range(n)]
    t = (time.perf_counter() - t0) / reps
    nc, threads = cpu_sample_size(a)
    nc = min(nc, n)

    def cpu_one(i):
        c = np.array(Image.open(io.BytesIO(cf[i])).convert("RGB"))
        s = np.array(Image.open(io.BytesIO(sf[i])).convert("RGB"))
        buf = io.BytesIO()
        Image.fromarray(orc.transfer(c, s, ms[i])[0]).save(buf, "JPEG")
        return buf.getvalue()

    t1 = time.perf_counter()
    with ThreadPoolExecutor(max_workers=min(threads, nc)) as ex:
        refs = list(ex.map(cpu_one, range(nc)))
    t_cpu = time.perf_counter() - t1
    for i in range(nc):
        assert outs[i] == refs[i], f"e2e_files: frame {i} differs from PIL decode + oracle + PIL encode"
    file_bytes = sum(len(x) for x in cf) + sum(len(x) for x in sf)
    return {"value": n * H * W / 1e6 / t, "unit": UNIT, "ms_per_step": 1e3 * t, "frames": n,
            "h2d_bytes_per_step": int(file_bytes + ms.size), "d2h_bytes_per_step": int(sum(len(x) for x in outs)),
            "input": "the step's frames as JPEG files (Pillow defaults: quality 75, 4:2:0)", "parity_checked_files": nc,
            "cpu_files_reference": {"value": nc * H * W / 1e6 / t_cpu, "unit": UNIT, "cores": min(threads, nc),
                                    "what": "PIL decode x2 + C port + PIL encode, one frame per thread"}}


# ---- clocks during the timed region -----------------------------------------------------------------
class ClockSampler:
    REASONS = {
        0x1: "gpu_idle", 0x2: "applications_clocks_setting", 0x4: "sw_power_cap", 0x8: "hw_slowdown",
        0x10: "sync_boost", 0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown",
        0x80: "hw_power_brake_slowdown", 0x100: "display_clock_setting",
    }

    def __init__(self, device_index):
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._thr = None
        try:
            import pynvml
            import torch

            pynvml.nvmlInit()
            uuid = str(torch.cuda.get_device_properties(device_index).uuid)
            try:
                self.h = pynvml.nvmlDeviceGetHandleByUUID(("GPU-" + uuid).encode())
            except Exception:
                self.h = pynvml.nvmlDeviceGetHandleByIndex(device_index)
            self.nv = pynvml
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def _loop(self):
        nv = self.nv
        while not self._stop.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                try:
                    r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in self.REASONS.items():
                    if r & bit and name != "gpu_idle":
                        self.reasons.add(name)
            except Exception:
                pass
            self._stop.wait(0.05)

    def start(self):
        if self.nv is not None:
            self._thr = threading.Thread(target=self._loop, daemon=True)
            self._thr.start()

    def stop(self):
        self._stop.set()
        if self._thr is not None:
            self._thr.join()
        med = float(np.median(self.samples)) if self.samples else None
        return {"sm_mhz": med, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(self.samples)}


# ---- GPU arm ------------------------------------------------------------------------------------------
def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


def run_gpu_arm(a):
    import torch
    import torch.distributed as dist

    import applied_image_processing_b200 as aip
    from applied_image_processing_b200 import _cabi
    from applied_image_processing_b200 import localized_style_transfer as lst

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the colour-transfer path has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    L = _cabi.lib()

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    if a.workload == "16k":
        return run_gpu_16k(a, torch, dist, world, rank, local, dev, barrier, max_over_ranks)
    n, H, W = a.frames, a.height, a.width
    labels = a.workload == "labels"
    cs, ss, ms = make_frames(a, n, 1000 + 100 * rank)
    # pinned host copies (e2e leg) and device-resident copies (value leg)
    hc, hs, hm = (torch.from_numpy(x).pin_memory() for x in (cs, ss, ms))
    ho = torch.empty_like(hc).pin_memory()
    dc, ds, dm = hc.to(dev), hs.to(dev), hm.to(dev)
    do = torch.empty_like(dc)
    sh, sw = sty_size(a)
    fused = labels and a.labels_fused
    plan = lst.get_plan(dev, H, W, sh, sw, slots=a.slots * (8 if fused else 1))
    plan.set_option(_cabi.OPT_FRONTEND, a.front)
    plan.set_option(_cabi.OPT_APPLY, a.apply)
    plan.set_option(_cabi.OPT_MATCH, a.match)
    plan.set_option(_cabi.OPT_LUT, a.lut)
    plan.set_option(_cabi.OPT_KEYTAB, a.keytab)
    st = torch.cuda.current_stream().cuda_stream

    def step():
        if fused:  # configs[4], all 8 labels of a.slots frames per launch group
            do.copy_(dc)
            _cabi.check(L.rct_transfer_labels_async(plan.handle, dc.data_ptr(), H * W * 3, ds.data_ptr(), sh * sw * 3,
                                                    dm.data_ptr(), H * W, do.data_ptr(), H * W * 3, n, 8, st))
            return
        if labels:  # configs[4]: every label is an independent exact transfer; untouched pixels keep the content
            do.copy_(dc)
            for lab in range(8):
                _cabi.check(L.rct_transfer_async(plan.handle, dc.data_ptr(), H * W * 3, ds.data_ptr(), sh * sw * 3,
                                                 dm.data_ptr(), H * W, do.data_ptr(), H * W * 3, n,
                                                 _cabi.MODE_LABEL, lab, st))
            return
        if a.bins:
            _cabi.check(L.rct_transfer_binned_async(plan.handle, dc.data_ptr(), H * W * 3, ds.data_ptr(), sh * sw * 3,
                                                    dm.data_ptr(), H * W, do.data_ptr(), H * W * 3, n, _cabi.MODE_MASK, 0,
                                                    a.bins, st))
            return
        if lanes:  # the frames split over concurrent lanes: plan + stream each, joined on the timing stream
            main = torch.cuda.current_stream()
            ev0 = torch.cuda.Event()
            ev0.record(main)
            per = n // len(lanes)
            for k, (lp, ls) in enumerate(lanes):
                ls.wait_event(ev0)
                f0 = k * per
                _cabi.check(L.rct_transfer_async(lp.handle, dc[f0:].data_ptr(), H * W * 3, ds[f0:].data_ptr(), sh * sw * 3,
                                                 dm[f0:].data_ptr(), H * W, do[f0:].data_ptr(), H * W * 3, per,
                                                 _cabi.MODE_MASK, 0, ls.cuda_stream))
                e = torch.cuda.Event()
                e.record(ls)
                main.wait_event(e)
            return
        _cabi.check(L.rct_transfer_async(plan.handle, dc.data_ptr(), H * W * 3, ds.data_ptr(), sh * sw * 3,
                                         dm.data_ptr(), H * W, do.data_ptr(), H * W * 3, n, _cabi.MODE_MASK, 0, st))

    lanes = None
    if a.lanes > 1 and not labels and not a.bins and n % a.lanes == 0:
        lanes = []
        for k in range(a.lanes):
            lp = lst.Plan(dev, H, W, sh, sw, slots=max(1, min(a.slots, n // a.lanes)))
            for opt, val in ((_cabi.OPT_FRONTEND, a.front), (_cabi.OPT_APPLY, a.apply), (_cabi.OPT_MATCH, a.match),
                             (_cabi.OPT_LUT, a.lut), (_cabi.OPT_KEYTAB, a.keytab)):
                lp.set_option(opt, val)
            lanes.append((lp, torch.cuda.Stream()))
    for _ in range(max(a.warmup, 3)):
        step()
    sampler = ClockSampler(local)
    L.rct_launch_count(1)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    sampler.start()
    barrier()
    e0.record()
    for _ in range(a.steps):
        step()
    e1.record()
    barrier()
    clocks = sampler.stop()
    launches = int(L.rct_launch_count(0))
    ms_total = max_over_ranks(e0.elapsed_time(e1))
    n_status = n * 8 if fused else n
    if lanes:
        for lp, ls in lanes:
            status = np.zeros(n // len(lanes), np.int32)
            _cabi.check(L.rct_plan_status(lp.handle, status.ctypes.data, n // len(lanes), ls.cuda_stream))
            assert not status.any(), f"unexpected empty-region status {status}"
    else:
        status = np.zeros(n_status, np.int32)
        _cabi.check(L.rct_plan_status(plan.handle, status.ctypes.data, n_status, st))
        assert not status.any(), f"unexpected empty-region status {status}"
    mp_step = world * n * H * W / 1e6
    ms_step = ms_total / a.steps
    value = mp_step / (ms_step * 1e-3)

    # ---- per-kernel device time: a second pass of the same steps with one CUDA event after every
    # kernel on the launching stream (RCT_OPT_TIMING = 2); the value above is measured without them ----
    lanes_used, lanes = lanes, None   # the per-kernel breakdown is taken on ONE stream (kernels of concurrent lanes overlap)
    plan.set_option(_cabi.OPT_TIMING, 2)
    plan.kernel_timing(reset=True)
    for _ in range(a.steps):
        step()
    barrier()
    kt = plan.kernel_timing(reset=True)
    plan.set_option(_cabi.OPT_TIMING, 0)
    lanes = lanes_used
    groups = max(1, kt.pop("groups"))
    ms_launch = {k: v / groups for k, v in kt.items()}           # average duration of one launch
    t_all = sum(ms_launch.values())
    frames_launch = min(a.slots, n)
    apply_name = {0: "k_apply", 1: "k_apply_kt" if a.keytab and not a.lut and not a.bins else "k_apply_warp<bulk,8>", 2: "k_apply_warp<cp.async,8>", 3: "k_apply_tma",
                  4: "k_apply_warp<bulk,16>", 5: "k_apply_warp<bulk,8>", 6: "k_apply_warp<bulk,8,3stage>", 7: "k_apply_warp<bulk,8,bulk-store>", 8: "k_apply_warp<bulk,8>"}[0 if labels else a.apply]
    if (sh, sw) != (H, W) and not labels:
        apply_name = "k_apply_warp_rs" if a.apply == 8 else "k_apply<resample>"

    # ---- roofline: the per-pixel kernels stream the images (algorithmic bytes per pixel: SURVEY 8d /
    # DESIGN.md section 5); the dominant one of them is the headline, every kernel's time is listed ----
    peak, peak_src = peaks()
    traffic_db = load_traffic()
    kern = {}
    for kname, bpp in PIXEL_KERNELS.items():
        if kname not in ms_launch:
            continue
        bpp = bpp - 3.0 * (1.0 - sh * sw / (H * W))  # stylised bytes counted at their true size
        gb = bpp * frames_launch * H * W / 1e9
        shown = apply_name if kname == "k_apply" else kname
        tr = traffic_db.get(a.kind, {}).get(shown)
        kern[shown] = {"ms_per_launch": ms_launch[kname], "alg_bytes_per_pixel": bpp,
                       "achieved_GBps": gb / (ms_launch[kname] * 1e-3), "frac": gb / (ms_launch[kname] * 1e-3) / peak,
                       "share_of_step": ms_launch[kname] / t_all,
                       "ncu_dram_bytes_per_launch": (tr["dram_bytes_per_frame"] * frames_launch if tr else None)}
    # the kernel the north_star sets a roofline target for is the fused apply pass (>= 60 % of HBM at 4K);
    # it is reported at the top level, every kernel's share of the step is listed beside it
    dom = apply_name if apply_name in kern else max(kern, key=lambda k: kern[k]["ms_per_launch"])
    roofline = {
        "bound": "hbm", "kernel": dom, "achieved": kern[dom]["achieved_GBps"], "peak": peak, "unit": "GB/s",
        "frac": kern[dom]["frac"], "traffic": kern[dom]["ncu_dram_bytes_per_launch"], "peak_source": peak_src,
        "algorithmic_bytes_per_launch": kern[dom]["alg_bytes_per_pixel"] * frames_launch * H * W,
        "traffic_source": TRAFFIC_FILE + " (ncu --set full, dram__bytes_read+write per launch, scaled by frames per launch)",
        "pixel_kernels": kern,
        "all_kernels_ms_per_launch": {k: round(v, 5) for k, v in sorted(ms_launch.items(), key=lambda kv: -kv[1])},
        "share_of_step": {k: round(v / t_all, 4) for k, v in sorted(ms_launch.items(), key=lambda kv: -kv[1])},
        "pipeline_achieved_GBps_per_gpu": (ALG_BYTES_PER_PIXEL["pipeline"] - 3.0 * (1.0 - sh * sw / (H * W)))
        * n * H * W / 1e9 / (ms_step * 1e-3),
    }
    roofline["pipeline_frac"] = roofline["pipeline_achieved_GBps_per_gpu"] / peak
    if dom == "k_apply_kt" and "k_keytab" in ms_launch:
        # the key table exists only to serve this pass: the same bytes over apply + table build (k_kt_mark / k_kt_fill)
        t_both = ms_launch["k_apply"] + ms_launch["k_keytab"]
        roofline["frac_with_table_build"] = roofline["algorithmic_bytes_per_launch"] / 1e9 / (t_both * 1e-3) / peak
        roofline["table_build_ms_per_launch"] = ms_launch["k_keytab"]

    # ---- e2e: the public numpy API, pinned host buffers, copies inside the timed region ----------------
    e2e = None
    e2e_pageable = None
    if not a.no_e2e:
        hcn, hsn, hmn, hon = hc.numpy(), hs.numpy(), hm.numpy(), ho.numpy()

        e2e_slots = a.e2e_slots or None
        if a.stage_sets:
            lst.get_plan(dev, H, W, sh, sw, slots=e2e_slots or lst.host_group_size(n, 4 * H * W + 3 * sh * sw)).set_option(
                _cabi.OPT_STAGE_SETS, a.stage_sets)

        def api_call():
            if labels:
                aip.regional_color_transfer_labels(hcn, hsn, hmn, 8, slots=e2e_slots, fused=bool(a.labels_fused), out=hon)
            elif a.bins:
                aip.regional_color_transfer_binned(hcn, hsn, hmn, a.bins, out=hon)
            else:
                aip.regional_color_transfer(hcn, hsn, hmn, out=hon, slots=e2e_slots)

        for _ in range(2):
            api_call()
        barrier()
        t0 = time.perf_counter()
        for _ in range(a.steps):
            api_call()
        torch.cuda.synchronize()
        t_e2e = max_over_ranks(time.perf_counter() - t0)
        barrier()
        e2e = {"value": mp_step * a.steps / t_e2e, "unit": UNIT,
               "h2d_bytes_per_step": int(hc.numel() + hs.numel() + hm.numel()),
               "d2h_bytes_per_step": int(ho.numel()), "ms_per_step": 1e3 * t_e2e / a.steps}
        assert np.array_equal(hon, do.cpu().numpy()), "host path and device path disagree"
        # the box's raw copy ceiling for this step, measured in the same run: the same pinned buffers copied H2D and D2H
        # on two streams with nothing else, all ranks at once (8 GPUs of this box share one host memory system: 136 GB/s
        # in total, profiles/r02_pcie_ceiling.json) -- e2e cannot exceed it, whatever the kernels do
        s_in, s_out = torch.cuda.Stream(), torch.cuda.Stream()
        dprobe = torch.empty_like(dc), torch.empty_like(ds), torch.empty_like(dm)

        def raw_copies():
            with torch.cuda.stream(s_in):
                dprobe[0].copy_(hc, non_blocking=True)
                dprobe[1].copy_(hs, non_blocking=True)
                dprobe[2].copy_(hm, non_blocking=True)
            with torch.cuda.stream(s_out):
                ho.copy_(do, non_blocking=True)

        raw_copies()
        barrier()
        t0 = time.perf_counter()
        for _ in range(5):
            raw_copies()
        torch.cuda.synchronize()
        t_raw = max_over_ranks(time.perf_counter() - t0) / 5
        barrier()
        del dprobe
        e2e["raw_copy_ceiling"] = {"value": mp_step / t_raw, "unit": UNIT, "ms_per_step": 1e3 * t_raw,
                                   "what": "this step's H2D + D2H bytes from / to the same pinned buffers on two streams, no kernels, all ranks at once"}
        e2e["frac_of_ceiling"] = e2e["value"] / e2e["raw_copy_ceiling"]["value"]
        # the same call with plain (pageable) numpy arrays: what the reference's caller holds
        # (np.array(Image.open(...)), LST:200-219); the driver stages such copies itself
        if not labels and not a.bins:
            pc, ps, pm, po = cs, ss, ms, np.empty_like(cs)

            def api_call_pageable():
                aip.regional_color_transfer(pc, ps, pm, out=po, slots=e2e_slots)

            api_call_pageable()
            barrier()
            t0 = time.perf_counter()
            ksteps = max(1, a.steps // 2)
            for _ in range(ksteps):
                api_call_pageable()
            torch.cuda.synchronize()
            t_pg = max_over_ranks(time.perf_counter() - t0)
            barrier()
            e2e_pageable = {"value": mp_step * ksteps / t_pg, "unit": UNIT, "ms_per_step": 1e3 * t_pg / ksteps,
                            "steps": ksteps, "host_buffers": "pageable numpy arrays in and out"}
            assert np.array_equal(po, hon)

    # ---- e2e_files: the reference's real boundary (LST:191-245: files in, file out), one C-ABI call per frame ------
    e2e_files = None
    if world == 1 and a.files and not labels and not a.bins and not a.no_e2e:
        try:
            e2e_files = files_leg(a, aip, cs, ss, ms, H, W)
        except AssertionError:
            raise  # a parity failure must fail the bench
        except Exception as e:  # noqa: BLE001 -- an auxiliary leg must not take the headline line down with it
            e2e_files = {"error": f"{type(e).__name__}: {e}"}

    # ---- CPU baseline: C port of the reference on this box's host cores (rank 0, N=1 only) ------------
    cpu = None
    parity_frames = 0
    if world == 1 and not a.no_cpu:
        nc, threads = cpu_sample_size(a)
        nc = min(nc, n)
        threads = min(threads, nc)
        t, cpu_outs = cpu_pass((cs[:nc], ss[:nc], ms[:nc]), threads, labels, keep=True, bins=a.bins)
        cpu = {"value": nc * H * W / 1e6 / t, "unit": UNIT, "cores": threads, "host_cores_available": host_threads(), "kind": "port",
               "sample": f"{nc} frame(s) of the step's batch, {threads} thread(s), one pass ({t:.1f} s)"}
        # parity at the benchmarked size: the GPU output of the same frames equals the oracle's byte for byte
        got = do[:nc].cpu().numpy()
        for i in range(nc):
            nbad = int((got[i] != cpu_outs[i]).sum())
            assert nbad == 0, f"bench parity: frame {i}: {nbad} bytes differ from the oracle"
        parity_frames = nc

    if a.keytab and not a.lut and not a.bins and not labels:
        # Key-table self-check, every slot of the last launch group: the table answers in fp32 (lg2.approx + an fma chain)
        # under an error bound of 2^-16 on the key; each distinct target colour is re-checked against the exact LUT here.
        # A toolchain or architecture that loosened lg2.approx must fail this run, not shift pixels silently.
        nslots = min(int(getattr(plan, "slots", 1)), n)
        checks = [plan.debug_keytab(k, key_error=(k == 0)) for k in range(nslots)]
        roofline["keytab_slot0"] = checks[0]
        roofline["keytab_slots_checked"] = nslots
        for k, c in enumerate(checks):
            assert c["colours_wrong"] == 0 and c["pixels_wrong"] == 0, f"bench: key table of slot {k} disagrees with the LUT: {c}"
        if checks[0]["valid"]:
            assert checks[0]["max_key_err"] < 2.0 ** -17, f"bench: fp32 key error {checks[0]['max_key_err']:.3g} >= 2^-17 (bound 2^-16)"
    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": a.steps, "warmup": max(a.warmup, 3),
            "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic",
            "config": config_dict(a, world),
            "clocks": clocks, "e2e": e2e, "e2e_pageable": e2e_pageable, "gpu_launches": launches, "roofline": roofline,
            "cpu_baseline": cpu, "parity_checked_frames": parity_frames, "e2e_files": e2e_files,
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def run_gpu_16k(a, torch, dist, world, rank, local, dev, barrier, max_over_ranks):
    """BASELINE.json configs[3]: ONE image, rows sharded over the ranks; the only exchange is the
    all-reduce (NCCL) of the two colour-cube histograms.  Strong scaling: total work is fixed."""
    from applied_image_processing_b200 import _cabi, sharded
    from tests.synth import ellipse_mask, synth_image

    H, W = a.height, a.width
    L = _cabi.lib()
    c = synth_image(H, W, 2000, a.kind)
    s = synth_image(H, W, 2001, a.kind)
    m = ellipse_mask(H, W, 0.4, 2002)
    tr = sharded.ShardedTransfer(H, W, H, W, dev)
    r0, r1 = tr.row0, tr.row0 + tr.rows
    hc, hs, hm = (torch.from_numpy(np.ascontiguousarray(x[r0:r1])).pin_memory() for x in (c, s, m))
    del c, s, m
    ho = torch.empty_like(hc).pin_memory()
    dc, ds, dm = hc.to(dev), hs.to(dev), hm.to(dev)
    do = torch.empty_like(dc)
    run = (lambda timed=False: tr.binned(dc, ds, dm, do, bins=a.bins, timed=timed)) if a.bins else (
        lambda timed=False: tr(dc, ds, dm, do, timed=timed))
    for _ in range(max(a.warmup, 3)):
        run()
    # everything slow and rank-dependent (NVML init of the clock sampler) happens BEFORE the barrier: the step contains
    # a collective, so a rank that enters the timed loop late makes every other rank wait inside its first all-reduce
    # (measured: 20-30 ms of skew = 4-6 ms per step over 5 steps, charged to the "all-reduce")
    sampler = ClockSampler(local)
    L.rct_launch_count(1)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    sampler.start()
    tr.host_s.clear()
    barrier()
    e0.record()
    t_host0 = time.perf_counter()
    for _ in range(a.steps):
        run()
    host_enqueue_ms = 1e3 * (time.perf_counter() - t_host0) / a.steps  # host time to ENQUEUE a step (no sync inside)
    e1.record()
    barrier()
    clocks = sampler.stop()
    launches = int(L.rct_launch_count(0))
    ms_step = max_over_ranks(e0.elapsed_time(e1)) / a.steps
    host_enqueue = {"total_ms_per_step_max_over_ranks": max_over_ranks(host_enqueue_ms)}
    for k in sorted(tr.host_s):
        host_enqueue[k + "_ms"] = max_over_ranks(1e3 * tr.host_s[k] / a.steps)
    assert tr.status() == 0
    mp_step = H * W / 1e6
    value = mp_step / (ms_step * 1e-3)
    # phase breakdown: a second pass with CUDA events around the phases and a host synchronisation per step (the value
    # above is measured without either)
    acc = ({"front_moments_ms": 0.0, "keys_ms": 0.0, "hist_ms": 0.0, "match_apply_ms": 0.0, "allreduce_ms": 0.0} if a.bins
           else {"hist_ms": 0.0, "allreduce_ms": 0.0, "solve_ms": 0.0, "apply_ms": 0.0})
    for _ in range(a.steps):
        run(timed=True)
        for k, v in (tr.binned_phase_ms() if a.bins else tr.phase_ms()).items():
            acc[k] += v
    barrier()
    phases = {k: max_over_ranks(v / a.steps) for k, v in acc.items()}
    peak, peak_src = peaks()
    npix_rank = (H + world - 1) // world * W
    if a.bins:  # phase 0 = front end + moments (7 B/px), phase 3 = LUT build + apply (10 B/px)
        kern = {
            "front+moments": {"ms_per_launch": phases["front_moments_ms"], "alg_bytes_per_pixel": 7.0,
                              "achieved_GBps": 7.0 * npix_rank / 1e9 / (phases["front_moments_ms"] * 1e-3)},
            "match+apply": {"ms_per_launch": phases["match_apply_ms"], "alg_bytes_per_pixel": 10.0,
                            "achieved_GBps": 10.0 * npix_rank / 1e9 / (phases["match_apply_ms"] * 1e-3)},
        }
    else:
        kern = {
            "k_hist": {"ms_per_launch": phases["hist_ms"], "alg_bytes_per_pixel": 7.0,
                       "achieved_GBps": 7.0 * npix_rank / 1e9 / (phases["hist_ms"] * 1e-3)},
            "k_apply": {"ms_per_launch": phases["apply_ms"], "alg_bytes_per_pixel": 10.0,
                        "achieved_GBps": 10.0 * npix_rank / 1e9 / (phases["apply_ms"] * 1e-3)},
        }
    dom = max(kern, key=lambda k: kern[k]["ms_per_launch"])
    hist_bytes = (sum(int(L.rct_plan_xchg_bytes(w, a.bins)) for w in range(3)) if a.bins
                  else int(L.rct_plan_hist_bytes(tr.plan.handle)))
    roofline = {
        "bound": "hbm", "kernel": dom, "achieved": kern[dom]["achieved_GBps"], "peak": peak, "unit": "GB/s",
        "frac": kern[dom]["achieved_GBps"] / peak, "traffic": None, "peak_source": peak_src, "kernels": kern,
        "phase_ms_max_over_ranks": phases, "host_enqueue": host_enqueue, "allreduce_bytes": hist_bytes,
        "allreduce_busbw_GBps": (2.0 * (world - 1) / world * hist_bytes / 1e9 / (phases["allreduce_ms"] * 1e-3)
                                 if world > 1 and phases["allreduce_ms"] > 0 else None),
        "pipeline_achieved_GBps_all_gpus": 10.0 * H * W / 1e9 / (ms_step * 1e-3),
    }
    # the collective alone, on the plan's buffer and on a plain torch tensor of the same size (68 GB/s at
    # N = 8 on the plan's buffer was never explained)
    if world > 1 and not a.bins:
        probe = {}
        plain = torch.zeros(hist_bytes // 4, dtype=torch.int32, device=dev)
        for name, buf in (("plan_buffer", tr.hist), ("torch_tensor", plain)):
            for _ in range(2):
                dist.all_reduce(buf, op=dist.ReduceOp.SUM)
            barrier()
            ea, eb = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            ea.record()
            for _ in range(5):
                dist.all_reduce(buf, op=dist.ReduceOp.SUM)
            eb.record()
            barrier()
            ms = max_over_ranks(ea.elapsed_time(eb) / 5)
            probe[name] = {"ms": ms, "busbw_GBps": 2.0 * (world - 1) / world * hist_bytes / 1e9 / (ms * 1e-3)}
        roofline["allreduce_alone"] = probe
    e2e = None
    if not a.no_e2e:
        def api_call():
            dc.copy_(hc, non_blocking=True)
            ds.copy_(hs, non_blocking=True)
            dm.copy_(hm, non_blocking=True)
            run()
            ho.copy_(do, non_blocking=True)
            torch.cuda.synchronize()

        api_call()
        barrier()
        t0 = time.perf_counter()
        for _ in range(a.steps):
            api_call()
        t_e2e = max_over_ranks(time.perf_counter() - t0)
        barrier()
        e2e = {"value": mp_step * a.steps / t_e2e, "unit": UNIT,
               "h2d_bytes_per_step": int(hc.numel() + hs.numel() + hm.numel()) * world,
               "d2h_bytes_per_step": int(ho.numel()) * world, "ms_per_step": 1e3 * t_e2e / a.steps}
    cpu = None
    if world == 1 and not a.no_cpu:
        hb = H // 16
        threads = max(1, min(host_threads(), 8))
        cs = np.stack([hc.numpy()[i * hb:(i + 1) * hb] for i in range(threads)])
        ss = np.stack([hs.numpy()[i * hb:(i + 1) * hb] for i in range(threads)])
        ms_ = np.stack([hm.numpy()[i * hb:(i + 1) * hb] for i in range(threads)])
        t = cpu_pass((cs, ss, ms_), threads, bins=a.bins)
        cpu = {"value": threads * hb * W / 1e6 / t, "unit": UNIT, "cores": threads, "kind": "port",
               "sample": f"{threads} bands of {hb} rows of the image, one per thread, one pass ({t:.1f} s)"}
    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": a.steps, "warmup": max(a.warmup, 3),
            "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic",
            "config": config_dict(a, world),
            "clocks": clocks, "e2e": e2e, "gpu_launches": launches, "roofline": roofline, "cpu_baseline": cpu,
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    a = parse_args()
    if a.impl == "reference":
        run_reference_arm(a)
    else:
        run_gpu_arm(a)


if __name__ == "__main__":
    main()
