"""Drop-in for the hot path of the reference ``Style_3DGS/localized_style_transfer.py`` (LST).

Same function names, positional order, defaults, return types and warning/empty-region
behaviour as LST:11-168 and the tail LST:222-238, but every function runs hand-written
sm_100a CUDA kernels through the C ABI in ``include/rct.h`` (``librct.so``).  numpy arrays
in -> numpy arrays out (the reference's convention); CUDA ``torch`` tensors in -> CUDA
tensors out, no host round trip.  There is no CPU fallback.

PyTorch is used only for device memory, streams and (in :mod:`.sharded`) torch.distributed.
"""
from __future__ import annotations

import collections
import ctypes as C
import os
import threading

import numpy as np
import torch

from . import _cabi
from ._cabi import MODE_ALL, MODE_LABEL, MODE_MASK, check

# ---- a-0: matrices for Reinhard's transform, computed exactly as LST:12-19 ------------------
RGB_TO_LMS = np.array(
    [[0.3811, 0.5783, 0.0402], [0.1967, 0.7244, 0.0782], [0.0241, 0.1288, 0.8444]]
)
LMS_TO_LAB = np.array(
    [[1 / np.sqrt(3), 0, 0], [0, 1 / np.sqrt(6), 0], [0, 0, 1 / np.sqrt(2)]]
) @ np.array([[1, 1, 1], [1, 1, -2], [1, -1, 0]])
LAB_TO_LMS = np.linalg.inv(LMS_TO_LAB)
LMS_TO_RGB = np.linalg.inv(RGB_TO_LMS)

_matrices_installed = False
_plan_lock = threading.Lock()
_plans: "collections.OrderedDict" = collections.OrderedDict()
# The plan cache is an LRU bounded by workspace bytes (a caller feeding arbitrary image sizes -- the
# reference's real use -- must not grow device memory without bound).  The most recent plan always stays.
PLAN_CACHE_BYTES = int(os.environ.get("RCT_PLAN_CACHE_BYTES", 24 << 30))


def _install_matrices():
    global _matrices_installed
    if not _matrices_installed:
        m = np.ascontiguousarray(
            np.stack([RGB_TO_LMS, LMS_TO_LAB, LAB_TO_LMS, LMS_TO_RGB]).astype(np.float64)
        )
        check(_cabi.lib().rct_set_matrices(m.ctypes.data_as(C.c_void_p)))
        _matrices_installed = True


def _require_cuda() -> torch.device:
    if not torch.cuda.is_available():
        raise RuntimeError(
            "applied_image_processing_b200 needs a CUDA device (B200, sm_100a); "
            "there is no CPU fallback for the colour-transfer path"
        )
    return torch.device("cuda", torch.cuda.current_device())


def _stream_ptr() -> int:
    return torch.cuda.current_stream().cuda_stream


def _is_tensor(x) -> bool:
    return isinstance(x, torch.Tensor)


class Plan:
    """A ``rct_plan`` plus the torch-owned workspace it lives in."""

    def __init__(self, device: torch.device, H, W, sh, sw, slots=1, flags=0):
        _install_matrices()
        L = _cabi.lib()
        self.device, self.H, self.W, self.sh, self.sw, self.slots, self.flags = device, H, W, sh, sw, slots, flags
        nbytes = L.rct_plan_workspace_bytes(H, W, slots, flags)
        if nbytes == 0:
            raise ValueError("bad plan geometry")
        self.nbytes = int(nbytes)
        with torch.cuda.device(device):
            self.workspace = torch.empty(nbytes, dtype=torch.uint8, device=device)
            handle = C.c_void_p()
            check(L.rct_plan_create(device.index, H, W, sh, sw, slots, flags, self.workspace.data_ptr(),
                                    nbytes, C.byref(handle)))
        self.handle = handle
        self.lock = threading.Lock()

    def __del__(self):
        try:
            if getattr(self, "handle", None):
                _cabi.lib().rct_plan_destroy(self.handle)
                self.handle = None
        except Exception:
            pass

    def on_current_stream(self):
        """Context for enqueueing on the caller's current stream: tells the caching allocator that the workspace (allocated on
        whatever stream was current when the plan was made) is in use there, so that a plan dropped from the LRU cache right
        after the call cannot have its memory handed out while the kernels still run."""
        import contextlib

        self.workspace.record_stream(torch.cuda.current_stream(self.device))
        return contextlib.nullcontext()

    def set_option(self, option: int, value: int):
        check(_cabi.lib().rct_plan_set_option(self.handle, option, value))

    def timing(self, reset=True):
        """Per-phase device time (ms) accumulated since the last reset, measured with CUDA events
        on the launching stream (needs ``set_option(OPT_TIMING, 1)``)."""
        ms = (C.c_double * 4)()
        groups = C.c_longlong(0)
        check(_cabi.lib().rct_plan_timing(self.handle, ms, C.byref(groups), int(reset)))
        return dict(memset_ms=ms[0], hist_ms=ms[1], solve_ms=ms[2], apply_ms=ms[3], groups=groups.value)

    def kernel_timing(self, reset=True):
        """Per-kernel device time (ms) accumulated since the last reset (``set_option(OPT_TIMING, 2)``:
        one CUDA event after every kernel, on the launching stream)."""
        L = _cabi.lib()
        ms = (C.c_double * _cabi.TK_COUNT)()
        groups = C.c_longlong(0)
        check(L.rct_plan_kernel_timing(self.handle, ms, C.byref(groups), int(reset)))
        out = {L.rct_kernel_timing_name(i).decode(): ms[i] for i in range(1, _cabi.TK_COUNT) if ms[i] > 0.0}
        out["groups"] = groups.value
        return out

    # -- introspection (parity tests) -------------------------------------------------------
    def info(self, slot=0) -> dict:
        fi = _cabi.FrameInfo()
        check(_cabi.lib().rct_plan_info(self.handle, slot, C.byref(fi)))
        return dict(
            n_target=fi.n_target, n_source=fi.n_source, d_target=fi.d_target, d_source=fi.d_source,
            t_mean=np.array(fi.t_mean), t_comp=np.array(fi.t_comp), s_mean=np.array(fi.s_mean),
            s_comp=np.array(fi.s_comp), status=fi.status,
        )

    def debug_hist(self, slot=0):
        ht = np.empty(1 << 24, np.uint32)
        hs = np.empty(1 << 24, np.uint32)
        check(_cabi.lib().rct_plan_debug_hist(self.handle, slot, ht.ctypes.data, hs.ctypes.data))
        return ht, hs

    def debug_target(self, slot=0) -> dict:
        d = self.info(slot)["d_target"]
        out = dict(colour=np.empty(d, np.uint32), key=np.empty(d, np.float64), j=np.empty(d, np.int64),
                   matched=np.empty(d, np.float64), cum=np.empty(d, np.uint32))
        check(_cabi.lib().rct_plan_debug_target(self.handle, slot, out["colour"].ctypes.data,
                                                out["key"].ctypes.data, out["j"].ctypes.data,
                                                out["matched"].ctypes.data, out["cum"].ctypes.data))
        return out

    def debug_binned(self, bins, slot=0) -> dict:
        """Histograms, key ranges and (FLAG_DEBUG plans) per-target-colour CDF index pairs of the binned mode."""
        d = self.info(slot)["d_target"] if (self.flags & _cabi.FLAG_DEBUG) else 0
        out = dict(hist_t=np.zeros(bins, np.uint32), hist_s=np.zeros(bins, np.uint32), range=np.zeros(4, np.float64),
                   colour=np.empty(d, np.uint32), code=np.empty(d, np.int64), matched=np.empty(d, np.float64))
        dbg = bool(self.flags & _cabi.FLAG_DEBUG)
        check(_cabi.lib().rct_plan_debug_binned(
            self.handle, slot, bins, out["hist_t"].ctypes.data, out["hist_s"].ctypes.data, out["range"].ctypes.data,
            out["colour"].ctypes.data if dbg else None, out["code"].ctypes.data if dbg else None,
            out["matched"].ctypes.data if dbg else None))
        return out

    def debug_keytab(self, slot=0, key_error=True) -> dict:
        """Self-check of the key table of the last launch group (rct_plan_debug_keytab)."""
        st = (C.c_ulonglong * 10)()
        err = C.c_double(0.0)
        check(_cabi.lib().rct_plan_debug_keytab(self.handle, slot, st, C.byref(err) if key_error else None))
        names = ("valid", "runs", "cells", "flagged_cells", "colours_pure", "colours_wrong", "colours_flagged",
                 "pixels_pure", "pixels_wrong", "pixels_flagged")
        out = {k: int(v) for k, v in zip(names, st)}
        out["max_key_err"] = float(err.value)
        return out

    def debug_lut(self, slot=0):
        lut = np.empty(1 << 24, np.uint32)
        check(_cabi.lib().rct_plan_debug_lut(self.handle, slot, lut.ctypes.data))
        return lut


def get_plan(device, H, W, sh, sw, slots=1, flags=0, lane=0) -> Plan:
    """Per-(device, geometry) plan cache (workspace cache behind a mutex), least recently
    used plans are dropped once the cached workspaces exceed PLAN_CACHE_BYTES.  A dropped plan that a
    caller still holds stays valid (it is destroyed when the last reference goes).  `lane` distinguishes
    the plans of launch groups that run concurrently on separate streams (see LANES)."""
    key = (device.index, H, W, sh, sw, slots, flags, lane)
    with _plan_lock:
        p = _plans.get(key)
        if p is None:
            need = int(_cabi.lib().rct_plan_workspace_bytes(H, W, slots, flags))
            total = sum(q.nbytes for q in _plans.values())
            while _plans and total + need > PLAN_CACHE_BYTES:
                _, old = _plans.popitem(last=False)
                total -= old.nbytes
                del old
            p = Plan(device, H, W, sh, sw, slots, flags)
            _plans[key] = p
        else:
            _plans.move_to_end(key)
        return p


def cached_plan_bytes() -> int:
    with _plan_lock:
        return sum(q.nbytes for q in _plans.values())


def clear_plans():
    with _plan_lock:
        _plans.clear()


def _warn(status: int):
    # the reference prints, never raises (LST:141-147)
    if status == _cabi.RCT_EMPTY_TARGET:
        print("Warning: No foreground pixels found.")
    elif status == _cabi.RCT_EMPTY_SOURCE:
        print("Warning: No background pixels found for color transfer.")
    elif status == 3:
        raise _cabi.RctError("palette LUT overflow (> 65536 runs of output colour): set OPT_LUT to 0 for this plan")


def _np_u8(a, name):
    a = np.asarray(a)
    if a.dtype != np.uint8:
        raise TypeError(f"{name} must be uint8, got {a.dtype}")
    return np.ascontiguousarray(a)


# Launch groups in flight at once on the tensor path.  The solve kernels are latency-bound (ncu: k_scatter issues on 17 %
# of its cycles), so two groups on two streams, each with its own plan, overlap their stalls: 8 + 8 frames of 4K take
# 11.03 ms instead of 11.68 ms back to back (measured on B200; four lanes are slower than two).
# Memory: every lane holds a plan workspace of its own for the geometry (about 1 GB per 4K slot, so 2 x 8 GB for batches
# of 8-frame groups at 4K), counted against PLAN_CACHE_BYTES like any other plan.  Set RCT_LANES=1 (or this attribute)
# where device memory matters more than the last 5 % of throughput.
LANES = max(1, int(os.environ.get("RCT_LANES", 2)))
_lane_streams: dict = {}


def _lane_stream(device, lane) -> torch.cuda.Stream:
    key = (device.index, lane)
    with _plan_lock:
        st = _lane_streams.get(key)
        if st is None:
            st = _lane_streams[key] = torch.cuda.Stream(device=device)
        return st


def host_group_size(n: int, frame_bytes: int = 58_060_800) -> int:
    """Frames per copy/compute group of the host (numpy) entry point: about 60 MB of input per group, at least four
    groups per call when there are that many frames.  Small groups keep the un-overlapped tail (the last group's compute
    and copy-out) short, but a group must be large enough for its kernels to run efficiently while the next one is
    copied in.  Measured on B200, MP/s end to end: 16 x 4K frames (58 MB each) in groups of 1 / 2: 6 751 / 6 467;
    32 x 1080p frames (14.5 MB each) in groups of 1 / 2 / 4 / 8: 3 369 / 4 745 / 5 733 / 5 550."""
    g = int(round(60e6 / max(1, frame_bytes)))
    g = max(1, min(8, g))
    return max(1, min(g, n // 4)) if n >= 4 else 1


def _transfer(content, stylised, mask, mode, label=0, out=None, slots=None, quiet=False, return_status=False):
    """Shared driver: one image [H,W,3] or a batch [N,H,W,3]; numpy or CUDA tensors.
    return_status (tensor path): also return the per-frame status array (0 ok / 1 no target pixels /
    2 no source pixels); fetching it synchronises the stream, so it is off by default there -- the numpy
    path always reads it and prints the reference's warnings (LST:141-147)."""
    L = _cabi.lib()
    if _is_tensor(content):
        dev = content.device
        if dev.type != "cuda":
            raise RuntimeError("tensor inputs must live on a CUDA device (no CPU fallback)")
        for t in (content, stylised) + ((mask,) if mask is not None else ()):
            if t.dtype != torch.uint8 or not t.is_contiguous() or t.device != dev:
                raise TypeError("inputs must be contiguous uint8 tensors on one device")
        batched = content.dim() == 4
        n = content.shape[0] if batched else 1
        H, W = content.shape[-3], content.shape[-2]
        sty_b = stylised.dim() == 4
        sh, sw = stylised.shape[-3], stylised.shape[-2]
        mask_b = mask is not None and mask.dim() == 3
        group = slots or min(n, 8)
        if out is None:
            out = torch.empty_like(content) if mode != MODE_LABEL else content.clone()
        nlanes = LANES if (LANES > 1 and n >= 2 * group and not return_status) else 1
        if nlanes > 1:
            # contiguous blocks of whole groups per lane, each lane its own plan and stream, joined on the caller's stream
            main = torch.cuda.current_stream(dev)
            start = torch.cuda.Event()
            start.record(main)
            ngroups = (n + group - 1) // group
            f0 = 0
            with torch.cuda.device(dev):
                for lane in range(nlanes):
                    g = ngroups // nlanes + (1 if lane < ngroups % nlanes else 0)
                    cnt = min(n - f0, g * group)
                    if cnt <= 0:
                        break
                    lp, ls = get_plan(dev, H, W, sh, sw, group, lane=lane), _lane_stream(dev, lane)
                    # the workspace was allocated on another stream: if the plan is evicted from the cache and freed
                    # while this lane still runs, the allocator must not hand the memory out before the lane is done
                    lp.workspace.record_stream(ls)
                    ls.wait_event(start)
                    with lp.lock:
                        check(L.rct_transfer_async(
                            lp.handle, content[f0:].data_ptr(), H * W * 3,
                            (stylised[f0:] if sty_b else stylised).data_ptr(), sh * sw * 3 if sty_b else 0,
                            (mask[f0:] if mask_b else mask).data_ptr() if mask is not None else None, H * W if mask_b else 0,
                            out[f0:].data_ptr(), H * W * 3, cnt, mode, label, ls.cuda_stream))
                    done = torch.cuda.Event()
                    done.record(ls)
                    main.wait_event(done)
                    f0 += cnt
            return out
        plan = get_plan(dev, H, W, sh, sw, group)
        with torch.cuda.device(dev), plan.lock, plan.on_current_stream():
            check(L.rct_transfer_async(
                plan.handle, content.data_ptr(), H * W * 3, stylised.data_ptr(), sh * sw * 3 if sty_b else 0,
                mask.data_ptr() if mask is not None else None, H * W if mask_b else 0, out.data_ptr(),
                H * W * 3, n, mode, label, _stream_ptr()))
            if return_status:
                status = np.zeros(n, np.int32)
                check(L.rct_plan_status(plan.handle, status.ctypes.data, n, _stream_ptr()))
        if return_status:
            if not quiet:
                for st_ in status:
                    _warn(int(st_))
            return out, status
        return out
    # numpy: the reference's own calling convention
    dev = _require_cuda()
    content = _np_u8(content, "content")
    stylised = _np_u8(stylised, "stylised")
    mask = _np_u8(mask, "mask") if mask is not None else None
    batched = content.ndim == 4
    n = content.shape[0] if batched else 1
    H, W = content.shape[-3], content.shape[-2]
    sh, sw = stylised.shape[-3], stylised.shape[-2]
    if batched and stylised.ndim != 4:
        stylised = np.ascontiguousarray(np.broadcast_to(stylised, (n,) + stylised.shape))
    if batched and mask is not None and mask.ndim != 3:
        mask = np.ascontiguousarray(np.broadcast_to(mask, (n,) + mask.shape))
    # host path: groups of ~n/8 frames so that copy-in, compute and copy-out of successive groups
    # overlap inside rct_transfer_host (a ring of staging sets, three streams) and the un-overlapped
    # head (first copy-in) and tail (last compute + copy-out) stay one small group each
    # (measured, 8 x 4K: groups of 1 -> 10.6 ms per call, 2 -> 11.7, 4 -> 14.1; PCIe floor 9.0)
    plan = get_plan(dev, H, W, sh, sw, slots or host_group_size(n, 4 * H * W + 3 * sh * sw))
    if out is None:
        out = np.empty_like(content) if mode != MODE_LABEL else content.copy()
    status = np.zeros(n, np.int32)
    with torch.cuda.device(dev), plan.lock, plan.on_current_stream():
        check(L.rct_transfer_host(plan.handle, content.ctypes.data, stylised.ctypes.data,
                                  mask.ctypes.data if mask is not None else None, out.ctypes.data, n, mode,
                                  label, status.ctypes.data))
    if not quiet:
        for s in status:
            _warn(int(s))
    return (out, status) if return_status else out


# ---- the reference's functions, same names and signatures ------------------------------------
def rgb_to_lab_pixels(pixels_uint8):
    """LST:67-77.  Nx3 uint8 RGB -> Nx3 float64 Reinhard lab."""
    L = _cabi.lib()
    _install_matrices()
    if _is_tensor(pixels_uint8):
        px = pixels_uint8.contiguous()
        lab = torch.empty(px.shape, dtype=torch.float64, device=px.device)
        with torch.cuda.device(px.device):
            check(L.rct_rgb_to_lab(px.data_ptr(), lab.data_ptr(), px.numel() // 3, _stream_ptr()))
        return lab
    dev = _require_cuda()
    px = torch.from_numpy(_np_u8(pixels_uint8, "pixels")).to(dev)
    return rgb_to_lab_pixels(px).cpu().numpy()


def lab_to_rgb_pixels(lab_pixels, return_float=False):
    """LST:80-89.  Nx3 float64 lab -> Nx3 uint8 RGB (clip, *255, truncate)."""
    L = _cabi.lib()
    _install_matrices()
    if _is_tensor(lab_pixels):
        lab = lab_pixels.to(torch.float64).contiguous()
        rgb = torch.empty(lab.shape, dtype=torch.uint8, device=lab.device)
        rgbf = torch.empty(lab.shape, dtype=torch.float64, device=lab.device) if return_float else None
        with torch.cuda.device(lab.device):
            check(L.rct_lab_to_rgb(lab.data_ptr(), rgb.data_ptr(), rgbf.data_ptr() if return_float else None,
                                   lab.numel() // 3, _stream_ptr()))
        return (rgb, rgbf) if return_float else rgb
    dev = _require_cuda()
    lab = torch.from_numpy(np.ascontiguousarray(lab_pixels, np.float64)).to(dev)
    r = lab_to_rgb_pixels(lab, return_float)
    return (r[0].cpu().numpy(), r[1].cpu().numpy()) if return_float else r.cpu().numpy()


def rgb_to_lab_image(image_uint8):
    """LST:22-40.  [H,W,3] uint8 -> [H,W,3] float64 lab."""
    shape = tuple(image_uint8.shape)
    return rgb_to_lab_pixels(image_uint8.reshape(-1, 3)).reshape(shape)


def lab_to_rgb_image(lab):
    """LST:43-61.  [H,W,3] lab -> [H,W,3] uint8."""
    shape = tuple(lab.shape)
    return lab_to_rgb_pixels(lab.reshape(-1, 3)).reshape(shape)


class RegionPCA:
    """What the hot path uses of sklearn's fitted ``PCA(n_components=1)`` (LST:92-96, 161)."""

    n_components = 1

    def __init__(self, mean, comp):
        self._mean, self._comp = mean, comp  # device float64 [3]

    @property
    def mean_(self):
        return self._mean.cpu().numpy()

    @property
    def components_(self):
        return self._comp.cpu().numpy().reshape(1, 3)

    def inverse_transform(self, X):
        """sklearn _base.py:205: X @ components_ + mean_."""
        L = _cabi.lib()
        as_np = not _is_tensor(X)
        m = (torch.from_numpy(np.ascontiguousarray(X, np.float64)).to(self._mean.device) if as_np
             else X.to(torch.float64)).contiguous().reshape(-1)
        lab = torch.empty((m.numel(), 3), dtype=torch.float64, device=m.device)
        with torch.cuda.device(m.device):
            check(L.rct_pca_inverse(m.data_ptr(), m.numel(), self._mean.data_ptr(), self._comp.data_ptr(),
                                    lab.data_ptr(), _stream_ptr()))
        return lab.cpu().numpy() if as_np else lab


def apply_pca(lab_data):
    """LST:92-96: ``PCA(n_components=1).fit_transform`` -> (projection [N,1], pca)."""
    L = _cabi.lib()
    _install_matrices()
    as_np = not _is_tensor(lab_data)
    dev = _require_cuda() if as_np else lab_data.device
    lab = (torch.from_numpy(np.ascontiguousarray(lab_data, np.float64)).to(dev) if as_np
           else lab_data.to(torch.float64)).contiguous()
    n = lab.shape[0]
    proj = torch.empty((n, 1), dtype=torch.float64, device=dev)
    mean = torch.empty(3, dtype=torch.float64, device=dev)
    comp = torch.empty(3, dtype=torch.float64, device=dev)
    scratch = torch.empty(L.rct_pca_scratch_bytes(), dtype=torch.uint8, device=dev)
    with torch.cuda.device(dev):
        check(L.rct_apply_pca(lab.data_ptr(), n, proj.data_ptr(), mean.data_ptr(), comp.data_ptr(),
                              scratch.data_ptr(), _stream_ptr()))
    pca = RegionPCA(mean, comp)
    return (proj.cpu().numpy() if as_np else proj), pca


def match_cdf(target_proj, source_proj, return_index=False):
    """LST:99-125: exact sort-based CDF matching along one dimension -> [N,1]."""
    L = _cabi.lib()
    as_np = not _is_tensor(target_proj)
    dev = _require_cuda() if as_np else target_proj.device
    t = (torch.from_numpy(np.ascontiguousarray(target_proj, np.float64)).to(dev) if as_np
         else target_proj.to(torch.float64)).contiguous().reshape(-1)
    s = (torch.from_numpy(np.ascontiguousarray(source_proj, np.float64)).to(dev) if not _is_tensor(source_proj)
         else source_proj.to(torch.float64)).contiguous().reshape(-1)
    matched = torch.empty(t.numel(), dtype=torch.float64, device=dev)
    jidx = torch.empty(t.numel(), dtype=torch.int64, device=dev) if return_index else None
    scratch = torch.empty(L.rct_match_scratch_bytes(t.numel(), s.numel()), dtype=torch.uint8, device=dev)
    with torch.cuda.device(dev):
        check(L.rct_match_cdf(t.data_ptr(), t.numel(), s.data_ptr(), s.numel(), matched.data_ptr(),
                              jidx.data_ptr() if return_index else None, scratch.data_ptr(), _stream_ptr()))
        if as_np:
            torch.cuda.current_stream().synchronize()
    matched = matched.reshape(-1, 1)
    if as_np:
        matched = matched.cpu().numpy()
        if return_index:
            jidx = jidx.cpu().numpy()
    return (matched, jidx) if return_index else matched


def color_transfer_foreground(foreground_img, background_img):
    """LST:128-168.  Recolour the non-black pixels of ``foreground_img`` so that their
    principal-axis distribution matches the non-black pixels of ``background_img``.

    The two images need not have the same shape: the reference gathers ALL non-black pixels of each
    (boolean gather, LST:134-139) and never resizes here.  Black pixels take no part in the statistics,
    the ranking or the write-back, so both pixel lists are padded with black to a common length and run
    as 1 x n images -- exactly the reference's semantics, no resample tables involved."""
    fs, bs = tuple(foreground_img.shape), tuple(background_img.shape)
    if len(fs) == 3 and len(bs) == 3 and fs[:2] != bs[:2]:
        nf, nb = fs[0] * fs[1], bs[0] * bs[1]
        n = max(nf, nb)
        if _is_tensor(foreground_img):
            f = torch.zeros((1, n, 3), dtype=torch.uint8, device=foreground_img.device)
            b = torch.zeros((1, n, 3), dtype=torch.uint8, device=foreground_img.device)
            f[0, :nf] = foreground_img.reshape(nf, 3)
            b[0, :nb] = background_img.reshape(nb, 3)
            return _transfer(f, b, None, MODE_ALL)[0, :nf].reshape(fs).contiguous()
        f = np.zeros((1, n, 3), np.uint8)
        b = np.zeros((1, n, 3), np.uint8)
        f[0, :nf] = _np_u8(foreground_img, "foreground_img").reshape(nf, 3)
        b[0, :nb] = _np_u8(background_img, "background_img").reshape(nb, 3)
        return np.ascontiguousarray(_transfer(f, b, None, MODE_ALL)[0, :nf].reshape(fs))
    return _transfer(foreground_img, background_img, None, MODE_ALL)


def regional_color_transfer(content, stylized, background_mask, slots=None, out=None, return_status=False):
    """The tail of ``run_localized_style_transfer`` (LST:222-238) in one fused call:
    NEAREST-resize ``stylized`` to the mask size if needed, mask algebra, colour transfer of
    the foreground, final combination.  Accepts one image or a batch (leading dim).
    ``return_status=True`` also returns the per-frame status codes (and prints the reference's
    empty-region warnings on the tensor path too; it synchronises the stream)."""
    return _transfer(content, stylized, background_mask, MODE_MASK, slots=slots, out=out, return_status=return_status)


def regional_color_transfer_binned(content, stylized, background_mask, bins=1024, slots=None, out=None,
                                   mode=MODE_MASK, label=0):
    """The same fused tail with the CDF-matching step (a-6) replaced by ``bins``-bin histogram matching
    (BASELINE.json configs[4] "1024-bin histograms"; ``rct_transfer_binned_async``).  NO COUNTERPART IN
    THE REFERENCE and not the parity path: an approximation of :func:`regional_color_transfer` (measured:
    within 1 uint8 LSB on all but isolated tail pixels (max 3-4 observed); 0.5-5 % of the pixels differ) that is bit-exact against
    the oracle's statement of the same definition.  numpy in -> numpy out, CUDA tensors in -> CUDA tensor out."""
    L = _cabi.lib()
    if not _is_tensor(content):
        dev = _require_cuda()
        c, s = (torch.from_numpy(_np_u8(x, nm)).to(dev) for x, nm in ((content, "content"), (stylized, "stylised")))
        m = torch.from_numpy(_np_u8(background_mask, "mask")).to(dev) if background_mask is not None else None
        if c.dim() == 4 and s.dim() == 3:
            s = s.unsqueeze(0).expand(c.shape[0], -1, -1, -1).contiguous()
        if c.dim() == 4 and m is not None and m.dim() == 2:
            m = m.unsqueeze(0).expand(c.shape[0], -1, -1).contiguous()
        res = regional_color_transfer_binned(c, s, m, bins, slots, None, mode, label).cpu().numpy()
        if out is not None:
            out[...] = res
            return out
        return res
    dev = content.device
    if dev.type != "cuda":
        raise RuntimeError("tensor inputs must live on a CUDA device (no CPU fallback)")
    for t in (content, stylized) + ((background_mask,) if background_mask is not None else ()):
        if t.dtype != torch.uint8 or not t.is_contiguous() or t.device != dev:
            raise TypeError("inputs must be contiguous uint8 tensors on one device")
    n = content.shape[0] if content.dim() == 4 else 1
    H, W = content.shape[-3], content.shape[-2]
    sh, sw = stylized.shape[-3], stylized.shape[-2]
    sty_b = stylized.dim() == 4
    mask_b = background_mask is not None and background_mask.dim() == 3
    plan = get_plan(dev, H, W, sh, sw, slots or min(n, 8))
    if out is None:
        out = torch.empty_like(content) if mode != MODE_LABEL else content.clone()
    with torch.cuda.device(dev), plan.lock, plan.on_current_stream():
        check(L.rct_transfer_binned_async(
            plan.handle, content.data_ptr(), H * W * 3, stylized.data_ptr(), sh * sw * 3 if sty_b else 0,
            background_mask.data_ptr() if background_mask is not None else None, H * W if mask_b else 0,
            out.data_ptr(), H * W * 3, n, mode, label, int(bins), _stream_ptr()))
    return out


def regional_color_transfer_labels(content, stylized, labels, n_labels, slots=None, fused=True, out=None):
    """BASELINE.json config 5 (no counterpart in the reference): for every label l the content
    pixels with that label are matched to the stylised pixels with that label.  numpy in -> numpy out
    (``out``: optional uint8 array to receive the result; a pinned one makes the copy-out a plain DMA)."""
    if not _is_tensor(content):
        # numpy calling convention: one upload, the n_labels passes on the device, one download
        # (per-pass host round trips would move every image 2 * n_labels times over PCIe)
        dev = _require_cuda()
        c, s, l = (torch.from_numpy(_np_u8(x, nm)).to(dev, non_blocking=True)
                   for x, nm in ((content, "content"), (stylized, "stylised"), (labels, "labels")))
        if c.dim() == 4 and s.dim() == 3:
            s = s.unsqueeze(0).expand(c.shape[0], -1, -1, -1).contiguous()
        if c.dim() == 4 and l.dim() == 2:
            l = l.unsqueeze(0).expand(c.shape[0], -1, -1).contiguous()
        res = regional_color_transfer_labels(c, s, l, n_labels, slots=slots, fused=fused)
        if out is None:
            return res.cpu().numpy()
        if out.dtype != np.uint8 or tuple(out.shape) != tuple(res.shape) or not out.flags.c_contiguous:
            raise ValueError("out must be a C-contiguous uint8 array of the content's shape")
        torch.from_numpy(out).copy_(res)  # straight into the caller's memory (DMA when it is pinned)
        return out
    if fused and 0 < n_labels <= 64:
        return _transfer_labels_fused(content, stylized, labels, n_labels, slots)
    out = None
    for lab in range(n_labels):
        out = _transfer(content, stylized, labels, MODE_LABEL, label=lab, out=out, slots=slots, quiet=True)
    return out


_LABEL_WS_BUDGET = 48 << 30  # bytes of workspace a fused label plan may take (of 180 GB)


def _transfer_labels_fused(content, stylized, labels, n_labels, frames_per_group=None):
    """All labels of a group of frames in one launch group: rct_transfer_labels_async with a plan of
    frames_per_group * n_labels slots (slot = frame * n_labels + label)."""
    L = _cabi.lib()
    dev = content.device
    for t in (content, stylized, labels):
        if t.dtype != torch.uint8 or not t.is_contiguous() or t.device != dev or dev.type != "cuda":
            raise TypeError("inputs must be contiguous uint8 tensors on one CUDA device")
    batched = content.dim() == 4
    n = content.shape[0] if batched else 1
    H, W = content.shape[-3], content.shape[-2]
    sh, sw = stylized.shape[-3], stylized.shape[-2]
    sty_b, lab_b = stylized.dim() == 4, labels.dim() == 3
    fpg = max(1, min(n, frames_per_group or 8))
    while fpg > 1 and L.rct_plan_workspace_bytes(H, W, fpg * n_labels, 0) > _LABEL_WS_BUDGET:
        fpg -= 1
    plan = get_plan(dev, H, W, sh, sw, fpg * n_labels)
    out = content.clone()  # pixels of labels >= n_labels keep the content
    with torch.cuda.device(dev), plan.lock, plan.on_current_stream():
        check(L.rct_transfer_labels_async(
            plan.handle, content.data_ptr(), H * W * 3, stylized.data_ptr(), sh * sw * 3 if sty_b else 0,
            labels.data_ptr(), H * W if lab_b else 0, out.data_ptr(), H * W * 3, n, n_labels, _stream_ptr()))
    return out


def extract_foreground_deeplab(content_img, threshold=0.5, model=None):
    """LST:171-188.  The DeepLabV3 network itself is outside the accelerated path (SURVEY.md
    section 2 #3; its weights are a network download the reference performs, LST:173): pass a loaded
    ``torchvision.models.segmentation`` model as ``model=`` and it runs as stock PyTorch on the GPU,
    followed by the fused softmax / threshold kernel (f-2, :mod:`.adjacent`).  Returns the
    reference's uint8 ``[1,H,W]`` background mask."""
    if model is None:
        raise NotImplementedError(
            "DeepLabV3 weights are not available offline: pass model= (a loaded torchvision "
            "deeplabv3_resnet101) or pass background_mask= to run_localized_style_transfer"
        )
    from . import adjacent

    dev = _require_cuda()
    x = torch.from_numpy(np.ascontiguousarray(np.asarray(content_img))).to(dev)
    x = x.permute(2, 0, 1).float().div(255).unsqueeze(0)  # torchvision F.to_tensor for uint8 HWC
    with torch.no_grad():
        logits = model.to(dev).eval()(x)["out"][0]
    return adjacent.deeplab_background_mask(logits, threshold)


def run_localized_style_transfer(
    content_img_path,
    style_img_path,
    output_path="../output",
    file_name="test",
    use_depth=False,
    depth_offset=0.5,
    depth_prominence=20,
    *,
    background_mask=None,
    stylized_image=None,
):
    """LST:191-245 with the two out-of-scope producers injected: ``background_mask`` (uint8
    [1,H,W] or [H,W], what extract_foreground_deeplab returns) and ``stylized_image`` (path or
    uint8 array, what adain_inference writes).  The in-scope tail (LST:218-243) is unchanged:
    resize -> masks -> colour transfer -> combine -> save JPEG (device encoder); returns the saved path."""
    from .jpeg import decode_image

    def _open_rgb(path):
        """LST:200, 219 `Image.open(path).convert("RGB")`: baseline JPEG decoded on the device, PNG on the host inside
        the C library (both bit-exact with Pillow).  A file the decoders refuse with a reason (progressive / CMYK JPEG,
        interlaced or 16-bit PNG, another container) is read by Pillow itself: format handling only, the pixels then
        take the same device path."""
        try:
            return decode_image(path, device=True)
        except ValueError:
            from PIL import Image

            return torch.from_numpy(np.array(Image.open(path).convert("RGB"))).to(_require_cuda())

    content_t = _open_rgb(content_img_path)
    if background_mask is None:
        background_mask = extract_foreground_deeplab(content_t.cpu().numpy())
    if stylized_image is None:
        raise NotImplementedError("AdaIN inference is outside the accelerated path; pass stylized_image=")
    background_mask = np.asarray(background_mask, np.uint8)
    if background_mask.ndim == 3:
        background_mask = background_mask[0]
    if isinstance(stylized_image, str):
        stylized_t = _open_rgb(stylized_image)
    else:
        stylized_t = torch.from_numpy(np.ascontiguousarray(np.asarray(stylized_image, np.uint8))).to(content_t.device)
    mask_t = torch.from_numpy(np.ascontiguousarray(background_mask)).to(content_t.device)
    # decode -> transfer -> encode without the pixels leaving the device
    combined_image, status = regional_color_transfer(content_t.contiguous(), stylized_t.contiguous(), mask_t, return_status=True)
    save_path = f"{output_path}/localized_style_transfer_result.jpg"
    # LST:243 `Image.fromarray(combined_image.astype(np.uint8)).save(save_path)`: the same bytes, encoded on the GPU
    from .jpeg import save_jpeg

    save_jpeg(combined_image, save_path)
    return save_path
