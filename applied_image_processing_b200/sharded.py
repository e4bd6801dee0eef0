"""One very large image row-sharded over the ranks of a ``torch.distributed`` group
(BASELINE.json configs[3]: one 15360x8640 image over 2/4/8 B200).

Protocol (identical result to the unsharded call, bit for bit, for any number of ranks):

1. every rank counts the colours of ITS rows into the two colour-cube histograms
   (``rct_shard_hist_async``: target = foreground of content, source = background of stylised);
2. one ``all_reduce(SUM)`` over the 2 x 2^24 uint32 histograms -- integer sums are independent of
   order and of the number of ranks (NCCL over NVLink/NVSwitch; ``gloo`` through a host copy
   for the CPU-side tests);
3. every rank derives the same statistics, PCA, exact CDF match and colour LUT from the summed
   histograms (``rct_shard_solve_async``; redundant, no second exchange);
4. every rank writes its own rows of the output (``rct_shard_apply_async``).

A rank holds only its band: rows ``[row0, row0+rows)`` of content / mask / output and the rows
of the stylised image its NEAREST row table refers to (``stylised_rows_needed``).  The C ABI
takes pointers to the FULL images, so band tensors are passed as ``band.data_ptr() - offset``.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch
import torch.distributed as dist

from . import _cabi
from ._cabi import MODE_MASK, check
from . import localized_style_transfer as lst


def row_band(H: int, rank: int, world: int) -> tuple[int, int]:
    """Contiguous band of rank ``rank``: (row0, rows); the remainder goes to the first ranks."""
    if not (0 <= rank < world):
        raise ValueError("rank out of range")
    base, rem = divmod(H, world)
    rows = base + (1 if rank < rem else 0)
    row0 = rank * base + min(rank, rem)
    return row0, rows


def resample_table(in_size: int, out_size: int) -> np.ndarray:
    """Pillow NEAREST source-index table (LST:223-230) through the C ABI (host function)."""
    tab = np.empty(out_size, np.int32)
    check(_cabi.lib().rct_resample_table(in_size, out_size, tab.ctypes.data))
    return tab


def stylised_rows_needed(sh: int, H: int, row0: int, rows: int) -> tuple[int, int]:
    """Rows [s0, s1) of the sh-row stylised image that output rows [row0, row0+rows) read."""
    if rows <= 0:
        return 0, 0
    if sh == H:
        return row0, row0 + rows
    tab = resample_table(sh, H)[row0:row0 + rows]
    return int(tab.min()), int(tab.max()) + 1


def _hist_view(plan) -> torch.Tensor:
    """The plan's histogram buffer as an int32 tensor (no copy).  uint32 sums wrap exactly like
    int32 sums, so SUM over int32 is the uint32 sum bit for bit."""
    L = _cabi.lib()
    ptr = L.rct_plan_hist_ptr(plan.handle)
    n = L.rct_plan_hist_bytes(plan.handle) // 4

    class _Buf:
        __cuda_array_interface__ = {"shape": (n,), "typestr": "<i4", "data": (ptr, False), "version": 3}

    return torch.as_tensor(_Buf(), device=plan.device)


class ShardedTransfer:
    """Row-sharded ``regional_color_transfer`` for one geometry; one instance per rank."""

    def __init__(self, H, W, sh, sw, device=None, group=None):
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.device = device if device is not None else torch.device("cuda", torch.cuda.current_device())
        self.H, self.W, self.sh, self.sw = H, W, sh, sw
        self.row0, self.rows = row_band(H, self.rank, self.world)
        self.s0, self.s1 = stylised_rows_needed(sh, H, self.row0, self.rows)
        self.plan = lst.Plan(self.device, H, W, sh, sw, slots=1)
        self.hist = _hist_view(self.plan)
        backend = dist.get_backend(group) if dist.is_initialized() else None
        self._via_host = backend == "gloo"
        self._ev = [torch.cuda.Event(enable_timing=True) for _ in range(5)]
        self.last_ms = None
        self.host_s = {}   # host seconds spent ENQUEUEING each phase / collective (perf_counter; the calls are async)

    def _host(self, name, t0):
        import time

        self.host_s[name] = self.host_s.get(name, 0.0) + (time.perf_counter() - t0)

    def allreduce(self):
        if self.world == 1:
            return
        if self._via_host:  # CPU-side tests: gloo has no device transport here
            h = self.hist.cpu()
            dist.all_reduce(h, op=dist.ReduceOp.SUM, group=self.group)
            self.hist.copy_(h)
        else:
            dist.all_reduce(self.hist, op=dist.ReduceOp.SUM, group=self.group)

    def __call__(self, content_band, stylised_band, mask_band, out_band=None, mode=MODE_MASK, label=0,
                 timed=False):
        """content_band / mask_band / out_band: this rank's rows ([rows,W,3] / [rows,W]) as CUDA
        uint8 tensors; stylised_band: rows [s0, s1) of the stylised image ([s1-s0, sw, 3])."""
        L = _cabi.lib()
        W, sw = self.W, self.sw
        for t, shape in ((content_band, (self.rows, W, 3)), (stylised_band, (self.s1 - self.s0, sw, 3)),
                         (mask_band, (self.rows, W))):
            if t is None and shape == (self.rows, W):
                continue  # MODE_ALL has no mask: NULL goes to the library, which refuses MASK / LABEL calls without one
            if tuple(t.shape) != shape or t.dtype != torch.uint8 or not t.is_contiguous() or not t.is_cuda:
                raise TypeError(f"expected a contiguous CUDA uint8 tensor of shape {shape}, got {tuple(t.shape)}")
        if out_band is None:
            out_band = torch.empty_like(content_band)
        # virtual full-image base pointers (only the band's rows are ever dereferenced)
        c_ptr = content_band.data_ptr() - self.row0 * W * 3
        m_ptr = None if mask_band is None else mask_band.data_ptr() - self.row0 * W
        o_ptr = out_band.data_ptr() - self.row0 * W * 3
        s_ptr = stylised_band.data_ptr() - self.s0 * sw * 3
        st = torch.cuda.current_stream()
        sp = st.cuda_stream
        ev = self._ev
        with torch.cuda.device(self.device), self.plan.lock:
            if timed:
                ev[0].record(st)
            import time

            t0 = time.perf_counter()
            check(L.rct_shard_hist_async(self.plan.handle, c_ptr, s_ptr, m_ptr, mode, label, self.row0,
                                         self.rows, sp))
            self._host("hist", t0)
            if timed:
                ev[1].record(st)
            t0 = time.perf_counter()
            self.allreduce()
            self._host("allreduce", t0)
            if timed:
                ev[2].record(st)
            t0 = time.perf_counter()
            check(L.rct_shard_solve_async(self.plan.handle, sp))
            self._host("solve", t0)
            if timed:
                ev[3].record(st)
            t0 = time.perf_counter()
            check(L.rct_shard_apply_async(self.plan.handle, c_ptr, s_ptr, m_ptr, o_ptr, mode, label,
                                          self.row0, self.rows, sp))
            self._host("apply", t0)
            if timed:
                ev[4].record(st)
        self._timed = timed
        return out_band

    # ---- binned mode: the only cross-GPU state is 56 integers, 4 doubles and 2 x bins counters ----------
    def _xchg(self, which, bins):
        cache = self.__dict__.setdefault("_xchg_views", {})
        if (which, bins) not in cache:
            cache[(which, bins)] = self._xchg_view(which, bins)
        return cache[(which, bins)]

    def _xchg_view(self, which, bins):
        L = _cabi.lib()
        ptr = L.rct_plan_xchg_ptr(self.plan.handle)
        nbytes = L.rct_plan_xchg_bytes(which, bins)
        n, typestr = {0: (nbytes // 8, "<i8"), 1: (nbytes // 8, "<f8"), 2: (nbytes // 4, "<i4")}[which]

        class _Buf:
            __cuda_array_interface__ = {"shape": (n,), "typestr": typestr, "data": (ptr, False), "version": 3}

        return torch.as_tensor(_Buf(), device=self.plan.device)

    def _allreduce_small(self, t, op):
        if self.world == 1:
            return
        if self._via_host:
            h = t.cpu()
            dist.all_reduce(h, op=op, group=self.group)
            t.copy_(h)
        else:
            dist.all_reduce(t, op=op, group=self.group)

    def binned(self, content_band, stylised_band, mask_band, out_band=None, bins=1024, mode=MODE_MASK, label=0,
               timed=False):
        """Row-sharded BINNED transfer (``rct_shard_binned_phase_async``): three tiny all-reduces (exact integer
        moments, key ranges, key histograms) instead of the 128 MiB colour cubes, no redundant solve -- every rank
        builds the LUT of its own band's colours only.  Same band conventions as ``__call__``.  An approximation of
        the exact path (see ``regional_color_transfer_binned``), bit-identical for any number of ranks."""
        L = _cabi.lib()
        W, sw = self.W, self.sw
        if out_band is None:
            out_band = torch.empty_like(content_band)
        c_ptr = content_band.data_ptr() - self.row0 * W * 3
        m_ptr = None if mask_band is None else mask_band.data_ptr() - self.row0 * W
        o_ptr = out_band.data_ptr() - self.row0 * W * 3
        s_ptr = stylised_band.data_ptr() - self.s0 * sw * 3
        st = torch.cuda.current_stream()
        sp = st.cuda_stream
        ops = (dist.ReduceOp.SUM, dist.ReduceOp.MAX, dist.ReduceOp.SUM)
        if not hasattr(self, "_evb"):
            self._evb = [torch.cuda.Event(enable_timing=True) for _ in range(8)]
        ev = self._evb
        with torch.cuda.device(self.device), self.plan.lock:
            for phase in range(4):
                if timed:
                    ev[2 * phase].record(st)
                import time

                t0 = time.perf_counter()
                check(L.rct_shard_binned_phase_async(self.plan.handle, phase, c_ptr, s_ptr, m_ptr, o_ptr, mode, label,
                                                     self.row0, self.rows, int(bins), sp))
                self._host(f"phase{phase}", t0)
                if timed:
                    ev[2 * phase + 1].record(st)
                if phase < 3:
                    t0 = time.perf_counter()
                    self._allreduce_small(self._xchg(phase, bins), ops[phase])
                    self._host("allreduce", t0)
        return out_band

    def binned_phase_ms(self):
        """Device time of the last timed binned call: compute of the four phases and the three exchanges (ms)."""
        ev = self._evb
        ev[7].synchronize()
        comp = [ev[2 * k].elapsed_time(ev[2 * k + 1]) for k in range(4)]
        comm = [ev[2 * k + 1].elapsed_time(ev[2 * k + 2]) for k in range(3)]
        return {"front_moments_ms": comp[0], "keys_ms": comp[1], "hist_ms": comp[2], "match_apply_ms": comp[3],
                "allreduce_ms": sum(comm)}

    def phase_ms(self):
        """Device time of the last timed call: hist / all-reduce / solve / apply (ms)."""
        self._ev[4].synchronize()
        names = ("hist_ms", "allreduce_ms", "solve_ms", "apply_ms")
        return {n: self._ev[i].elapsed_time(self._ev[i + 1]) for i, n in enumerate(names)}

    def status(self) -> int:
        s = np.zeros(1, np.int32)
        check(_cabi.lib().rct_plan_status(self.plan.handle, s.ctypes.data, 1,
                                          torch.cuda.current_stream().cuda_stream))
        return int(s[0])


def regional_color_transfer_sharded(content, stylized, background_mask, group=None, bins=0):
    """Convenience wrapper for callers that hold the FULL numpy images on every rank: slices
    this rank's band, runs the sharded protocol and returns this rank's output rows
    (numpy, [rows, W, 3]) together with (row0, rows).  bins > 0 selects the binned mode."""
    content = np.ascontiguousarray(content, np.uint8)
    stylized = np.ascontiguousarray(stylized, np.uint8)
    mask = np.ascontiguousarray(background_mask, np.uint8)
    H, W = mask.shape
    sh, sw = stylized.shape[:2]
    dev = torch.device("cuda", torch.cuda.current_device())
    t = ShardedTransfer(H, W, sh, sw, dev, group)
    r0, r1 = t.row0, t.row0 + t.rows
    args = (torch.from_numpy(content[r0:r1]).to(dev), torch.from_numpy(stylized[t.s0:t.s1]).to(dev),
            torch.from_numpy(mask[r0:r1]).to(dev))
    out = t.binned(*args, bins=bins) if bins else t(*args)
    st = t.status()
    lst._warn(st)
    return out.cpu().numpy(), (t.row0, t.rows)
