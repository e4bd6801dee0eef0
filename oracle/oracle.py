"""TEST INFRASTRUCTURE ONLY -- ctypes front-end of the C oracle (oracle/rct_oracle.c).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
may import this module.  The product package (applied_image_processing_b200) never does.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

MODE_MASK, MODE_LABEL, MODE_ALL = 0, 1, 2


def reference_matrices() -> np.ndarray:
    """The four Reinhard matrices, computed with numpy exactly as LST:12-19 does."""
    rgb_to_lms = np.array(
        [[0.3811, 0.5783, 0.0402], [0.1967, 0.7244, 0.0782], [0.0241, 0.1288, 0.8444]]
    )
    lms_to_lab = np.array(
        [[1 / np.sqrt(3), 0, 0], [0, 1 / np.sqrt(6), 0], [0, 0, 1 / np.sqrt(2)]]
    ) @ np.array([[1, 1, 1], [1, 1, -2], [1, -1, 0]])
    lab_to_lms = np.linalg.inv(lms_to_lab)
    lms_to_rgb = np.linalg.inv(rgb_to_lms)
    return np.ascontiguousarray(
        np.stack([rgb_to_lms, lms_to_lab, lab_to_lms, lms_to_rgb]).astype(np.float64)
    )


def build(force: bool = False) -> str:
    so = os.path.join(_HERE, "liborc.so")
    src = os.path.join(_HERE, "rct_oracle.c")
    if force or not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-B", "liborc.so"], stdout=subprocess.DEVNULL)
    return so


class _Info(C.Structure):
    _fields_ = [
        ("n_t", C.c_int64), ("n_s", C.c_int64),
        ("t_mean", C.c_double * 3), ("t_comp", C.c_double * 3),
        ("s_mean", C.c_double * 3), ("s_comp", C.c_double * 3),
    ]


def lib():
    global _LIB
    if _LIB is None:
        _LIB = C.CDLL(build())
        _LIB.orc_log10.restype = C.c_double
        _LIB.orc_log10.argtypes = [C.c_double]
        _LIB.orc_pow10.restype = C.c_double
        _LIB.orc_pow10.argtypes = [C.c_double]
        _LIB.orc_transfer.restype = C.c_int
        _LIB.orc_transfer_binned.restype = C.c_int
        _LIB.orc_match_cdf.restype = C.c_int
        _LIB.orc_apply_pca.restype = C.c_int
        m = reference_matrices()
        _LIB.orc_set_matrices(m.ctypes.data_as(C.c_void_p))
    return _LIB


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def resample_table(in_size: int, out_size: int) -> np.ndarray:
    tab = np.empty(out_size, np.int32)
    lib().orc_resample_table(C.c_int(in_size), C.c_int(out_size), _p(tab))
    return tab


def rgb_to_lab_pixels(pixels_uint8: np.ndarray) -> np.ndarray:
    px = np.ascontiguousarray(pixels_uint8, np.uint8).reshape(-1, 3)
    lab = np.empty((px.shape[0], 3), np.float64)
    lib().orc_rgb_to_lab(_p(px), C.c_size_t(px.shape[0]), _p(lab))
    return lab


def lab_to_rgb_pixels(lab: np.ndarray, want_float: bool = False):
    lab = np.ascontiguousarray(lab, np.float64).reshape(-1, 3)
    out = np.empty((lab.shape[0], 3), np.uint8)
    rgbf = np.empty((lab.shape[0], 3), np.float64) if want_float else None
    lib().orc_lab_to_rgb(_p(lab), C.c_size_t(lab.shape[0]), _p(out), _p(rgbf))
    return (out, rgbf) if want_float else out


def apply_pca(lab: np.ndarray):
    lab = np.ascontiguousarray(lab, np.float64).reshape(-1, 3)
    proj = np.empty((lab.shape[0], 1), np.float64)
    mean = np.empty(3, np.float64)
    comp = np.empty(3, np.float64)
    st = lib().orc_apply_pca(_p(lab), C.c_size_t(lab.shape[0]), _p(proj), _p(mean), _p(comp))
    if st:
        raise ValueError("empty input")
    return proj, mean, comp


def match_cdf(target_proj: np.ndarray, source_proj: np.ndarray, want_index: bool = False):
    t = np.ascontiguousarray(target_proj, np.float64).ravel()
    s = np.ascontiguousarray(source_proj, np.float64).ravel()
    matched = np.empty(t.shape[0], np.float64)
    j = np.empty(t.shape[0], np.int64)
    st = lib().orc_match_cdf(_p(t), C.c_int64(t.shape[0]), _p(s), C.c_int64(s.shape[0]),
                             _p(matched), _p(j))
    if st:
        raise ValueError("empty input")
    return (matched.reshape(-1, 1), j) if want_index else matched.reshape(-1, 1)


def transfer(content, stylised, mask=None, mode=MODE_MASK, label=0, out=None, debug=False):
    """Run the oracle pipeline.  Returns (out, status, info dict)."""
    content = np.ascontiguousarray(content, np.uint8)
    stylised = np.ascontiguousarray(stylised, np.uint8)
    if mode == MODE_ALL and mask is None and content.shape[:2] != stylised.shape[:2] and out is None and not debug:
        # color_transfer_foreground(fg, bg) with different shapes (LST:134-139): the reference gathers
        # all non-black pixels of each image and never resizes.  Black pixels are excluded from everything,
        # so padding both pixel lists with black to one length restates exactly that.
        nf, nb = content.shape[0] * content.shape[1], stylised.shape[0] * stylised.shape[1]
        n = max(nf, nb)
        f = np.zeros((1, n, 3), np.uint8)
        b = np.zeros((1, n, 3), np.uint8)
        f[0, :nf] = content.reshape(nf, 3)
        b[0, :nb] = stylised.reshape(nb, 3)
        o, st, d = transfer(f, b, None, mode=MODE_ALL)
        return np.ascontiguousarray(o[0, :nf].reshape(content.shape)), st, d
    H, W = content.shape[:2]
    sh, sw = stylised.shape[:2]
    if mask is not None:
        mask = np.ascontiguousarray(mask, np.uint8)
        assert mask.shape == (H, W)
    elif mode != MODE_ALL:
        raise ValueError("mask required")
    if out is None:
        out = np.zeros((H, W, 3), np.uint8)
    info = _Info()
    d = {}
    if debug:
        d["key"] = np.empty((H, W), np.float64)
        d["j"] = np.empty((H, W), np.int64)
        d["matched"] = np.empty((H, W), np.float64)
        d["hist_t"] = np.empty(1 << 24, np.uint32)
        d["hist_s"] = np.empty(1 << 24, np.uint32)
        d["rgbf"] = np.full((H, W, 3), np.nan, np.float64)
    st = lib().orc_transfer(
        _p(content), C.c_int(H), C.c_int(W), _p(stylised), C.c_int(sh), C.c_int(sw), _p(mask),
        C.c_int(mode), C.c_int(label), _p(out), C.byref(info),
        _p(d.get("key")), _p(d.get("j")), _p(d.get("matched")), _p(d.get("hist_t")),
        _p(d.get("hist_s")), _p(d.get("rgbf")),
    )
    d.update(
        n_t=info.n_t, n_s=info.n_s,
        t_mean=np.array(info.t_mean), t_comp=np.array(info.t_comp),
        s_mean=np.array(info.s_mean), s_comp=np.array(info.s_comp),
    )
    return out, st, d


def transfer_binned(content, stylised, mask=None, bins=1024, mode=MODE_MASK, label=0, out=None, debug=False):
    """The pipeline with the BINNED CDF match (orc_match_binned in rct_oracle.c): an approximation
    with no counterpart in the reference, whose definition this oracle and the CUDA path share.
    Returns (out, status, info); debug adds key / code (= b * 65536 + b') / matched per pixel and
    the two `bins`-bin histograms + key ranges."""
    content = np.ascontiguousarray(content, np.uint8)
    stylised = np.ascontiguousarray(stylised, np.uint8)
    H, W = content.shape[:2]
    sh, sw = stylised.shape[:2]
    if mask is not None:
        mask = np.ascontiguousarray(mask, np.uint8)
        assert mask.shape == (H, W)
    elif mode != MODE_ALL:
        raise ValueError("mask required")
    if out is None:
        out = np.zeros((H, W, 3), np.uint8)
    info = _Info()
    d = {"hist_t": np.zeros(bins, np.uint32), "hist_s": np.zeros(bins, np.uint32), "range": np.zeros(4, np.float64)}
    if debug:
        d["key"] = np.empty((H, W), np.float64)
        d["code"] = np.empty((H, W), np.int64)
        d["matched"] = np.empty((H, W), np.float64)
    st = lib().orc_transfer_binned(
        _p(content), C.c_int(H), C.c_int(W), _p(stylised), C.c_int(sh), C.c_int(sw), _p(mask),
        C.c_int(mode), C.c_int(label), C.c_int(bins), _p(out), C.byref(info),
        _p(d.get("key")), _p(d.get("code")), _p(d.get("matched")), _p(d["hist_t"]), _p(d["hist_s"]), _p(d["range"]),
    )
    d.update(
        n_t=info.n_t, n_s=info.n_s,
        t_mean=np.array(info.t_mean), t_comp=np.array(info.t_comp),
        s_mean=np.array(info.s_mean), s_comp=np.array(info.s_comp),
    )
    return out, st, d
