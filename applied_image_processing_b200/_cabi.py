"""ctypes binding of librct.so (include/rct.h).  No CPU fallback: a missing library or a
missing CUDA device raises -- the product path never routes around the kernels."""
from __future__ import annotations

import ctypes as C
import os
import threading

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("RCT_LIB") or os.path.join(_HERE, "librct.so")  # RCT_LIB: tuning builds (tools/)

RCT_OK, RCT_EMPTY_TARGET, RCT_EMPTY_SOURCE = 0, 1, 2
MODE_MASK, MODE_LABEL, MODE_ALL = 0, 1, 2
FLAG_DEBUG = 1
OPT_AGGREGATE = 0
OPT_TIMING = 1
OPT_PART_CAP = 2
OPT_FRONTEND = 3
OPT_APPLY = 4
OPT_MATCH = 5
OPT_STAGE_SETS = 6
OPT_LUT = 7
OPT_KEYTAB = 8
JPEG_SERIAL = 1
TK_COUNT = 20

_lib = None
_lock = threading.Lock()


class FrameInfo(C.Structure):
    _fields_ = [
        ("n_target", C.c_int64), ("n_source", C.c_int64),
        ("d_target", C.c_int64), ("d_source", C.c_int64),
        ("t_mean", C.c_double * 3), ("t_comp", C.c_double * 3),
        ("s_mean", C.c_double * 3), ("s_comp", C.c_double * 3),
        ("status", C.c_int),
    ]


_vp, _sz, _i, _u, _i64 = C.c_void_p, C.c_size_t, C.c_int, C.c_uint, C.c_int64

# name -> (restype, argtypes); every symbol include/rct.h declares
SIGNATURES = {
    "rct_version": (_i, []),
    "rct_last_error": (C.c_char_p, []),
    "rct_set_matrices": (_i, [_vp]),
    "rct_get_matrices": (_i, [_vp]),
    "rct_resample_table": (_i, [_i, _i, _vp]),
    "rct_plan_workspace_bytes": (_sz, [_i, _i, _i, _u]),
    "rct_plan_create": (_i, [_i, _i, _i, _i, _i, _i, _u, _vp, _sz, C.POINTER(_vp)]),
    "rct_plan_destroy": (_i, [_vp]),
    "rct_plan_set_option": (_i, [_vp, _i, _i]),
    "rct_plan_timing": (_i, [_vp, _vp, _vp, _i]),
    "rct_plan_kernel_timing": (_i, [_vp, _vp, _vp, _i]),
    "rct_kernel_timing_name": (C.c_char_p, [_i]),
    "rct_transfer_async": (_i, [_vp, _vp, _sz, _vp, _sz, _vp, _sz, _vp, _sz, _i, _i, _i, _vp]),
    "rct_plan_status": (_i, [_vp, _vp, _i, _vp]),
    "rct_transfer_labels_async": (_i, [_vp, _vp, _sz, _vp, _sz, _vp, _sz, _vp, _sz, _i, _i, _vp]),
    "rct_transfer_host": (_i, [_vp, _vp, _vp, _vp, _vp, _i, _i, _i, _vp]),
    "rct_shard_hist_async": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _i, _vp]),
    "rct_plan_hist_ptr": (_vp, [_vp]),
    "rct_plan_hist_bytes": (_sz, [_vp]),
    "rct_shard_solve_async": (_i, [_vp, _vp]),
    "rct_shard_apply_async": (_i, [_vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _vp]),
    "rct_transfer_binned_async": (_i, [_vp, _vp, _sz, _vp, _sz, _vp, _sz, _vp, _sz, _i, _i, _i, _i, _vp]),
    "rct_shard_binned_phase_async": (_i, [_vp, _i, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _vp]),
    "rct_plan_xchg_ptr": (_vp, [_vp]),
    "rct_plan_xchg_bytes": (_sz, [_i, _i]),
    "rct_plan_debug_binned": (_i, [_vp, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp]),
    "rct_rgb_to_lab": (_i, [_vp, _vp, _sz, _vp]),
    "rct_lab_to_rgb": (_i, [_vp, _vp, _vp, _sz, _vp]),
    "rct_pca_scratch_bytes": (_sz, []),
    "rct_apply_pca": (_i, [_vp, _sz, _vp, _vp, _vp, _vp, _vp]),
    "rct_pca_inverse": (_i, [_vp, _sz, _vp, _vp, _vp, _vp]),
    "rct_match_scratch_bytes": (_sz, [_sz, _sz]),
    "rct_match_cdf": (_i, [_vp, _sz, _vp, _sz, _vp, _vp, _vp, _vp]),
    "rct_match_cdf_static": (_i, [_vp, _sz, _vp, _sz, _vp, _vp, _vp]),
    "rct_adain_mask_blend": (_i, [_vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _i, _sz, _vp]),
    "rct_softmax_threshold_mask": (_i, [_vp, _i, _sz, _i, C.c_float, _vp, _vp]),
    "rct_argmax_labels": (_i, [_vp, _i, _sz, _vp, _vp]),
    "rct_warp_image": (_i, [_vp, _vp, _vp, _i, _i, _i, _i, _vp]),
    "rct_blend_images": (_i, [_vp, _vp, C.c_double, _vp, _sz, _vp]),
    "rct_warp_blend": (_i, [_vp, _vp, _vp, C.c_double, _vp, _i, _i, _i, _i, _vp]),
    "rct_jpeg_workspace_bytes": (_sz, [_i, _i, _i]),
    "rct_jpeg_max_bytes": (_sz, [_i, _i]),
    "rct_jpeg_encode_async": (_i, [_vp, _sz, _i, _i, _i, _i, _vp, _vp, _sz, _sz, _vp, _vp]),
    "rct_jpeg_decode_info": (_i, [_vp, _sz, C.POINTER(_i), C.POINTER(_i), C.POINTER(_i), C.POINTER(_sz)]),
    "rct_jpeg_decode_async": (_i, [_vp, _sz, _vp, _sz, _vp, _i, _vp]),
    "rct_files_out_capacity": (_sz, [_i, _i]),
    "rct_transfer_files_host": (_i, [_vp, _vp, _sz, _vp, _sz, _vp, _i, _vp, _sz, C.POINTER(_sz), C.POINTER(_i)]),
    "rct_png_info": (_i, [_vp, _sz, C.POINTER(_i), C.POINTER(_i)]),
    "rct_png_decode_host": (_i, [_vp, _sz, _vp, _sz]),
    "rct_plan_info": (_i, [_vp, _i, C.POINTER(FrameInfo)]),
    "rct_plan_debug_hist": (_i, [_vp, _i, _vp, _vp]),
    "rct_plan_debug_target": (_i, [_vp, _i, _vp, _vp, _vp, _vp, _vp]),
    "rct_plan_debug_lut": (_i, [_vp, _i, _vp]),
    "rct_plan_debug_keytab": (_i, [_vp, _i, _vp, _vp]),
    "rct_launch_count": (C.c_longlong, [_i]),
}


def lib():
    """Load librct.so (built in-tree by __graft_entry__.build() / csrc/Makefile)."""
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is None:
            if not os.path.exists(LIB_PATH):
                raise RuntimeError(
                    f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; "
                    "g.build()'` (or make -C applied_image_processing_b200/csrc). "
                    "There is no CPU fallback for the colour-transfer path."
                )
            L = C.CDLL(LIB_PATH)
            for name, (res, args) in SIGNATURES.items():
                fn = getattr(L, name)
                fn.restype = res
                fn.argtypes = args
            _lib = L
    return _lib


class RctError(RuntimeError):
    pass


def check(rc: int) -> int:
    """Negative return codes raise with the library's own message; 0/1/2 pass through."""
    if rc < 0:
        msg = lib().rct_last_error()
        raise RctError(f"librct error {rc}: {msg.decode() if msg else '?'}")
    return rc
