"""rct_match_cdf_static: the CDF match for callers that are not allowed to allocate device memory.

Who that is: a plug-in living inside an application that owns the device heap and exposes no allocator, a service whose
memory pool is sized once at start-up, a real-time loop where the device-wide synchronisation a cudaMalloc / cudaFree
may cause is not acceptable.  They pass inputs and outputs only; the sort's work arrays sit in a fixed 34 MiB array
inside librct.so (one per GPU, shared by every call, limit RCT_MATCH_STATIC_MAX = 2^20 entries per side).

What is checked, through the C ABI (ctypes, raw device pointers):
  * two calls in a row, the second on different and shorter data so that whatever the first left in the shared array
    would show, both equal to rct_match_cdf with caller-provided scratch BIT FOR BIT (the tolerance of every match_cdf
    test in this suite: the path is compare / sort / lerp only) and equal to the CPU oracle;
  * the steady-state call leaves the amount of free device memory unchanged;
  * inputs beyond the limit are refused with a message that names the limit, and nothing is written.
The test allocates its tensors with torch because IT may; the entry point under test never does.

(Named test_zz_no_alloc so that pytest collects it after every other file: `-x` stops at the first failure and a
failure here must not hide the rest of the suite.  It passes alone in a fresh process and after the other tests.)"""
import ctypes

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

LIMIT = 1 << 20  # RCT_MATCH_STATIC_MAX of include/rct.h


def _keys(n, seed, ties):
    rng = np.random.default_rng(seed)
    k = rng.normal(0.0, 3.0, n)
    if ties:  # long tie groups: the ranks inside a group decide the result
        k = np.round(k * 8.0) / 8.0
    return k


def _reference_entry(L, t, s, stream):
    """The existing entry point, scratch from the caller (torch here)."""
    import torch

    m = torch.full((t.numel(),), float("nan"), dtype=torch.float64, device=t.device)
    j = torch.full((t.numel(),), -1, dtype=torch.int64, device=t.device)
    scratch = torch.empty(L.rct_match_scratch_bytes(t.numel(), s.numel()), dtype=torch.uint8, device=t.device)
    rc = L.rct_match_cdf(t.data_ptr(), t.numel(), s.data_ptr(), s.numel(), m.data_ptr(), j.data_ptr(),
                         scratch.data_ptr(), stream)
    assert rc == 0, L.rct_last_error()
    torch.cuda.synchronize()
    return m.cpu().numpy(), j.cpu().numpy()


def test_static_scratch_match_equals_the_allocating_entry_point_twice_in_a_row():
    import torch

    from applied_image_processing_b200 import _cabi
    from oracle import oracle as orc

    L = _cabi.lib()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    # call 1: large, continuous keys, target longer than source; call 2: shorter, heavy ties, source longer
    cases = [(300_000, 170_001, False), (40_003, 90_000, True)]
    free_before_second = None
    for call, (nt, ns, ties) in enumerate(cases):
        t_h, s_h = _keys(nt, 10 + call, ties), _keys(ns, 20 + call, ties) * 1.7 + 0.5
        t = torch.from_numpy(t_h).to(dev)
        s = torch.from_numpy(s_h).to(dev)
        want_m, want_j = _reference_entry(L, t, s, stream)
        m = torch.full((nt,), float("nan"), dtype=torch.float64, device=dev)
        j = torch.full((nt,), -1, dtype=torch.int64, device=dev)
        torch.cuda.synchronize()
        if call == 1:
            free_before_second = torch.cuda.mem_get_info(dev)[0]
        rc = L.rct_match_cdf_static(t.data_ptr(), nt, s.data_ptr(), ns, m.data_ptr(), j.data_ptr(), stream)
        assert rc == 0, L.rct_last_error()
        torch.cuda.synchronize()
        if call == 1:  # set-up happened in call 0: the steady-state call takes no device memory
            assert torch.cuda.mem_get_info(dev)[0] == free_before_second
        got_m, got_j = m.cpu().numpy(), j.cpu().numpy()
        assert np.array_equal(got_j, want_j), f"call {call}: CDF index differs from rct_match_cdf"
        assert np.array_equal(got_m.view(np.uint64), want_m.view(np.uint64)), f"call {call}: matched values differ"
        om, oj = orc.match_cdf(t_h, s_h, want_index=True)
        assert np.array_equal(got_m.view(np.uint64), np.asarray(om).reshape(-1).view(np.uint64)), f"call {call}: vs oracle"
        assert np.array_equal(got_j, np.asarray(oj).reshape(-1))


def test_static_scratch_match_at_the_limit_and_beyond():
    import torch

    from applied_image_processing_b200 import _cabi

    L = _cabi.lib()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    # exactly at the limit on both sides: accepted and right
    t = torch.from_numpy(_keys(LIMIT, 1, False)).to(dev)
    s = torch.from_numpy(_keys(LIMIT, 2, True)).to(dev)
    want_m, want_j = _reference_entry(L, t, s, stream)
    m = torch.empty(LIMIT, dtype=torch.float64, device=dev)
    j = torch.empty(LIMIT, dtype=torch.int64, device=dev)
    assert L.rct_match_cdf_static(t.data_ptr(), LIMIT, s.data_ptr(), LIMIT, m.data_ptr(), j.data_ptr(), stream) == 0, \
        L.rct_last_error()
    torch.cuda.synchronize()
    assert np.array_equal(m.cpu().numpy().view(np.uint64), want_m.view(np.uint64))
    assert np.array_equal(j.cpu().numpy(), want_j)
    # one entry more on either side: refused, the message names the limit, the output is untouched
    m.fill_(-7.0)
    torch.cuda.synchronize()
    for nt, ns in ((LIMIT + 1, 16), (16, LIMIT + 1)):
        rc = L.rct_match_cdf_static(t.data_ptr(), nt, s.data_ptr(), ns, m.data_ptr(), None, stream)
        assert rc == -2  # RCT_ERR_ARG
        msg = L.rct_last_error().decode()
        assert "RCT_MATCH_STATIC_MAX" in msg and str(LIMIT) in msg, msg
    torch.cuda.synchronize()
    assert bool((m == -7.0).all())
    # empty inputs and null pointers are argument errors like in rct_match_cdf
    assert L.rct_match_cdf_static(t.data_ptr(), 0, s.data_ptr(), 8, m.data_ptr(), None, stream) == -2
    assert L.rct_match_cdf_static(None, 8, s.data_ptr(), 8, m.data_ptr(), None, stream) == -2
