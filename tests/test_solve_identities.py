"""Integer / floating-point identities the solve kernels lean on, restated in Python so that they are checked without a
GPU (the kernels themselves are compared with the oracle by the `-m gpu` suite).

* k_lab_moments' fast path (csrc/rct_math.cuh, fast_add / fast_finish): the fixed-point value is taken with a bias of
  2^44 by a magic-number fma, split into a 22-bit and a 24-bit half, accumulated as unsigned 32x32->64 products and the
  bias removed when the partial sums are widened -- must give the same integers as sum(cnt * q_i * q_j) with
  q = rint(lab * 2^40) (C5 of DESIGN.md), and no partial sum may pass 2^64 within 128 terms.
* k_match's np.interp bracket (csrc/rct_sort.cuh, lin_bracket): walking one way from the estimate floor(x * (n - 1)) with
  both ends carried must land on the bracket the two-loop form (and numpy's searchsorted) finds, with the same
  xm = lin(m) and x1 = lin(m + 1).
"""
import random
import struct

import numpy as np


def _bits(x):
    return struct.unpack("<Q", struct.pack("<d", x))[0]


def _fast_moments(entries):
    U, V, A, B, C, w = [0] * 3, [0] * 3, [0] * 6, [0] * 6, [0] * 6, 0
    for lab, cnt in entries:
        qh, ql = [], []
        for x in lab:
            # x * 2^40 is exact, so the single rounding of the addition is the fma's rounding
            y = x * 2.0**40 + (2.0**52 + 2.0**44)
            b = _bits(y)
            hi, lo = b >> 32, b & 0xFFFFFFFF
            assert (hi >> 14) == (0x43300000 >> 14)  # fast_add's range test
            qh.append((b >> 24) & 0x3FFFFF)
            ql.append(lo & 0xFFFFFF)
        u = [cnt * a for a in qh]
        v = [cnt * a for a in ql]
        assert max(u + v) < 2**32  # 32-bit multiplicands
        for i in range(3):
            U[i] += u[i]
            V[i] += v[i]
        k = 0
        for i in range(3):
            for j in range(i, 3):
                A[k] += u[i] * qh[j]
                B[k] += u[i] * ql[j] + v[i] * qh[j]
                C[k] += v[i] * ql[j]
                k += 1
        w += cnt
    assert max(U + V + A + B + C) < 2**64  # no 64-bit accumulator overflows
    s1 = [(U[i] << 24) + V[i] - (w << 44) for i in range(3)]
    s2, k = [], 0
    for i in range(3):
        for j in range(i, 3):
            s2.append((A[k] << 48) + (B[k] << 24) + C[k] - ((s1[i] + s1[j]) << 44) - (w << 88))
            k += 1
    return s1, s2


def _direct_moments(entries):
    s1, s2 = [0] * 3, [0] * 6
    for lab, cnt in entries:
        q = [round(x * 2.0**40) for x in lab]  # Python's round: nearest, ties to even == __double2ll_rn
        for i in range(3):
            s1[i] += q[i] * cnt
        k = 0
        for i in range(3):
            for j in range(i, 3):
                s2[k] += q[i] * q[j] * cnt
                k += 1
    return s1, s2


def test_biased_unsigned_moment_accumulation_is_exact():
    rng = random.Random(5)
    tie = lambda: 0.5 * 2.0**-40 * rng.randint(-9, 9)  # halves of the fixed-point unit: the ties-to-even cases
    for trial in range(120):
        entries = []
        for _ in range(128):  # kFastTerms additions, the most a thread makes between two flushes
            lab = [rng.uniform(-10.5, 0.2), rng.uniform(-5.0, 5.0), rng.choice([rng.uniform(-4.4, 4.4), tie()])]
            entries.append((lab, 127 if trial % 2 else rng.randint(1, 127)))
        assert _fast_moments(entries) == _direct_moments(entries)
    # extremes of a-4's range: log10 clamped at 1e-6 on all three channels, and white
    lo = [-18 / 3**0.5, 0.0, 0.0]
    assert _fast_moments([(lo, 127)] * 128) == _direct_moments([(lo, 127)] * 128)


def _lin(n, k):
    if n == 1:
        return 0.0
    if k == n - 1:
        return 1.0
    return float(np.float64(k) * (np.float64(1.0) / np.float64(n - 1)))


def _bracket_two_loops(n, x):  # the form the oracle uses (C8)
    m = min(max(int(x * float(n - 1)), 0), n - 1)
    while m + 1 <= n - 1 and _lin(n, m + 1) <= x:
        m += 1
    while m > 0 and _lin(n, m) > x:
        m -= 1
    return m


def _bracket_one_way(n, x):  # lin_bracket
    m = min(max(int(x * float(n - 1)), 0), n - 1)
    xm, x1 = _lin(n, m), 1.0
    if xm <= x:
        while m + 1 <= n - 1:
            x1 = _lin(n, m + 1)
            if not x1 <= x:
                break
            m += 1
            xm = x1
    else:
        while m > 0 and xm > x:
            m -= 1
            x1 = xm
            xm = _lin(n, m)
    return m, xm, x1


def test_one_way_bracket_equals_the_two_loop_form_and_numpy():
    rng = np.random.default_rng(11)
    for n in [2, 3, 5, 7, 10, 97, 1000, 4099, 3317760, 8294399, 2**31 - 2]:
        grid_k = np.unique(np.concatenate([rng.integers(0, n, 200), [0, 1, n - 2, n - 1]]).clip(0, n - 1))
        xs = [0.0, 1.0] + list(rng.random(200))
        for k in grid_k:  # the knots themselves and their neighbours one ulp away
            v = _lin(n, int(k))
            xs += [v, float(np.nextafter(v, 0.0)), float(np.nextafter(v, 1.0))]
        # the other grid's knots, as k_match feeds them (lin of a slightly different length)
        for n2 in (n - 1, n + 1, n + 61):
            if n2 >= 2:
                xs += [_lin(n2, int(k)) for k in rng.integers(0, n2, 50)]
        lin_small = np.array([_lin(n, k) for k in range(n)]) if n <= 4099 else None
        for x in xs:
            x = min(max(x, 0.0), 1.0)
            m, xm, x1 = _bracket_one_way(n, x)
            assert m == _bracket_two_loops(n, x)
            assert xm == _lin(n, m)
            if m < n - 1:
                assert x1 == _lin(n, m + 1)
            if lin_small is not None:  # np.interp's own bracket
                assert m == min(int(np.searchsorted(lin_small, x, side="right")) - 1, n - 1)
