"""kg_chi_points_noalloc (include/kgcuda.h): the point kernel with its staging buffers in a fixed device array instead of device
allocations -- no counterpart in the reference, added for callers that cannot call cudaMalloc.
Called through the C-ABI (ctypes on libkgcuda.so) and checked against kg_chi_points on the same inputs: same kernel, same arithmetic,
so the results are required to be bit-identical (the tolerance of the other chi tests, 1e-9, is the documented bound). Two calls in a
row with different sizes show that the shared array's old contents do not leak into the next call. Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _p(a):
    return None if a is None else a.ctypes.data_as(C.POINTER(C.c_double))


def _both(model, pair, T, g, b, coeff=None):
    L = model.L
    out = []
    for fn in (L.kg_chi_points, L.kg_chi_points_noalloc):
        chi, R = np.full(g.size, -7.0), np.full(g.size, -7.0)
        st = np.full(g.size, 12345, np.int32)
        rc = fn(model.h, pair, float(T), g.size, _p(g), _p(b), _p(coeff), _p(chi), _p(R), st.ctypes.data_as(C.POINTER(C.c_int32)))
        assert rc == 0, L.kg_last_error(model.h)
        out.append((chi, R, st))
    return out


def test_noalloc_matches_chi_points_two_calls_in_a_row():
    from kineticgas_b200 import DeviceModel, mie_desc
    model = DeviceModel([mie_desc(3.405e-10, 118.0 * 1.38064852e-23, 6.0, 12.0)], [6.63e-26], True)
    rng = np.random.default_rng(7)
    # first call: 20000 points incl. the short-cuts b < 1e-3 and b > 100; second call: fewer points, other T (stale contents beyond n
    # and stale status words must not matter)
    for n, T in ((20000, 300.0), (777, 45.0)):
        g = np.ascontiguousarray(rng.uniform(0.05, 4.0, n))
        b = np.ascontiguousarray(rng.uniform(0.0, 4.5, n))
        b[:5] = [0.0, 5e-4, 150.0, 1.0, 2.5]
        (chi0, R0, st0), (chi1, R1, st1) = _both(model, 0, T, g, b)
        assert np.array_equal(st0, st1) and not st1.any()
        far = b > 100                                                      # chi = 0 there and R is not evaluated (NaN) in either entry point
        assert np.all(np.isfinite(chi1)) and np.all(np.isfinite(R1[~far])) and np.array_equal(np.isnan(R0), np.isnan(R1))
        assert np.max(np.abs(chi1 - chi0)) <= 1e-9 * max(1.0, np.max(np.abs(chi0)))
        assert np.max(np.abs(R1[~far] / R0[~far] - 1.0)) <= 1e-9
        assert np.array_equal(chi0, chi1) and np.array_equal(R0, R1, equal_nan=True)       # same kernel on the same inputs: identical bits


def test_noalloc_with_effective_coefficients_and_limits():
    from kineticgas_b200 import DeviceModel, mie_desc
    model = DeviceModel([mie_desc(2.96e-10, 33.0 * 1.38064852e-23, 6.0, 9.0)], [3.3e-27], True)
    L = model.L
    nmax = L.kg_chi_points_noalloc_max()
    assert nmax == 65536
    rng = np.random.default_rng(11)
    g = np.ascontiguousarray(rng.uniform(0.1, 3.0, 4096)); b = np.ascontiguousarray(rng.uniform(0.01, 3.0, 4096))
    coeff = np.array([2.7, -2.4, 0, 0, 0, 0], float)               # a call-specific coefficient block goes through the array as well
    (chi0, R0, st0), (chi1, R1, st1) = _both(model, 0, 120.0, g, b, coeff)
    assert np.array_equal(chi0, chi1) and np.array_equal(R0, R1) and np.array_equal(st0, st1)
    # and without coefficients right after: the block of the previous call is not picked up
    (chi2, _, _), (chi3, _, _) = _both(model, 0, 120.0, g, b)
    assert np.array_equal(chi2, chi3) and not np.array_equal(chi2, chi0)
    # beyond the fixed maximum: refused with a message, nothing launched
    big = np.ones(nmax + 1)
    out = np.empty(nmax + 1)
    rc = L.kg_chi_points_noalloc(model.h, 0, 300.0, big.size, _p(big), _p(big), None, _p(out), None, None)
    assert rc == -1
    assert b"exceeds the fixed scratch" in L.kg_last_error(model.h)
    # exactly the maximum works
    gm = np.ascontiguousarray(rng.uniform(0.1, 3.0, nmax)); bm = np.ascontiguousarray(rng.uniform(0.01, 3.0, nmax))
    (c0, r0, s0), (c1, r1, s1) = _both(model, 0, 300.0, gm, bm)
    assert np.array_equal(c0, c1) and np.array_equal(r0, r1)
