"""Pins the C oracle (oracle/kg_oracle.c) against (1) golden vectors generated from the unmodified reference build
(tests/golden/chi_w_reference.npz, oracle/make_golden.py) and (2) the reference build itself (oracle/_ref) when present."""
import ctypes as C
import os
import numpy as np
import pytest

from kineticgas_b200.fluids import mie_parameters, BOLTZMANN, AVOGADRO
from oracle import port, ref

GOLD = np.load(os.path.join(os.path.dirname(__file__), "golden", "chi_w_reference.npz"))


def _pair(comp):
    ids, Mw, s, e, la, lr = mie_parameters(comp)
    return port.mie_pair(s[0], e[0] * BOLTZMANN, la[0], lr[0]), s[0], Mw[0] * 1e-3 / AVOGADRO


@pytest.mark.parametrize("key", sorted(k[:-4] for k in GOLD.files if k.endswith("_chi")))
def test_chi_R_golden(key):
    comp, T = key.split("_")[0], float(key.split("_")[1])
    p, s, _ = _pair(comp)
    g, b = GOLD[key + "_g"], GOLD[key + "_b"]
    chi = port.chi(p, T, g, b * s)
    R = port.get_R(p, T, g, b * s) / s
    # same compiler, same libm, same operation order: bit-exact
    assert np.array_equal(chi, GOLD[key + "_chi"])
    assert np.array_equal(R, GOLD[key + "_R"])


@pytest.mark.parametrize("key", [k for k in GOLD.files if k.startswith("W_")])
def test_w_golden(key):
    _, comp, T, l, r = key.split("_")
    p, s, m = _pair(comp)
    w = port.w_integral(p, float(T), int(l), int(r))
    om = w * port.lib().kgo_omega_prefactor(s, float(T), m, m, 1)
    assert w == GOLD[key][0]
    assert abs(om / GOLD[key][1] - 1) < 1e-15


def test_lattice_regimes():
    # cpp/Spherical.cpp:60-84
    expect = {300.0: (8, 5, 0.5, 1 / 32, 4, 16), 60.0: (6, 3.5, 0.125, 1 / 32, 4, 16), 40.0: (6, 3.5, 0.125, 1 / 32, 4, 8),
              30.0: (3.5, 3.5, 1 / 16, 1 / 64, 8, 8), 10.0: (3.5, 3.5, 1 / 32, 1 / 128, 16, 16)}
    for T, e in expect.items():
        L = port.lattice_for_T(T)
        assert (L.end_g, L.end_b, L.dg, L.db, L.rlg, L.rlb) == e
        assert ref.lattice_regime(T)[:6] == e


needs_ref = pytest.mark.skipif(not ref.available(), reason="oracle/_ref/libkgref.so not built (needs /root/reference)")


@needs_ref
def test_integrate2d_table_vs_reference():
    rng = np.random.default_rng(3)
    for scale in (1e-3, 1e-5, 1.0):
        tab = np.ascontiguousarray(rng.normal(size=(25, 193)) * scale)
        n = C.c_int()
        v_ref = ref.lib().kgref_integrate2d_table(tab.ctypes.data_as(ref.dp), 25, 193, 1e-7, 1e-7, 8.0, 5.0, 0.5, 0.03125, 4, 16, 1e-5, C.byref(n))
        v, n2 = port.integrate2d_table(tab, 1e-7, 1e-7, 8.0, 5.0, 0.5, 0.03125, 4, 16, 1e-5)
        assert v == v_ref and n2 == n.value


@needs_ref
def test_tanh_sinh_vs_reference():
    for which, a in ((0, 0.0), (1, 1.0), (1, -3.0)):
        for h, tol in ((1e-3, 1e-5), (7.5e-3, 1e-6)):
            assert port.lib().kgo_tanh_sinh_test(which, a, h, tol) == ref.lib().kgref_tanh_sinh_test(which, a, h, tol)


@needs_ref
@pytest.mark.parametrize("comp,T", [("AR", 300.0), ("CO2", 180.0), ("NC10", 500.0), ("HE", 20.0)])
def test_chi_R_vs_reference(comp, T):
    ids, Mw, s, e, la, lr = mie_parameters(comp)
    m = ref.RefMie(Mw, s, e, la, lr, is_idealgas=True, py_constants=False)
    p, sig, _ = _pair(comp)
    rng = np.random.default_rng(5)
    g = rng.uniform(0.005, 8, 80)
    b = rng.uniform(0.002, 5, 80)
    assert np.array_equal(port.chi(p, T, g, b * sig), m.chi(0, 0, T, g, b * sig))
    assert np.array_equal(port.get_R(p, T, g, b * sig), m.get_R(0, 0, T, g, b * sig))


@needs_ref
def test_w_vs_reference_live():
    ids, Mw, s, e, la, lr = mie_parameters("N2")
    m = ref.RefMie(Mw, s, e, la, lr, is_idealgas=True, py_constants=False)
    p, sig, _ = _pair("N2")
    assert port.w_integral(p, 400.0, 2, 4) == m.w_integral(0, 0, 400.0, 2, 4)
