"""The SAFT-VR Mie model of kineticgas_b200/csrc/kg_saft.cuh, compiled for the HOST (tests/host/eos_host.cpp: the same generic code the
device kernels instantiate) and checked on CPU against thermopack 2.2.3 -- the EoS pykingas constructs (pykingas/MieKinGas.py:55-65) --
through the fixture tests/golden/eos_reference.npz (oracle/make_golden.py eos, generated here from the wheel vendored in the reference
tree). Pins the model term by term: Barker-Henderson diameters, hard-sphere and dispersion Helmholtz energy, total residual, the total
d mu_i / d n_j (what the RET code asks for) and the pressure. No GPU needed."""
import ctypes as C
import os

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "tests", "host", "libeos_host.so")
G = dict(np.load(os.path.join(ROOT, "tests", "golden", "eos_reference.npz")))
RGAS = 6.02214076e23 * 1.380649e-23
TAGS = ["ar", "co2c1n2", "c1nc10", "five"]
dp = C.POINTER(C.c_double)


def P(a):
    return a.ctypes.data_as(dp)


@pytest.fixture(scope="module")
def L():
    if not os.path.exists(LIB):
        pytest.skip("tests/host/libeos_host.so not built (python -c 'import __graft_entry__ as g; g.build()')")
    return C.CDLL(LIB)


def model(tag):
    m, sig, eps = G[tag + "_m"].copy(), np.ascontiguousarray(G[tag + "_sigma"]), np.ascontiguousarray(G[tag + "_eps"])
    la1, lr1 = G[tag + "_la1"], G[tag + "_lr1"]
    la = np.ascontiguousarray(3 + np.sqrt((la1[:, None] - 3) * (la1[None, :] - 3)))
    lr = np.ascontiguousarray(3 + np.sqrt((lr1[:, None] - 3) * (lr1[None, :] - 3)))
    return m, sig, eps, la, lr


def helmholtz(L, prm, T, V, n, terms):
    m, sig, eps, la, lr = prm
    nc = m.size
    F = C.c_double()
    g, H = np.zeros(nc + 1), np.zeros((nc + 1, nc + 1))
    L.kgsaft_host_helmholtz(nc, P(m), P(sig), P(eps), P(la), P(lr), C.c_double(T), C.c_double(V), P(np.ascontiguousarray(n, float)), terms,
                            C.byref(F), P(g), P(H))
    return F.value, g, H


@pytest.mark.parametrize("tag", TAGS)
def test_barker_henderson_diameters(L, tag):
    m, sig, eps, la, lr = model(tag)
    for k in range(4):
        d = np.zeros(m.size)
        L.kgsaft_host_diameters(m.size, P(m), P(sig), P(eps), P(la), P(lr), C.c_double(G[tag + "_T"][k]), P(d))
        assert np.max(np.abs(d / G[tag + "_d"][k] - 1)) < 1e-11


@pytest.mark.parametrize("tag", TAGS)
def test_helmholtz_terms_dmudn_and_pressure(L, tag):
    prm = model(tag)
    nc = prm[0].size
    worst = {"hs": 0.0, "disp": 0.0, "total": 0.0, "dmudn": 0.0, "p": 0.0}
    for k in range(G[tag + "_T"].size):
        T, V, x = G[tag + "_T"][k], G[tag + "_Vm"][k], G[tag + "_x"][k]
        a_hs, a_disp, a_tot = G[tag + "_terms"][k]
        for name, bits, ref in (("hs", 1, a_hs), ("disp", 14, a_disp), ("total", 31, a_tot)):
            F, g, H = helmholtz(L, prm, T, V, x, bits)           # one mole in total: F = a
            worst[name] = max(worst[name], abs(F - ref) / (abs(a_hs) + abs(a_disp)))      # the total passes through zero along an isotherm
        dmudn = RGAS * T * (H[:nc, :nc] + np.diag(1.0 / x))
        ref = G[tag + "_dmudn"][k]
        worst["dmudn"] = max(worst["dmudn"], np.max(np.abs(dmudn - ref)) / np.max(np.abs(ref)))
        p = RGAS * T / V - RGAS * T * g[nc]
        worst["p"] = max(worst["p"], abs(p / G[tag + "_p"][k] - 1))
    print(tag, worst)
    assert worst["hs"] < 1e-10 and worst["disp"] < 1e-10 and worst["total"] < 1e-9
    assert worst["dmudn"] < 1e-9 and worst["p"] < 1e-9
