"""GPU parity of the classical cross sections Q^(l)(E) and the classical second virial coefficient (kg_cross_sections,
kg_second_virial; Spherical::cross_section / second_virial, cpp/Spherical.cpp:104-160) against the unmodified reference build
(tests/golden/xsec_reference.npz, `python -m oracle.make_golden xsec`). Tolerance: 1e-6 relative (north_star)."""
import os
import numpy as np
import pytest

from kineticgas_b200.fluids import mie_parameters, BOLTZMANN, AVOGADRO

pytestmark = pytest.mark.gpu
GOLD = np.load(os.path.join(os.path.dirname(__file__), "golden", "xsec_reference.npz"))


def _model(comp):
    from kineticgas_b200 import DeviceModel, mie_desc
    ids, Mw, s, e, la, lr = mie_parameters(comp)
    return DeviceModel([mie_desc(s[0], e[0] * BOLTZMANN, la[0], lr[0])], [Mw[0] * 1e-3 / AVOGADRO], True)


@pytest.mark.parametrize("comp", ["AR", "CO2", "LJF", "H2"])
def test_cross_sections_vs_reference(comp):
    m = _model(comp)
    Q = m.cross_sections(0, GOLD["E"], GOLD["l"])
    err = np.abs(Q / GOLD[comp + "_Q"] - 1)
    print(comp, "max rel err", float(err.max()))
    assert err.max() < 1e-6


@pytest.mark.parametrize("comp", ["AR", "CO2", "LJF", "H2"])
def test_second_virial_vs_reference(comp):
    m = _model(comp)
    B = m.second_virial(0, GOLD["T"])
    assert np.max(np.abs(B / GOLD[comp + "_B2"] - 1)) < 1e-9


def test_hard_sphere_limit_of_reduced_cross_section():
    """At high energy the l = 1 cross section of a steep repulsive core tends to the hard-sphere value pi d^2 with d the classical
    turning distance: Q*(1) = Q / (pi sigma^2) falls monotonically with E and is below 1 at E = 40 eps."""
    m = _model("AR")
    ids, Mw, s, e, la, lr = mie_parameters("AR")
    Q = m.cross_sections(0, GOLD["E"], [1])[:, 0] / (np.pi * s[0] ** 2)
    assert np.all(np.diff(Q) < 0) and Q[-1] < 1.0
