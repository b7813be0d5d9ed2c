"""Goldens the reference ITSELF holds, used two ways:

  (CPU, not gpu)  the local build of the unmodified reference (oracle/_ref/libkgref.so) reproduces them at the reference's own
                  tolerance -- this pins the oracle build to upstream's published numbers;
  (GPU)           kineticgas_b200.MieKinGas reproduces them at the north_star tolerance of 1e-6 relative.

  * /root/reference/tests/data/viscosity_MieKinGas.csv (1 050 rows: AR, H2/O2, CO2/C1/N2; T = 150 ... 800 K, rho* = 0.01 ... 0.7),
    compared by /root/reference/tests/test_unchanged_output.py:51-94 through viscosity(T, 1/rho, x, N=2) with
    check_eq_rel(tol=1e-8) (/root/reference/tests/tools.py:6,15-18); rho = rho* / (N_A sigma_11^3). Repacked as
    tests/golden/ref_viscosity_csv.npz by `python -m oracle.make_golden refcsv`. The shear viscosity needs no EoS call.
  * the known answers of /root/reference/tests/test_issues.py:21-29 (reduced viscosity of LJF-like Mie fluids with
    lr = 20, 30, 40, 50 at T* = 2, rho* = 0.1, tolerance 1e-4 absolute).
"""
import os
import numpy as np
import pytest

from kineticgas_b200.fluids import mie_parameters, BOLTZMANN, AVOGADRO
from oracle import ref

CSV = np.load(os.path.join(os.path.dirname(__file__), "golden", "ref_viscosity_csv.npz"))
MIXES = ["AR", "H2,O2", "CO2,C1,N2"]
# /root/reference/tests/test_issues.py:21-24
ISSUE25_VISC = {20: 0.24135277203767233, 30: 0.2425968999401442, 40: 0.24306376868255214, 50: 0.24328026411238798}

needs_ref = pytest.mark.skipif(not ref.available(), reason="oracle/_ref/libkgref.so not built (needs /root/reference)")


def _rows(mix):
    sel = CSV["mix"] == mix
    nc = len(mix.split(","))
    ids, Mw, s, e, la, lr = mie_parameters(mix)
    rho = CSV["rho_star"][sel] / (AVOGADRO * s[0] ** 3)          # mol / m^3 (pykingas get_reducing_units: rho = 1 / (N_A sigma^3))
    return CSV["T"][sel], 1.0 / rho, CSV["x"][sel][:, :nc], CSV["value"][sel]


def _rel(a, b):
    return np.abs(a - b) / (0.5 * (a + b))                         # check_eq_rel, /root/reference/tests/tools.py:15-18


@needs_ref
@pytest.mark.parametrize("mix", MIXES)
def test_reference_build_reproduces_upstream_viscosity_csv(mix):
    """The oracle build against upstream's own golden file at upstream's own tolerance (1e-8 relative)."""
    ids, Mw, s, e, la, lr = mie_parameters(mix)
    m = ref.RefMie(Mw, s, e, la, lr, is_idealgas=False)
    T, Vm, x, val = _rows(mix)
    out = np.array([m.viscosity(float(T[k]), float(Vm[k]), [0.5, 0.5] if m.single else list(x[k]), 2) for k in range(T.size)])
    err = _rel(out, val)
    print(mix, "rows", T.size, "max rel err vs upstream CSV", float(err.max()))
    assert err.max() < 1e-8


def _reduced_visc_units(sigma, eps, mass):
    # pykingas/py_KineticGas.py get_reducing_units: visc = sqrt(m eps) / sigma^2
    return np.sqrt(mass * eps) / sigma ** 2


@needs_ref
@pytest.mark.parametrize("lr_", [20, 30, 40, 50])
def test_reference_build_reproduces_issue25_known_answers(lr_):
    ids, Mw, s, e, la, lr = mie_parameters("LJF")
    m = ref.RefMie([10.0], s, e, la, [float(lr_)], is_idealgas=False)
    T = 2.0 * e[0]
    rho = 0.1 / (AVOGADRO * s[0] ** 3)
    visc = m.viscosity(T, 1.0 / rho, [0.5, 0.5], 2) / _reduced_visc_units(s[0], e[0] * BOLTZMANN, 10.0e-3 / AVOGADRO)
    assert abs(visc - ISSUE25_VISC[lr_]) < 1e-4


@pytest.mark.gpu
@pytest.mark.parametrize("mix", MIXES)
def test_gpu_reproduces_upstream_viscosity_csv(mix):
    """One batched call per mixture through kineticgas_b200.MieKinGas.viscosity (C-ABI: kg_tcache_create + kg_transport_batch)."""
    from kineticgas_b200 import MieKinGas
    kin = MieKinGas(mix, N=2, is_idealgas=False)
    T, Vm, x, val = _rows(mix)
    xx = np.ones((T.size, 1)) if mix == "AR" else x
    out = kin.viscosity(T, Vm, xx, N=2)
    err = np.abs(out / val - 1)
    print(mix, "rows", T.size, "max rel err vs upstream CSV", float(err.max()))
    assert err.max() < 1e-6          # north_star tolerance


@pytest.mark.gpu
@pytest.mark.parametrize("lr_", [20, 30, 40, 50])
def test_gpu_reproduces_issue25_known_answers(lr_):
    from kineticgas_b200 import MieKinGas
    ids, Mw, s, e, la, lr = mie_parameters("LJF")
    kin = MieKinGas("LJF", mole_weights=[10.0], lr=[float(lr_)], N=2, is_idealgas=False)
    T = 2.0 * e[0]
    rho = 0.1 / (AVOGADRO * s[0] ** 3)
    visc = kin.viscosity(T, 1.0 / rho, [1.0], N=2) / _reduced_visc_units(s[0], e[0] * BOLTZMANN, 10.0e-3 / AVOGADRO)
    assert abs(visc - ISSUE25_VISC[lr_]) < 1e-4
