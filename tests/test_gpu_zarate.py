"""GPU parity of the Ortiz de Zarate frames of reference (ideal gas): interdiffusion in 'zarate', 'zarate_x', 'zarate_w',
thermal diffusion coefficients in 'zarate' and the Soret coefficient S_T = D^-1 D_T (cpp/KineticGas_properties.cpp:179-193,
262-277, 325-330, 397-418) against fixtures from the UNMODIFIED reference build (oracle/make_golden.py zarate)."""
import os
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

GOLD = dict(np.load(os.path.join(os.path.dirname(__file__), "golden", "zarate_reference.npz")))
TOL = 1e-6
CASES = {"h2_o2": ("H2,O2", 2, 1), "co2_c1_n2": ("CO2,C1,N2", 2, 2)}
_models = {}


def rel(a, b):
    a, b = np.asarray(a, float), np.asarray(b, float)
    return float(np.max(np.abs(a - b)) / np.max(np.abs(b)))


def _model(name):
    from kineticgas_b200 import MieKinGas
    if name not in _models:
        _models[name] = MieKinGas(CASES[name][0], N=CASES[name][1], is_idealgas=True)
    return _models[name]


@pytest.mark.parametrize("name,k", [(n, k) for n, c in sorted(CASES.items()) for k in range(c[2])])
def test_zarate_frames(name, k):
    kin = _model(name)
    N = CASES[name][1]
    key = f"{name}_{k}"
    v = GOLD[key + "_in"]
    T, Vm, dep, x = v[0], v[1], int(v[2]), v[3:]
    eos = {"dmudn": GOLD[key + "_dmudn"][None]}
    for frame in ("zarate", "zarate_x", "zarate_w"):
        D = kin.interdiffusion(T, Vm, x, N=N, frame_of_reference=frame, dependent_idx=dep, use_binary=False, eos_inputs=eos)
        assert D.shape == GOLD[key + "_D_" + frame].shape
        assert rel(D, GOLD[key + "_D_" + frame]) < TOL, frame
    DT = kin.thermal_diffusion_coeff(T, Vm, x, N=N, frame_of_reference="zarate", dependent_idx=dep, eos_inputs=eos)
    assert DT.shape == GOLD[key + "_DT_zarate"].shape
    assert rel(DT, GOLD[key + "_DT_zarate"]) < TOL
    S = kin.soret_coefficient(T, Vm, x, N=N, dependent_idx=dep, eos_inputs=eos)
    assert rel(S, GOLD[key + "_soret"]) < TOL
    S2 = kin.soret_coefficient(T, Vm, x, N=N, use_zarate=False, eos_inputs=eos)
    assert S2.shape == (len(x), len(x))


def test_zarate_needs_ideal_gas():
    from kineticgas_b200 import MieKinGas
    kin = MieKinGas("H2,O2", N=2, is_idealgas=False)
    eos = {"dmudn": GOLD["h2_o2_0_dmudn"][None]}
    with pytest.raises(RuntimeError, match="Zarate relation"):
        kin.interdiffusion(312.5, 0.026, [0.25, 0.75], N=2, frame_of_reference="zarate", eos_inputs=eos)
    with pytest.raises(RuntimeError, match="Zarate relation"):
        kin.soret_coefficient(312.5, 0.026, [0.25, 0.75], N=2, eos_inputs=eos)
