"""GPU parity of the Feynman-Hibbs Mie path (QuantumMie: T-dependent effective coefficients, 4 and 6 power-law terms)
against fixtures from the UNMODIFIED reference build (tests/golden/qmie_reference.npz, oracle/make_golden.py qmie)."""
import os
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

GOLD = np.load(os.path.join(os.path.dirname(__file__), "golden", "qmie_reference.npz"))
LR = [(1, 1), (1, 2), (1, 3), (2, 2), (2, 3), (2, 4), (3, 3)]
KEYS = sorted(k[:-6] for k in GOLD.files if k.endswith("_omega"))


def _model(key):
    from kineticgas_b200 import QuantumMie
    comp, fh, T = key.split("_")
    return QuantumMie(comp, FH_orders=int(fh[2:]), N=2, is_idealgas=True), float(T)


@pytest.mark.parametrize("key", KEYS)
def test_effective_coefficients(key):
    kin, T = _model(key)
    ce = GOLD[key + "_ceff"]
    assert np.allclose(kin._coeff_T([T])[0, 0, :ce.size], ce, rtol=1e-13, atol=0)


@pytest.mark.parametrize("key", KEYS)
def test_chi_points(key):
    kin, T = _model(key)
    chi = kin.dev.chi_points(0, T, GOLD[key + "_g"], GOLD[key + "_b"], coeff=kin._coeff_T([T])[0, 0])
    assert np.max(np.abs(chi - GOLD[key + "_chi"])) < 1e-9


@pytest.mark.parametrize("key", KEYS)
def test_omega(key):
    kin, T = _model(key)
    om = kin.omega_table([T], [l for l, r in LR], [r for l, r in LR])[0, 0]
    # north_star: 1e-6 relative
    assert np.max(np.abs(om / GOLD[key + "_omega"] - 1)) < 1e-6


@pytest.mark.parametrize("key", KEYS)
def test_ideal_gas_properties(key):
    kin, T = _model(key)
    Vm = 8.314 * T / 1e5
    eos = {"dmudn": np.array([[[8.314 * T]]]), "cp_ideal": np.array([[2.5 * 8.31446261815324 + 1.0]])}
    eta = kin.viscosity(T, Vm, [1.0], N=2)
    lam = kin.thermal_conductivity(T, Vm, [1.0], N=2, idealgas=False, contributions="tdi", eos_inputs=eos)
    D = kin.interdiffusion(T, Vm, [1.0], N=2, eos_inputs=eos)
    assert abs(eta / float(GOLD[key + "_eta"]) - 1) < 1e-6
    ref = GOLD[key + "_lambda"]
    assert abs((lam["t"] + lam["d"] + lam["i"]) / ref.sum() - 1) < 1e-6
    assert abs(D / GOLD[key + "_D"][0, 0] - 1) < 1e-6


def test_nonideal_models_without_rdf_are_refused():
    """Potentials without an RDF model keep the reference's error (cpp/KineticGas.h:259); QuantumMie has one since the extended
    Sutherland RDF runs on the device (tests/test_gpu_qmie_ret.py)."""
    from kineticgas_b200 import QuantumMie, HFD_B2
    kin = QuantumMie("HE", FH_orders=1, N=2, is_idealgas=False)
    assert kin.viscosity(100.0, 1e-3, [1.0], N=2) > 0
    gen = HFD_B2("HE", N=2)
    gen.is_idealgas = False
    with pytest.raises(RuntimeError, match="RDF model is not implemented"):
        gen.viscosity(100.0, 1e-3, [1.0], N=2)
