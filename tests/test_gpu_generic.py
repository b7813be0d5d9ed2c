"""GPU parity of the non-power-law potentials (LJ/spline, LJTS, modified Tang-Toennies, HFD-B2, Patkowski and their
Feynman-Hibbs corrected variants) through kg_model_create_ex / kg_potential_points / kg_chi_points / kg_omega_batch,
against fixtures from the UNMODIFIED reference build (tests/golden/generic_reference.npz, oracle/make_golden.py generic)."""
import os
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

GOLD = np.load(os.path.join(os.path.dirname(__file__), "golden", "generic_reference.npz"))
from test_genpot_host import MODELS, POINT_KEYS, split_key  # noqa: E402

W_KEYS = sorted(k for k in GOLD.files if k.startswith("W_"))
_cache = {}


def model(key):
    import kineticgas_b200 as kg
    if key not in _cache:
        cls, comp, fh = MODELS[key]
        _cache[key] = getattr(kg, cls)() if comp is None else getattr(kg, cls)(comp, FH_order=fh)
    return _cache[key]


@pytest.mark.parametrize("pk", POINT_KEYS)
def test_potential_points(pk):
    key, T = split_key(pk)
    kin = model(key)
    u, du = kin.dev.potential_points(0, T, GOLD[pk + "_r"])
    eps, sig = kin.epsilon_ij[0, 0], kin.sigma_ij[0, 0]
    assert np.max(np.abs(u - GOLD[pk + "_u"]) / np.maximum(np.abs(GOLD[pk + "_u"]), 1e-6 * eps)) < 1e-8
    assert np.max(np.abs(du - GOLD[pk + "_du"]) / np.maximum(np.abs(GOLD[pk + "_du"]), 1e-6 * eps / sig)) < 1e-7


@pytest.mark.parametrize("pk", POINT_KEYS)
def test_chi_and_R_points(pk):
    key, T = split_key(pk)
    kin = model(key)
    chi, R = kin.dev.chi_points(0, T, GOLD[pk + "_g"], GOLD[pk + "_b"], want_R=True)
    assert np.max(np.abs(chi - GOLD[pk + "_chi"])) < 1e-8
    assert np.max(np.abs(R / GOLD[pk + "_R"] - 1)) < 1e-7


@pytest.mark.parametrize("wk", W_KEYS)
def test_w_and_omega(wk):
    _, rest = wk.split("_", 1)
    key, T, l, r = rest.rsplit("_", 3)
    kin = model(key)
    om, w = kin.dev.omega_batch([float(T)], [int(l)], [int(r)], want_w=True)
    # north_star: 1e-6 relative
    assert abs(w[0, 0, 0] / GOLD[wk][0] - 1) < 1e-6
    assert abs(om[0, 0, 0] / GOLD[wk][1] - 1) < 1e-6


def test_ideal_gas_viscosity_runs_on_generic_model():
    kin = model("tt_ar")
    kin.set_quantum_active(False)           # classical collision integrals here; the partial-wave engine is tests/test_gpu_quantum.py
    eta = kin.viscosity(300.0, 0.024, [1.0], N=2)
    assert 2.0e-5 < eta < 2.6e-5            # argon at 300 K: 22.7 uPa s


LJS = dict(np.load(os.path.join(os.path.dirname(__file__), "golden", "ljs_transport_reference.npz")))    # oracle/make_golden.py ljs


@pytest.mark.parametrize("k", range(4))
def test_ljs_ret_state_point(k):
    """Revised Enskog theory for the LJ/spline fluid: LJSpline::saft_rdf, LJSpline::omega (correlation branch for T* > 0.4),
    EWCA transfer lengths, K-factors, viscosity and the three thermal-conductivity contributions (N = 2)."""
    import kineticgas_b200 as kg
    kin = _cache.setdefault("ljs_ret", kg.LJSpline(N=2, is_idealgas=False))
    key = f"ljs_{k}"
    T, Vm, rho = LJS[key + "_in"][:3]
    r = kin._run(T, Vm, [1.0], 2, 1, ["rdf", "K", "K_prime"])
    assert np.max(np.abs(r["rdf"][0] / LJS[key + "_rdf"] - 1)) < 1e-9
    assert np.max(np.abs(r["K"][0] / LJS[key + "_K"] - 1)) < 1e-6
    assert np.max(np.abs(r["K_prime"][0] / LJS[key + "_Kp"] - 1)) < 1e-6
    assert np.max(np.abs(kin.get_mtl(rho, T, [1.0]) / LJS[key + "_mtl"] - 1)) < 1e-6
    assert np.max(np.abs(kin.get_etl(rho, T, [1.0]) / LJS[key + "_etl"] - 1)) < 1e-6
    om, _, _, is_corr = kin.tcache([T], 2, 1 | 2).fetch()
    ls, rs = kin.tcache([T], 2, 1 | 2).lr()
    for idx, (l, rr) in enumerate(zip(ls.tolist(), rs.tolist())):
        name = key + f"_omega_{l}_{rr}"
        if name in LJS:
            assert abs(om[0, 0, idx] / float(LJS[name]) - 1) < 1e-6, (l, rr, bool(is_corr[0, 0, idx]))
    assert abs(kin.viscosity(T, Vm, [1.0], N=2) / float(LJS[key + "_viscosity"]) - 1) < 1e-6
    c = LJS[key + "_contribs"]
    lam = kin.thermal_conductivity(T, Vm, [1.0], N=2, idealgas=False, contributions="tdi", eos_inputs={"cp_ideal": LJS[key + "_cp"][None, :]})
    for i, name in enumerate("tdi"):
        assert abs(lam[name] - c[i]) <= 1e-6 * abs(c).sum()
