"""CPU tests of the generic-potential host layer: the piecewise polynomial-times-exponential form the device evaluates
(kg_genpot_describe, built by kineticgas_b200/csrc/kg_genpot.cu, Feynman-Hibbs correction folded in per temperature)
against u(r) and du/dr of the UNMODIFIED reference classes LJSpline, LJTS, ModTangToennis, HFD_B2, Patowski and their
FH_Corrected variants (tests/golden/generic_reference.npz, oracle/make_golden.py generic). No GPU needed."""
import os
import numpy as np
import pytest

from kineticgas_b200 import LJSpline, LJTS, ModTangToennis, HFD_B2, Patowski  # noqa: F401
from kineticgas_b200.model import describe_generic, generic_desc

GOLD = np.load(os.path.join(os.path.dirname(__file__), "golden", "generic_reference.npz"))
# key -> (class name, fluid, FH order)
MODELS = {"ljs": ("LJSpline", None, 0), "ljts": ("LJTS", None, 0), "tt_ar": ("ModTangToennis", "AR", 0),
          "tt_ne_fh1": ("ModTangToennis", "NE", 1), "tt_he_fh2": ("ModTangToennis", "HE", 2), "hfd_he": ("HFD_B2", "HE", 0),
          "hfd_he_fh1": ("HFD_B2", "HE", 1), "hfd_he_fh2": ("HFD_B2", "HE", 2), "pat_h2": ("Patowski", "H2", 0),
          "pat_h2_fh1": ("Patowski", "H2", 1)}
POINT_KEYS = sorted(k[:-4] for k in GOLD.files if k.endswith("_chi"))


def split_key(k):
    key, T = k.rsplit("_", 1)
    return key, float(T)


def host_desc(key):
    """The kg_potential_desc_ex block the host mirror would hand to kg_model_create_ex, without creating a device model."""
    import kineticgas_b200.kingas as kk
    cls, comp, fh = MODELS[key]
    obj = object.__new__(getattr(kk, cls))
    obj._build_device_model = lambda: None
    if comp is None:
        obj.__init__()
    else:
        obj.__init__(comp, FH_order=fh)
    return obj, generic_desc(obj.KIND, obj.sigma_ij[0, 0], obj.epsilon_ij[0, 0], obj._params, obj.FH_order, obj._fh_D if obj.FH_order else 0.0)


def eval_desc(D, r_m, sigma):
    r = r_m / sigma * D["len_scale"]
    for pc in D["pieces"]:
        if r < pc["r_hi"]:
            U = dU = 0.0
            for t in pc["terms"]:
                P = sum(c * r ** (t["pmin"] + k) for k, c in enumerate(t["pc"]))
                Dd = sum(c * r ** (t["dmin"] + k) for k, c in enumerate(t["dc"]))
                E = np.exp(sum(c * r ** (t["qmin"] + k) for k, c in enumerate(t["qc"]))) if len(t["qc"]) else 1.0
                U += P * E
                dU += Dd * E
            return U * D["e_scale"], dU * D["e_scale"] * D["len_scale"] / sigma
    return 0.0, 0.0


@pytest.mark.parametrize("key", sorted(MODELS))
def test_model_scales_match_reference(key):
    obj, _ = host_desc(key)
    sig, eps, m, fhD = GOLD[key + "_scales"]
    assert abs(obj.sigma_ij[0, 0] / sig - 1) < 1e-14 and abs(obj.epsilon_ij[0, 0] / eps - 1) < 1e-14
    assert abs(obj.m[0] / m - 1) < 1e-12 and abs(obj._fh_D / fhD - 1) < 1e-9


@pytest.mark.parametrize("pk", POINT_KEYS)
def test_effective_potential_matches_reference(pk):
    key, T = split_key(pk)
    obj, d = host_desc(key)
    D = describe_generic(d, T)
    sig, eps = obj.sigma_ij[0, 0], obj.epsilon_ij[0, 0]
    r = GOLD[pk + "_r"]
    mine = np.array([eval_desc(D, x, sig) for x in r])
    u_ref, du_ref = GOLD[pk + "_u"], GOLD[pk + "_du"]
    # relative to the larger of the value and 1e-6 of the well depth (the zero crossings)
    assert np.max(np.abs(mine[:, 0] - u_ref) / np.maximum(np.abs(u_ref), 1e-6 * eps)) < 1e-8
    assert np.max(np.abs(mine[:, 1] - du_ref) / np.maximum(np.abs(du_ref), 1e-6 * eps / sig)) < 1e-7


def test_describe_rejects_bad_blocks():
    d = generic_desc(3, 3e-10, 1e-21, [1.0, 2.0])          # Tang-Toennies needs 16 parameters
    with pytest.raises(RuntimeError):
        describe_generic(d)
