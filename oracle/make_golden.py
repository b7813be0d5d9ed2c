"""Generate tests/golden/*.npz from the UNMODIFIED reference build (oracle/_ref/libkgref.so). Run in the container
that has /root/reference:  python -m oracle.make_golden
TEST INFRASTRUCTURE ONLY."""
import os
import sys
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import ref  # noqa: E402
from kineticgas_b200.fluids import mie_parameters  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")


def chi_w_golden():
    out = {}
    rng = np.random.default_rng(20261104)
    cases = [("AR", 300.0), ("AR", 60.0), ("CO2", 250.0), ("LJF", 40.0), ("LJF", 25.0), ("H2", 15.0)]
    for comp, T in cases:
        ids, Mw, s, e, la, lr = mie_parameters(comp)
        m = ref.RefMie(Mw, s, e, la, lr, is_idealgas=True)
        m.set_omega_correlation_active(False)
        g = np.concatenate([rng.uniform(0.01, 8, 60), [1e-7, 0.25, 3.0]])
        b = np.concatenate([rng.uniform(0.002, 5, 60), [0.7, 1e-7, 2.0]])
        key = f"{comp}_{T:g}"
        out[key + "_g"] = g
        out[key + "_b"] = b
        out[key + "_chi"] = m.chi(0, 0, T, g, b * s[0])
        out[key + "_R"] = m.get_R(0, 0, T, g, b * s[0]) / s[0]
    wcases = [("AR", 300.0, [(1, 1), (1, 4), (2, 2), (3, 3)]), ("CO2", 250.0, [(1, 1), (2, 3)]), ("AR", 60.0, [(1, 5)]),
              ("LJF", 45.0, [(2, 2)]), ("LJF", 30.0, [(1, 1)])]
    for comp, T, lrs in wcases:
        ids, Mw, s, e, la, lr = mie_parameters(comp)
        m = ref.RefMie(Mw, s, e, la, lr, is_idealgas=True)
        m.set_omega_correlation_active(False)
        for (l, r) in lrs:
            out[f"W_{comp}_{T:g}_{l}_{r}"] = np.array([m.w_integral(0, 0, T, l, r), m.omega(0, 0, l, r, T)])
            print("W", comp, T, l, r, out[f"W_{comp}_{T:g}_{l}_{r}"], flush=True)
    np.savez(os.path.join(GOLD, "chi_w_reference.npz"), **out)


if __name__ == "__main__":
    os.makedirs(GOLD, exist_ok=True)
    which = sys.argv[1:] or ["chi_w"]
    if "chi_w" in which:
        chi_w_golden()
