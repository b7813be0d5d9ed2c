"""GPU parity of the chi / Omega kernels (through the C-ABI) against the C oracle (oracle/kg_oracle.c, itself pinned
bit-exactly to the reference build, tests/test_oracle_vs_reference.py)."""
import numpy as np
import pytest

from kineticgas_b200.fluids import mie_parameters, BOLTZMANN, AVOGADRO

pytestmark = pytest.mark.gpu


def _model(comp):
    from kineticgas_b200 import DeviceModel, mie_desc
    ids, Mw, s, e, la, lr = mie_parameters(comp)
    return DeviceModel([mie_desc(s[0], e[0] * BOLTZMANN, la[0], lr[0])], [Mw[0] * 1e-3 / AVOGADRO], True), (s[0], e[0] * BOLTZMANN, la[0], lr[0])


@pytest.mark.parametrize("comp,T", [("AR", 300.0), ("AR", 40.0), ("CO2", 250.0), ("LJF", 15.0)])
def test_chi_points_vs_oracle(oracle_port, comp, T):
    m, (s, eps, la, lr) = _model(comp)
    p = oracle_port.mie_pair(s, eps, la, lr)
    rng = np.random.default_rng(1)
    g = np.concatenate([rng.uniform(0.01, 8, 300), [1e-7, 0.5, 2.0]])
    b = np.concatenate([rng.uniform(0.002, 5, 300), [0.5, 1e-7, 150.0]])
    chi_gpu, R_gpu = m.chi_points(0, T, g, b, want_R=True)
    chi_ref = oracle_port.chi(p, T, g, b * s)
    R_ref = oracle_port.get_R(p, T, g, b * s) / s
    ok = b < 100
    # tolerance: pointwise |chi_gpu - chi_ref| <= 1e-9, R to 1e-10 relative
    assert np.max(np.abs(chi_gpu - chi_ref)) < 1e-9
    sel = ok & (b >= 1e-3)
    assert np.max(np.abs(R_gpu[sel] / R_ref[sel] - 1)) < 1e-10


@pytest.mark.parametrize("comp,T", [("AR", 300.0), ("CO2", 250.0), ("AR", 60.0), ("LJF", 45.0)])
def test_w_integral_vs_oracle(oracle_port, comp, T):
    m, (s, eps, la, lr) = _model(comp)
    p = oracle_port.mie_pair(s, eps, la, lr)
    ls, rs = [1, 1, 2, 3], [1, 4, 2, 3]
    om, w = m.omega_batch([T], ls, rs, want_w=True)
    for k, (l, r) in enumerate(zip(ls, rs)):
        w_ref = oracle_port.w_integral(p, T, l, r)
        # north_star tolerance: 1e-6 relative on Omega
        assert abs(w[0, 0, k] / w_ref - 1) < 1e-6, (l, r, w[0, 0, k], w_ref)


@pytest.mark.parametrize("comp,T", [("AR", 300.0), ("CO2", 250.0), ("H2", 120.0)])
def test_transfer_lengths_vs_oracle(oracle_port, comp, T):
    m, (s, eps, la, lr) = _model(comp)
    p = oracle_port.mie_pair(s, eps, la, lr)
    mtl, etl = m.transfer_lengths([T])
    # 1e-6 relative (north_star); these feed K-factors as tl^3
    assert abs(mtl[0, 0] / oracle_port.tl_ewca(p, T, 0) - 1) < 1e-7
    assert abs(etl[0, 0] / oracle_port.tl_ewca(p, T, 1) - 1) < 1e-7


def test_strided_quadrature_matches_node_by_node():
    """The Euler-Maclaurin (stride 8) theta integrals against the reference's node-by-node sums on whole adaptive lattices
    of all five T regimes: same W to 1e-9, same number of reference nodes (i.e. the same stop index in every integral),
    about 1/7 of the node evaluations."""
    from kineticgas_b200 import DeviceModel, mie_desc
    from kineticgas_b200.model import set_exact_quadrature
    from kineticgas_b200.fluids import BOLTZMANN, AVOGADRO
    m = DeviceModel([mie_desc(3e-10, 30.0 * BOLTZMANN, 6.0, 12.0)], [1e-3 / AVOGADRO], True)
    T = [9.5, 14.0, 25.0, 40.0, 60.0, 300.0, 2500.0]
    l, r = [1, 1, 2, 2, 3, 4], [1, 4, 2, 3, 3, 4]
    res = {}
    try:
        for exact in (1, 0):
            set_exact_quadrature(exact)
            p = m.omega_plan(T, l, r)
            p.run()
            res[exact] = (p.fetch(want_w=True)[1], p.quadrature_stats())
            p.close()
    finally:
        set_exact_quadrature(0)
    w_exact, s_exact = res[1]
    w_em, s_em = res[0]
    print(s_exact, s_em, float(np.max(np.abs(w_em / w_exact - 1))))
    assert np.max(np.abs(w_em / w_exact - 1)) < 1e-9
    assert s_em["reference_nodes"] == s_exact["reference_nodes"] and s_em["theta_integrals"] == s_exact["theta_integrals"]
    assert s_exact["executed_nodes"] == s_exact["reference_nodes"]
    assert s_em["executed_nodes"] < 0.25 * s_em["reference_nodes"]
    assert s_em["redo_integrals"] < 0.1 * s_em["theta_integrals"]


def test_full_size_config1_strided_equals_node_by_node():
    """BASELINE configs[1] at full size (LJF, 4096-point T* grid 0.3-100, the ten (l,r) with 1 <= l <= r <= 4): every one of the
    40 960 collision integrals from the strided sums equals the node-by-node evaluation to 5e-9, the reference node totals are
    identical (same stop index in all ~3e7 theta integrals, same refinement decisions), and the reduced integrals have the known LJ 12-6 limits."""
    from kineticgas_b200 import DeviceModel, mie_desc
    from kineticgas_b200.model import set_exact_quadrature
    from kineticgas_b200.fluids import BOLTZMANN, AVOGADRO
    sig, eps, mass = 3e-10, 30.0 * BOLTZMANN, 1e-3 / AVOGADRO
    m = DeviceModel([mie_desc(sig, eps, 6.0, 12.0)], [mass], True)
    Ts = np.geomspace(0.3, 100.0, 4096)
    T = 30.0 * Ts
    lr = [(l, r) for l in range(1, 5) for r in range(l, 5)]
    l, r = [a for a, _ in lr], [b for _, b in lr]
    res = {}
    try:
        for exact in (0, 1):
            set_exact_quadrature(exact)
            p = m.omega_plan(T, l, r)
            p.run()
            res[exact] = (p.fetch(), p.quadrature_stats())
            p.close()
    finally:
        set_exact_quadrature(0)
    om, st = res[0]
    om_x, st_x = res[1]
    assert np.all(np.isfinite(om)) and np.all(om > 0)
    # strided (stride 16 accepted at 1e-7 against stride 32, else stride 8, else node by node) against node by node: typical 3e-11, worst of
    # the 40 960 integrals 1.2e-9 -- three orders of magnitude inside the 1e-6 contract, which tests/test_gpu_regimes.py checks against the reference
    assert np.max(np.abs(om / om_x - 1)) < 5e-9
    assert st["reference_nodes"] == st_x["reference_nodes"] and st["theta_integrals"] == st_x["theta_integrals"]
    # reduced integrals Omega* = Omega / Omega_HS (cpp/MieKinGas.h:52-60): Omega*(2,2) of LJ 12-6 is 1.593 at T* = 1 and
    # 0.5882 at T* = 100 (Hirschfelder, Curtiss & Bird table I-M; the reference's coarse lattice agrees to ~1 %)
    def omega_hs(ll, rr, Tk):
        w = np.pi * sig ** 2 * 0.5 * (rr + 1) * np.prod(np.arange(2, rr + 1, dtype=float))
        if ll % 2 == 0:
            w *= 1.0 - 1.0 / (ll + 1.0)
        return np.sqrt(BOLTZMANN * Tk / (np.pi * mass)) * w
    k22 = lr.index((2, 2))
    star = om[:, 0, k22] / omega_hs(2, 2, T)
    assert abs(star[np.argmin(np.abs(Ts - 1.0))] / 1.593 - 1) < 0.02
    assert abs(star[-1] / 0.5882 - 1) < 0.02
    assert np.all(np.diff(star[Ts > 1.0]) < 1e-3)          # decreasing in T* above the well depth (up to lattice noise)
