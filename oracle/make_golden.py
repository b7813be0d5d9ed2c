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
        m = ref.RefMie(Mw, s, e, la, lr, is_idealgas=True, py_constants=False)
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
        m = ref.RefMie(Mw, s, e, la, lr, is_idealgas=True, py_constants=False)
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


# ---------------------------------------------------------------------------------------------------------------
# state-point assembly fixtures: the reference's own rdf / K-factors / matrices / transport coefficients
# ---------------------------------------------------------------------------------------------------------------
R_GAS = 8.31446261815324


def synthetic_eos(T, Vm, x, Mw, B0=-1.0e-4):
    """EoS-derived inputs used identically by the reference (through the table-backed GenericEoS of oracle/ref_driver.cpp)
    and by the GPU path: an ideal-gas mixture plus a symmetric second-virial-like term. n = x (one mole), V = Vm."""
    x = np.asarray(x, float)
    nc = x.size
    B = B0 * (1.0 + 0.1 * (np.arange(nc)[:, None] + np.arange(nc)[None, :]))
    dmudn = R_GAS * T * (np.diag(1.0 / x) + 2.0 * B / Vm)
    cp_id = R_GAS * (2.5 + 0.35 * np.arange(1, nc + 1) + 1e-3 * (T - 300.0) * np.sqrt(np.asarray(Mw) / 16.0))
    dvdn = Vm * (1.0 + 0.05 * (np.arange(nc) - (nc - 1) / 2.0))
    return dmudn, cp_id, dvdn


TRANSPORT_CASES = [
    # name, comps, N, is_idealgas, [(T, rho_star (of sigma_00), x)], frames
    ("ar_single", "AR", 2, False, [(300.0, 0.05, [1.0]), (300.0, 0.4, [1.0]), (150.0, 0.2, [1.0])]),
    ("ar_single_ideal", "AR", 3, True, [(300.0, 0.05, [1.0])]),
    ("c1_nc10_N3", "C1,NC10", 3, False, [(400.0, 0.1, [0.3, 0.7]), (400.0, 0.45, [0.9, 0.1]), (650.0, 0.3, [0.5, 0.5])]),
    ("h2_o2_ideal", "H2,O2", 2, True, [(312.5, 0.01, [0.25, 0.75])]),
    ("co2_c1_n2", "CO2,C1,N2", 2, False, [(312.5, 0.3, [0.2, 0.3, 0.5]), (475.0, 0.6, [0.6, 0.1, 0.3])]),
]


def transport_golden(cases=TRANSPORT_CASES, fname="transport_reference.npz"):
    out = {}
    for name, comps, N, ideal, points in cases:
        ids, Mw, s, e, la, lr = mie_parameters(comps)
        m = ref.RefMie(Mw, s, e, la, lr, is_idealgas=ideal)
        nc = m.nc
        single = m.single
        for k, (T, rho_star, x) in enumerate(points):
            rho = rho_star / s[0] ** 3
            Vm = ref.AVOGADRO / rho
            xi = [0.5, 0.5] if single else x
            dmudn, cp_id, dvdn = synthetic_eos(T, Vm, x, Mw)
            key = f"{name}_{k}"
            out[key + "_in"] = np.array([T, Vm, rho] + list(xi))
            out[key + "_dmudn"], out[key + "_cp"], out[key + "_dvdn"] = dmudn, cp_id, dvdn
            m.set_eos(dmudn=dmudn, cp_id=cp_id, p=1e5, vm=Vm, dvdn=dvdn)
            out[key + "_rdf"] = m.rdf(rho, T, xi)
            K, Kp = m.K_factors(rho, T, xi)
            out[key + "_K"], out[key + "_Kp"] = K, Kp
            if not ideal:
                out[key + "_mtl"], out[key + "_etl"] = m.mtl(rho, T, xi), m.etl(rho, T, xi)
            M, v = m.matrix_vector(0, rho, T, xi, N)
            out[key + "_cond_matrix"], out[key + "_cond_vector"] = M, v
            M, v = m.matrix_vector(2, rho, T, xi, N)
            out[key + "_visc_matrix"], out[key + "_visc_vector"] = M, v
            out[key + "_viscosity"] = np.array(m.viscosity(T, Vm, xi, N))
            out[key + "_contribs"] = m.thermal_conductivity_contribs(T, Vm, xi, N)
            for frame, fname_ in ((0, "CoM"), (1, "CoN"), (2, "CoV"), (3, "solvent")):
                solv = 0 if frame == 3 else -1
                out[key + f"_D_{fname_}"] = m.interdiffusion(T, Vm, xi, N, frame=frame, dep=-1, solvent=solv, do_compress=True)
                out[key + f"_DT_{fname_}"] = m.thermal_diffusion_coeff(T, Vm, xi, N, frame=frame, dep=-1, solvent=solv)
            out[key + "_D_CoN_full"] = m.interdiffusion(T, Vm, xi, N, frame=1, dep=0, solvent=-1, do_compress=False)
            print(key, "eta", out[key + "_viscosity"], "lambda", out[key + "_contribs"], flush=True)
    np.savez(os.path.join(GOLD, fname), **out)


if __name__ == "__main__" and "transport" in sys.argv[1:]:
    transport_golden()


# ---------------------------------------------------------------------------------------------------------------
# Feynman-Hibbs Mie (QuantumMie, configs[3]): Omega by quadrature with T-dependent coefficients, ideal-gas properties
# ---------------------------------------------------------------------------------------------------------------
QMIE_LR = [(1, 1), (1, 2), (1, 3), (2, 2), (2, 3), (2, 4), (3, 3)]
QMIE_CASES = [("HE", 1, [25.0, 60.0, 300.0]), ("H2", 1, [40.0, 150.0]), ("HE", 2, [80.0])]


def qmie_golden(fname="qmie_reference.npz"):
    from kineticgas_b200.fluids import MIE, MIE_FH
    out = {}
    rng = np.random.default_rng(99)
    for comp, fh, Ts in QMIE_CASES:
        sig, epsk, la, lr = MIE_FH[(comp, fh)]
        Mw = MIE[comp][0]
        m = ref.RefQuantumMie([Mw], [sig], [epsk], [la], [lr], fh, is_idealgas=True)
        for T in Ts:
            key = f"{comp}_FH{fh}_{T:g}"
            m.set_internals(T)
            out[key + "_ceff"] = m.c_eff(0, 0)
            g = rng.uniform(0.05, 4, 24)
            b = rng.uniform(0.01, 3, 24)
            out[key + "_g"], out[key + "_b"] = g, b
            out[key + "_chi"] = m.chi(0, 0, T, g, b * sig)
            out[key + "_omega"] = np.array([m.omega(0, 0, l, r, T) for (l, r) in QMIE_LR])
            Vm = 8.314 * T / 1e5
            m.set_eos(dmudn=np.array([[8.314 * T]]), cp_id=np.array([2.5 * 8.31446261815324 + 1.0]), p=1e5, vm=Vm, dvdn=np.array([Vm]))
            out[key + "_eta"] = np.array(m.viscosity(T, Vm, [0.5, 0.5], 2))
            out[key + "_lambda"] = m.thermal_conductivity_contribs(T, Vm, [0.5, 0.5], 2)
            out[key + "_D"] = m.interdiffusion(T, Vm, [0.5, 0.5], 2, frame=1)
            print(key, out[key + "_omega"], out[key + "_eta"], out[key + "_lambda"], out[key + "_D"], flush=True)
    np.savez(os.path.join(GOLD, fname), **out)


if __name__ == "__main__" and "qmie" in sys.argv[1:]:
    qmie_golden()

FIVE_CASES = [("five_comp", "C1,C2,C3,N2,CO2", 2, False, [(400.0, 0.25, [0.35, 0.2, 0.1, 0.15, 0.2]), (400.0, 0.5, [0.05, 0.05, 0.3, 0.5, 0.1])])]

if __name__ == "__main__" and "five" in sys.argv[1:]:
    transport_golden(FIVE_CASES, "transport5_reference.npz")


# ---------------------------------------------------------------------------------------------------------------
# non-power-law potentials (LJ/spline, LJTS, modified Tang-Toennies, HFD-B2, Patkowski, Feynman-Hibbs corrected)
# ---------------------------------------------------------------------------------------------------------------
GENERIC_CASES = [  # (key, kind, fluid, fh_order, temperatures of the chi points, [(T, l, r)] of the W / Omega values)
    ("ljs", 1, "", 0, [300.0, 60.0], [(300.0, 1, 1), (300.0, 2, 2), (100.0, 1, 1)]),
    ("ljts", 2, "", 0, [300.0], [(300.0, 1, 1), (150.0, 2, 2)]),
    ("tt_ar", 3, "AR", 0, [300.0], [(300.0, 1, 1), (300.0, 2, 2)]),
    ("tt_ne_fh1", 3, "NE", 1, [100.0, 40.0], [(100.0, 1, 1), (100.0, 2, 2)]),
    ("tt_he_fh2", 3, "HE", 2, [80.0], [(80.0, 2, 2)]),
    ("hfd_he", 4, "HE", 0, [300.0], [(300.0, 1, 1)]),
    ("hfd_he_fh1", 4, "HE", 1, [75.0, 30.0], [(75.0, 1, 1), (75.0, 2, 2)]),
    ("hfd_he_fh2", 4, "HE", 2, [75.0], [(75.0, 2, 2)]),
    ("pat_h2", 5, "H2", 0, [300.0], [(300.0, 2, 2)]),
    ("pat_h2_fh1", 5, "H2", 1, [100.0, 50.0], [(100.0, 1, 1), (100.0, 2, 2)]),
]


def generic_golden(fname="generic_reference.npz"):
    import time
    out = {}
    rng = np.random.default_rng(20261105)
    for key, kind, comp, fh, chiT, wcases in GENERIC_CASES:
        m = ref.RefGeneric(kind, comp, fh)
        out[key + "_scales"] = np.array([m.sigma0, m.eps0, m.m0, m.fh_D])
        rr = np.concatenate([np.linspace(0.35, 3.0, 30), [5.0, 30.0, 1000.0]]) * m.sigma0
        for T in chiT:
            g = np.concatenate([rng.uniform(0.05, 8, 28), [1e-7, 0.25, 3.0, 7.5]])
            b = np.concatenate([rng.uniform(0.002, 4, 28), [0.7, 1e-7, 2.0, 0.01]])
            k = f"{key}_{T:g}"
            out[k + "_r"] = rr
            out[k + "_u"] = m.potential(rr, T)
            out[k + "_du"] = m.potential_r(rr, T)
            out[k + "_g"], out[k + "_b"] = g, b
            out[k + "_chi"] = m.chi(0, 0, T, g, b * m.sigma0)
            out[k + "_R"] = m.get_R(0, 0, T, g, b * m.sigma0) / m.sigma0
        for (T, l, r) in wcases:
            t0 = time.time()
            out[f"W_{key}_{T:g}_{l}_{r}"] = np.array([m.w_integral(0, 0, T, l, r), m.omega(0, 0, l, r, T)])
            print("W", key, T, l, r, out[f"W_{key}_{T:g}_{l}_{r}"], f"{time.time() - t0:.1f} s", flush=True)
    np.savez(os.path.join(GOLD, fname), **out)


if __name__ == "__main__" and "generic" in sys.argv[1:]:
    generic_golden()


def kt_golden(fname="kt_reference.npz"):
    """Thermal diffusion ratios k_T of the state points of TRANSPORT_CASES (inputs are those of transport_reference.npz)."""
    out = {}
    for name, comps, N, ideal, points in TRANSPORT_CASES:
        if N < 2:
            continue
        ids, Mw, s, e, la, lr = mie_parameters(comps)
        m = ref.RefMie(Mw, s, e, la, lr, is_idealgas=ideal)
        for k, (T, rho_star, x) in enumerate(points):
            rho = rho_star / s[0] ** 3
            Vm = ref.AVOGADRO / rho
            xi = [0.5, 0.5] if m.single else x
            dmudn, cp_id, dvdn = synthetic_eos(T, Vm, x, Mw)
            m.set_eos(dmudn=dmudn, cp_id=cp_id, p=1e5, vm=Vm, dvdn=dvdn)
            out[f"{name}_{k}_kT"] = m.thermal_diffusion_ratio(T, Vm, xi, N)
            print(name, k, out[f"{name}_{k}_kT"], flush=True)
    np.savez(os.path.join(GOLD, fname), **out)


if __name__ == "__main__" and "kt" in sys.argv[1:]:
    kt_golden()


def ljs_transport_golden(fname="ljs_transport_reference.npz"):
    """RET state points of the reference's LJSpline class: rdf, transfer lengths, K-factors, eta, lambda (t, d, i)."""
    out = {}
    m = ref.RefGeneric(1, "", 0)
    sig = m.sigma0
    Mw = [m.m0 * ref.AVOGADRO * 1e3]
    pts = [(150.0, 0.05), (150.0, 0.35), (300.0, 0.2), (45.0, 0.01)]       # last: T* = 0.36 < 0.4, all Omega by quadrature
    for k, (T, rho_star) in enumerate(pts):
        rho = rho_star / sig ** 3
        Vm = ref.AVOGADRO / rho
        xi = [0.5, 0.5]
        dmudn, cp_id, dvdn = synthetic_eos(T, Vm, [1.0], Mw)
        m.set_eos(dmudn=dmudn, cp_id=cp_id, p=1e5, vm=Vm, dvdn=dvdn)
        key = f"ljs_{k}"
        out[key + "_in"] = np.array([T, Vm, rho] + xi)
        out[key + "_cp"] = cp_id
        out[key + "_rdf"] = m.rdf(rho, T, xi)
        K, Kp = m.K_factors(rho, T, xi)
        out[key + "_K"], out[key + "_Kp"] = K, Kp
        out[key + "_mtl"], out[key + "_etl"] = m.mtl(rho, T, xi), m.etl(rho, T, xi)
        out[key + "_viscosity"] = np.array(m.viscosity(T, Vm, xi, 2))
        out[key + "_contribs"] = m.thermal_conductivity_contribs(T, Vm, xi, 2)
        for (l, r) in ((1, 1), (1, 2), (1, 3), (2, 2), (2, 3)):
            out[key + f"_omega_{l}_{r}"] = np.array(m.omega(0, 0, l, r, T, model=True))
        print(key, "eta", out[key + "_viscosity"], "lambda", out[key + "_contribs"], "rdf", out[key + "_rdf"][0, 0], flush=True)
    np.savez(os.path.join(GOLD, fname), **out)


if __name__ == "__main__" and "ljs" in sys.argv[1:]:
    ljs_transport_golden()


# ---------------------------------------------------------------------------------------------------------------
# Non-ideal QuantumMie: extended Sutherland RDF (cpp/ExtendedSutherland.cpp:315-333), transfer lengths, K-factors and the
# Revised Enskog properties at finite density
# ---------------------------------------------------------------------------------------------------------------
QMIE_RET_CASES = [
    # name, comps, FH order, N, [(T, rho_star of sigma_00, x)]
    ("he_fh1", "HE", 1, 2, [(40.0, 0.05, [1.0]), (40.0, 0.4, [1.0]), (150.0, 0.25, [1.0])]),
    ("h2_fh2", "H2", 2, 2, [(60.0, 0.3, [1.0])]),
    ("he_h2_fh1", "HE,H2", 1, 2, [(80.0, 0.2, [0.3, 0.7]), (200.0, 0.45, [0.8, 0.2])]),
]


def qmie_ret_golden(fname="qmie_ret_reference.npz"):
    from kineticgas_b200.fluids import MIE, MIE_FH
    out = {}
    for name, comps, fh, N, points in QMIE_RET_CASES:
        ids = comps.split(",")
        Mw = [MIE[c][0] for c in ids]
        sig, epsk, la, lr = [[MIE_FH[(c, fh)][k] for c in ids] for k in range(4)]
        m = ref.RefQuantumMie(Mw, sig, epsk, la, lr, fh, is_idealgas=False)
        for k, (T, rho_star, x) in enumerate(points):
            rho = rho_star / sig[0] ** 3
            Vm = ref.AVOGADRO / rho
            xi = [0.5, 0.5] if m.single else x
            dmudn, cp_id, dvdn = synthetic_eos(T, Vm, x, Mw, B0=-5.0e-6)
            m.set_eos(dmudn=dmudn, cp_id=cp_id, p=1e5, vm=Vm, dvdn=dvdn)
            key = f"{name}_{k}"
            out[key + "_in"] = np.array([T, Vm, rho] + list(xi))
            out[key + "_dmudn"], out[key + "_cp"], out[key + "_dvdn"] = dmudn, cp_id, dvdn
            out[key + "_rdf"] = m.rdf(rho, T, xi)
            K, Kp = m.K_factors(rho, T, xi)
            out[key + "_K"], out[key + "_Kp"] = K, Kp
            out[key + "_mtl"], out[key + "_etl"] = m.mtl(rho, T, xi), m.etl(rho, T, xi)
            out[key + "_viscosity"] = np.array(m.viscosity(T, Vm, xi, N))
            out[key + "_contribs"] = m.thermal_conductivity_contribs(T, Vm, xi, N)
            out[key + "_D_CoN"] = m.interdiffusion(T, Vm, xi, N, frame=1, dep=-1, solvent=-1, do_compress=True)
            print(key, "rdf", out[key + "_rdf"].ravel(), "eta", out[key + "_viscosity"], "lambda", out[key + "_contribs"], flush=True)
    np.savez(os.path.join(GOLD, fname), **out)


if __name__ == "__main__" and "qmie_ret" in sys.argv[1:]:
    qmie_ret_golden()


# ---------------------------------------------------------------------------------------------------------------
# Ortiz de Zarate frames (ideal gas): D in zarate / zarate_x / zarate_w, D_T (zarate), Soret coefficient
# ---------------------------------------------------------------------------------------------------------------
ZARATE_CASES = [("h2_o2", "H2,O2", 2, [(312.5, [0.25, 0.75])], -1), ("co2_c1_n2", "CO2,C1,N2", 2, [(350.0, [0.2, 0.3, 0.5]), (500.0, [0.6, 0.1, 0.3])], 1)]


def zarate_golden(fname="zarate_reference.npz"):
    out = {}
    for name, comps, N, points, dep in ZARATE_CASES:
        ids, Mw, s, e, la, lr = mie_parameters(comps)
        m = ref.RefMie(Mw, s, e, la, lr, is_idealgas=True)
        for k, (T, x) in enumerate(points):
            Vm = 8.314 * T / 1e5
            dmudn, cp_id, dvdn = synthetic_eos(T, Vm, x, Mw)
            m.set_eos(dmudn=dmudn, cp_id=cp_id, p=1e5, vm=Vm, dvdn=dvdn)
            key = f"{name}_{k}"
            out[key + "_in"] = np.array([T, Vm, dep] + list(x))
            out[key + "_dmudn"] = dmudn
            for fr, nm_ in ((4, "zarate"), (5, "zarate_x"), (6, "zarate_w")):
                out[key + "_D_" + nm_] = m.interdiffusion(T, Vm, x, N, frame=fr, dep=dep, solvent=-1, do_compress=True)
            out[key + "_DT_zarate"] = m.thermal_diffusion_coeff(T, Vm, x, N, frame=4, dep=dep, solvent=-1)
            out[key + "_soret"] = m.soret_coefficient(T, Vm, x, N, dep)
            print(key, out[key + "_D_zarate"].ravel(), out[key + "_DT_zarate"], out[key + "_soret"], flush=True)
    np.savez(os.path.join(GOLD, fname), **out)


if __name__ == "__main__" and "zarate" in sys.argv[1:]:
    zarate_golden()


# ---------------------------------------------------------------------------------------------------------------
# configs[1] / configs[3] on every lattice regime (cpp/Spherical.cpp:60-84): Omega from the reference build at a
# down-sampled headline grid, the regime edges, and cold Feynman-Hibbs temperatures.  ~2 core-hours: run with
#   python -m oracle.make_golden regimes [nproc]
# Results are appended to tests/golden/regimes_reference.jsonl as they finish (restartable), then packed into npz.
# ---------------------------------------------------------------------------------------------------------------
REGIME_LR = [(l, r) for l in range(1, 5) for r in range(l, 5)]
REGIME_EDGES = [19.99, 20.0, 34.99, 35.0, 49.99, 50.0, 69.99, 70.0]
REGIME_QM_T = [10.0, 11.7, 13.3, 15.0, 16.6, 18.2, 19.95, 30.0]
REGIME_QM_LR = [(1, 1), (1, 2), (1, 3), (2, 2), (2, 3), (2, 4), (3, 3)]


def config1_sample_indices(n=4096):
    """48 indices of the configs[1] grid T = 30 geomspace(0.3, 100, n): 12 below 20 K (incl. the coldest point 9.0 K), 8 in 20-35 K,
    6 in 35-50 K, 6 in 50-70 K, 16 above (incl. the hottest point)."""
    T = 30.0 * np.geomspace(0.3, 100.0, n)
    out = []
    for lo, hi, k in ((0.0, 20.0, 12), (20.0, 35.0, 8), (35.0, 50.0, 6), (50.0, 70.0, 6), (70.0, 1e9, 16)):
        idx = np.flatnonzero((T >= lo) & (T < hi))
        out += [int(idx[int(round(q))]) for q in np.linspace(0, idx.size - 1, k)]
    return out


def _regime_worker(task):
    import time
    kind, comp, T, l, r = task
    t0 = time.time()
    if kind == "mie":
        ids, Mw, s, e, la, lr = mie_parameters(comp)
        m = ref.RefMie(Mw, s, e, la, lr, is_idealgas=True, py_constants=False)
        m.set_omega_correlation_active(False)
    else:
        from kineticgas_b200.fluids import MIE, MIE_FH
        sig, epsk, la, lr = MIE_FH[(comp, 1)]
        m = ref.RefQuantumMie([MIE[comp][0]], [sig], [epsk], [la], [lr], 1, is_idealgas=True, py_constants=False)
    val = m.omega(0, 0, l, r, T)
    return {"kind": kind, "comp": comp, "T": T, "l": l, "r": r, "omega": val, "cpu_s": time.time() - t0}


def regimes_golden(nproc=8, fname="regimes_reference.npz"):
    import json
    import multiprocessing as mp
    Tgrid = 30.0 * np.geomspace(0.3, 100.0, 4096)
    idx = config1_sample_indices()
    tasks = [("mie", "LJF", float(Tgrid[i]), l, r) for i in idx for (l, r) in REGIME_LR]
    tasks += [("mie", "LJF", T, l, r) for T in REGIME_EDGES for (l, r) in REGIME_LR]
    tasks += [("qmie", c, T, l, r) for c in ("HE", "H2") for T in REGIME_QM_T for (l, r) in REGIME_QM_LR]
    log = os.path.join(GOLD, "regimes_reference.jsonl")
    done = {}
    if os.path.exists(log):
        for line in open(log):
            d = json.loads(line)
            done[(d["kind"], d["comp"], d["T"], d["l"], d["r"])] = d
    todo = [t for t in tasks if t not in done]
    todo.sort(key=lambda t: t[2])              # coldest (most expensive) first
    print(len(tasks), "tasks,", len(todo), "to do", flush=True)
    with mp.get_context("spawn").Pool(nproc) as pool, open(log, "a") as f:
        for d in pool.imap_unordered(_regime_worker, todo, chunksize=1):
            done[(d["kind"], d["comp"], d["T"], d["l"], d["r"])] = d
            f.write(json.dumps(d) + "\n"); f.flush()
    out = {"grid_index": np.array(idx), "grid_T": Tgrid[idx], "lr": np.array(REGIME_LR), "edges_T": np.array(REGIME_EDGES),
           "qm_T": np.array(REGIME_QM_T), "qm_lr": np.array(REGIME_QM_LR)}
    out["grid_omega"] = np.array([[done[("mie", "LJF", float(Tgrid[i]), l, r)]["omega"] for (l, r) in REGIME_LR] for i in idx])
    out["edges_omega"] = np.array([[done[("mie", "LJF", T, l, r)]["omega"] for (l, r) in REGIME_LR] for T in REGIME_EDGES])
    for c in ("HE", "H2"):
        out[f"qm_{c}_omega"] = np.array([[done[("qmie", c, T, l, r)]["omega"] for (l, r) in REGIME_QM_LR] for T in REGIME_QM_T])
    out["cpu_core_seconds"] = np.array(sum(d["cpu_s"] for d in done.values()))
    np.savez(os.path.join(GOLD, fname), **out)
    print("saved", fname, "core-s", out["cpu_core_seconds"], flush=True)


if __name__ == "__main__" and "regimes" in sys.argv[1:]:
    regimes_golden(int(sys.argv[2]) if len(sys.argv) > 2 else 8)


# ---------------------------------------------------------------------------------------------------------------
# Goldens the REFERENCE ITSELF holds: tests/data/viscosity_MieKinGas.csv (1 050 rows, compared at 1e-8 by
# tests/test_unchanged_output.py:51-94 through viscosity(T, 1/rho, x, N=2); rho = rho* / (N_A sigma_11^3),
# pykingas get_reducing_units). The rows are repacked as arrays (the CSV cannot travel to the GPU box);
# no EoS call is involved in the shear viscosity.
# ---------------------------------------------------------------------------------------------------------------
def refcsv_golden(fname="ref_viscosity_csv.npz"):
    import ast
    import csv
    src = "/root/reference/tests/data/viscosity_MieKinGas.csv"
    mix, T, rho, x, val = [], [], [], [], []
    with open(src) as f:
        for row in csv.DictReader(f):
            xs = ast.literal_eval(row["x"])
            mix.append(row["mix"]); T.append(float(row["T"])); rho.append(float(row["rho*"])); val.append(float(row["value"]))
            x.append(list(xs) + [np.nan] * (3 - len(xs)))
    np.savez(os.path.join(GOLD, fname), mix=np.array(mix), T=np.array(T), rho_star=np.array(rho), x=np.array(x), value=np.array(val),
             source=np.array(src))
    print("saved", fname, len(mix), "rows", sorted(set(mix)))


if __name__ == "__main__" and "refcsv" in sys.argv[1:]:
    refcsv_golden()


# ---------------------------------------------------------------------------------------------------------------
# classical cross sections Q^(l)(E) and second virial coefficients (cpp/Spherical.cpp:104-160) from the reference build
# ---------------------------------------------------------------------------------------------------------------
XSEC_FLUIDS = ["AR", "CO2", "LJF", "H2"]
XSEC_L = [1, 2, 3, 4]


def xsec_golden(fname="xsec_reference.npz"):
    out = {"l": np.array(XSEC_L), "E": np.geomspace(0.05, 40.0, 14), "T": np.array([80.0, 150.0, 300.0, 1000.0])}
    for comp in XSEC_FLUIDS:
        ids, Mw, s, e, la, lr = mie_parameters(comp)
        m = ref.RefMie(Mw, s, e, la, lr, is_idealgas=True, py_constants=False)
        out[comp + "_Q"] = np.array([[m.cross_section(0, 0, l, E) for l in XSEC_L] for E in out["E"]])
        out[comp + "_B2"] = np.array([m.second_virial(0, 0, T) for T in out["T"]])
        print(comp, out[comp + "_Q"][0], out[comp + "_B2"], flush=True)
    np.savez(os.path.join(GOLD, fname), **out)


if __name__ == "__main__" and "xsec" in sys.argv[1:]:
    xsec_golden()


# ---------------------------------------------------------------------------------------------------------------
# BASELINE configs[0]: the EoS-derived inputs of MieKinGas('AR') viscosity_tp / thermal_conductivity_tp at 300 K, 1 bar, from the
# thermopack 2.2.3 wheel vendored in the reference tree (saftvrmie('AR'), what pykingas constructs: pykingas/MieKinGas.py:56-60).
# ---------------------------------------------------------------------------------------------------------------
def config0_eos(fname="config0_eos.npz"):
    import zipfile
    whl = "/root/reference/cpp/external/thermopack/addon/pycThermopack/wheelhouse/thermopack-2.2.3-py3-none-manylinux_2_28_x86_64.whl"
    dst = "/tmp/kg_thermopack_wheel"
    zipfile.ZipFile(whl).extractall(dst)
    sys.path.insert(0, dst)
    from thermopack.saftvrmie import saftvrmie
    eos = saftvrmie("AR")
    T, p = 300.0, 1e5
    Vm, = eos.specific_volume(T, p, [1.0], eos.VAPPH)
    cp = eos.idealenthalpysingle(T, 1, dhdt=True)[1]
    _, dmudn = eos.chemical_potential_tv(T, Vm, [1.0], dmudn=True)
    np.savez(os.path.join(GOLD, fname), T=T, p=p, Vm=Vm, cp_ideal=cp, dmudn=np.asarray(dmudn), source=np.array("thermopack 2.2.3 saftvrmie('AR')"))
    print("configs[0] EoS inputs:", Vm, cp, dmudn)


if __name__ == "__main__" and "eos0" in sys.argv[1:]:
    config0_eos()


# ---------------------------------------------------------------------------------------------------------------
# the EoS-derived inputs (thermopack saftvrmie, the EoS pykingas constructs). Two products, both from the thermopack
# 2.2.3 wheel vendored in the reference tree (it cannot travel to the GPU box):
#   kineticgas_b200/saft_db.json      the SAFT-VR Mie parameters, binary parameters, molar masses and ideal-gas Cp correlations thermopack's
#                                     database holds for the fluids of kineticgas_b200/fluids.py (what `saftvrmie(comps)` would load)
#   tests/golden/eos_reference.npz    dmudn, Cp_ideal, p, dV/dn and Vm(T,p) of thermopack at seeded state points of four mixtures, plus the
#                                     Helmholtz terms used to pin the device model term by term
# ---------------------------------------------------------------------------------------------------------------
def _thermopack():
    import zipfile
    whl = "/root/reference/cpp/external/thermopack/addon/pycThermopack/wheelhouse/thermopack-2.2.3-py3-none-manylinux_2_28_x86_64.whl"
    dst = "/tmp/kg_thermopack_wheel"
    if not os.path.exists(dst):
        zipfile.ZipFile(whl).extractall(dst)
    if dst not in sys.path:
        sys.path.insert(0, dst)
    if not hasattr(np, "product"):
        np.product = np.prod                      # the wheel predates numpy 2
    from thermopack.saftvrmie import saftvrmie
    return saftvrmie


def eos_database(fname="saft_db.json"):
    import json
    from kineticgas_b200.fluids import MIE
    saftvrmie = _thermopack()
    comps = list(MIE)
    db = {"source": "thermopack 2.2.3 saftvrmie database (default parameter reference), read through the python wrapper", "pure": {}, "binary": {}}
    for c in comps:
        eos = saftvrmie(c)
        m, s, e, la, lr = eos.get_pure_fluid_param(1)
        t, par = eos.get_ideal_cp(1)
        db["pure"][c] = {"m": m, "sigma": s, "eps_div_k": e, "la": la, "lr": lr, "mw": float(eos.compmoleweight(1)),
                         "cp_type": int(t), "cp_par": [float(v) for v in par]}
    # binary parameters: one child process per pair (thermopack stops the process on pairs it cannot initialise)
    import subprocess
    code = ("import sys, numpy as np; np.product = np.prod; sys.path.insert(0, '/tmp/kg_thermopack_wheel'); "
            "from thermopack.saftvrmie import saftvrmie; e = saftvrmie(sys.argv[1]); "
            "print('BIN', float(e.get_eps_kij(1, 2)), float(e.get_sigma_lij(1, 2)), float(e.get_lr_gammaij(1, 2)))")
    db["binary_unavailable"] = []
    for i, a in enumerate(comps):
        for b in comps[i + 1:]:
            r = subprocess.run([sys.executable, "-c", code, a + "," + b], capture_output=True, text=True)
            line = [l for l in r.stdout.splitlines() if l.startswith("BIN")]
            if not line:
                db["binary_unavailable"].append(a + "," + b)
                continue
            k, l, g = (float(v) for v in line[0].split()[1:])
            if k or l or g:
                db["binary"][a + "," + b] = {"kij": k, "lij": l, "gammaij": g}
    json.dump(db, open(os.path.join(ROOT, "kineticgas_b200", fname), "w"), indent=1)
    print("saft_db.json:", len(db["pure"]), "fluids,", len(db["binary"]), "pairs with binary parameters")


def eos_reference(fname="eos_reference.npz"):
    saftvrmie = _thermopack()
    rng = np.random.default_rng(20251105)
    out = {}
    for tag, comps in (("ar", "AR"), ("co2c1n2", "CO2,C1,N2"), ("c1nc10", "C1,NC10"), ("five", "CO2,C1,N2,AR,KR")):
        eos = saftvrmie(comps)
        nc = eos.nc
        pure = [eos.get_pure_fluid_param(i + 1) for i in range(nc)]
        sig = np.array([[eos.sigma_ij(i + 1, j + 1) for j in range(nc)] for i in range(nc)])
        eps = np.array([[eos.epsilon_ij(i + 1, j + 1) for j in range(nc)] for i in range(nc)])
        npts = 24
        T = rng.uniform(200.0, 700.0, npts)
        if tag == "c1nc10":
            T = rng.uniform(450.0, 800.0, npts)
        x = rng.dirichlet(np.ones(nc), npts) if nc > 1 else np.ones((npts, 1))
        eta = rng.uniform(0.002, 0.45, npts)                        # segment packing fraction (pi/6) rho sum_i x_i m_i sigma_ii^3
        mvec = np.array([q[0] for q in pure])
        Vm = (np.pi / 6) * 6.02214076e23 * (x * mvec * np.diag(sig) ** 3).sum(axis=1) / eta
        dm, cp, p, dv, terms = [], [], [], [], []
        for k in range(npts):
            n = list(x[k])
            mu, dmudn = eos.chemical_potential_tv(T[k], Vm[k], n, dmudn=True)
            dm.append(np.asarray(dmudn))
            cp.append([eos.idealenthalpysingle(T[k], j + 1, dhdt=True)[1] for j in range(nc)])
            pk, = eos.pressure_tv(T[k], Vm[k], n)
            p.append(pk)
            a_hs, = eos.a_hard_sphere(T[k], Vm[k], n)
            a_d, = eos.a_dispersion(T[k], Vm[k], n)
            A, = eos.helmholtz_tv(T[k], Vm[k], n, property_flag="R")
            terms.append([a_hs, a_d, A / (eos.Rgas * T[k])])
        # Vm(T, p) on the vapour side and dV/dn there, at moderate pressures
        Tp = rng.uniform(250.0 if tag != "c1nc10" else 600.0, 700.0, npts)
        pp = 10.0 ** rng.uniform(4.0, 6.3, npts)
        vm_tp, dvdn_tp = [], []
        for k in range(npts):
            v, dvdn = eos.specific_volume(Tp[k], pp[k], list(x[k]), eos.VAPPH, dvdn=True)
            vm_tp.append(v); dvdn_tp.append(np.asarray(dvdn))
        out.update({f"{tag}_comps": np.array(comps), f"{tag}_m": np.array([q[0] for q in pure]), f"{tag}_sigma": sig, f"{tag}_eps": eps,
                    f"{tag}_la1": np.array([q[3] for q in pure]), f"{tag}_lr1": np.array([q[4] for q in pure]),
                    f"{tag}_T": T, f"{tag}_Vm": Vm, f"{tag}_x": x, f"{tag}_dmudn": np.array(dm), f"{tag}_cp": np.array(cp), f"{tag}_p": np.array(p),
                    f"{tag}_terms": np.array(terms), f"{tag}_Tp": Tp, f"{tag}_pp": pp, f"{tag}_vm_tp": np.array(vm_tp), f"{tag}_dvdn_tp": np.array(dvdn_tp),
                    f"{tag}_d": np.array([eos.hard_sphere_diameters(t)[0] for t in T[:4]])})
        print(tag, "done")
    np.savez(os.path.join(GOLD, fname), **out)


if __name__ == "__main__" and "eos" in sys.argv[1:]:
    eos_database()
    eos_reference()


# ---------------------------------------------------------------------------------------------------------------
# quantum-mechanical collision integrals of the unmodified reference build (cpp/Quantum.cpp) for the ab-initio potentials
# whose models are quantum by default (pykingas/multiparam.py): phase shifts (Numerov, JKWB, the reference's switch), cross sections Q^(n)(E)
# and Omega^(n,s)(T) with quantum_active = true.  tests/golden/quantum_reference.npz
# ---------------------------------------------------------------------------------------------------------------
def quantum_reference(fname="quantum_reference.npz", only=None, with_omega=True):
    import ctypes as C
    from oracle import ref
    L = ref.lib()
    L.kgref_generic_create.restype = C.c_void_p
    L.kgref_generic_create.argtypes = [C.c_int, C.c_char_p, C.c_int, C.c_char_p, C.c_int]
    L.kgref_set_quantum_active.argtypes = [C.c_void_p, C.c_int]
    L.kgref_phase_shift.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_double, C.c_int, C.POINTER(C.c_double)]
    L.kgref_quantum_cross_section.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_double, C.POINTER(C.c_double)]
    L.kgref_omega.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_double, C.POINTER(C.c_double)]
    rng = np.random.default_rng(4)
    out = {}
    val = C.c_double()
    for tag, kind, comp in (("hfd_he", 4, b"HE"), ("mtt_ne", 3, b"NE"), ("pat_h2", 5, b"H2")):
        if only and tag not in only:
            continue
        err = C.create_string_buffer(512)
        h = L.kgref_generic_create(kind, comp, 0, err, 512)
        assert h, err.value
        assert L.kgref_set_quantum_active(h, 1) == 0
        # Numerov phase shifts: low l over a wide energy range, and l near the centrifugal cut-off
        l = np.concatenate([rng.integers(0, 12, 40), rng.integers(12, 80, 24)]).astype(np.int32)
        E = np.concatenate([10.0 ** rng.uniform(-3.2, 2.0, 40), 10.0 ** rng.uniform(0.5, 2.05, 24)])
        num = np.zeros(l.size)
        for k in range(l.size):
            assert L.kgref_phase_shift(h, 0, 0, int(l[k]), float(E[k]), 1, C.byref(val)) == 0
            num[k] = val.value
        # JKWB: the energies and angular momenta where the reference switches to it, and a few below
        lj = rng.integers(0, 400, 40).astype(np.int32)
        Ej = np.concatenate([10.0 ** rng.uniform(2.08, 3.0, 30), 10.0 ** rng.uniform(0.0, 2.0, 10)])
        jk = np.zeros(lj.size)
        for k in range(lj.size):
            assert L.kgref_phase_shift(h, 0, 0, int(lj[k]), float(Ej[k]), 2, C.byref(val)) == 0
            jk[k] = val.value
        # cross sections Q^(n)(E), n = 0..4 (n >= 3 takes the general branch of cross_section_kernel)
        nq = np.tile(np.arange(0, 5), 6).astype(np.int32)
        Eq = np.repeat(10.0 ** np.linspace(-3.5, 2.4, 6), 5)
        Q = np.zeros(nq.size)
        for k in range(nq.size):
            L.kgref_set_quantum_active(h, 0); L.kgref_set_quantum_active(h, 1)      # empties the bucketed cross-section cache
            assert L.kgref_quantum_cross_section(h, 0, 0, int(nq[k]), float(Eq[k]), C.byref(val)) == 0
            Q[k] = val.value
        # Omega^(n,s)(T). The reference memoises cross sections in buckets of 0.01 eps (CrossSectionPoint, cpp/KineticGasCache.h:125-136:
        # round(100 E)), first come first served, so its own Omega depends on which energies were asked for before (1e-3 at low T, where
        # the nodes of one integral already share buckets). Two references are therefore stored:
        #   omega_native : Quantum::omega of a FRESH object per value (kgref_omega_model), kept where no two nodes of the integral can share a bucket
        #   omega_exact  : the reference's quadrature (quantum_omega, cpp/Quantum.cpp:1242-1259 + simpson_inf, Integration.cpp:346-373) restated
        #                  here over the reference's own cross_section with the cache emptied before every call
        L.kgref_omega_model.argtypes = L.kgref_omega.argtypes
        L.kgref_generic_scales.argtypes = [C.c_void_p, C.POINTER(C.c_double)]
        sc = (C.c_double * 4)()
        L.kgref_generic_scales(h, sc)
        eps_J, mu = sc[1], sc[2] / 2
        KB = 1.38064852e-23
        w2 = (C.c_double * 2)()
        L.kgref_symmetry_weights.argtypes = [C.c_void_p, C.c_int, C.c_int, C.POINTER(C.c_double)]
        assert L.kgref_symmetry_weights(h, 0, 0, w2) == 0

        def q_uncached(n, E):
            L.kgref_set_quantum_active(h, 0); L.kgref_set_quantum_active(h, 1)          # state change empties the cache (cpp/Quantum.cpp:1524-1526)
            assert L.kgref_quantum_cross_section(h, 0, 0, int(n), float(E), C.byref(val)) == 0
            return val.value

        def omega_exact(n, s, T):
            kern = lambda x: np.exp(-x) * x ** (s + 1) * q_uncached(n, x * KB * T / eps_J)

            def simpson(a, b):
                dx = (b - a) / 10
                v = kern(a) + kern(b)
                for i in range(1, 5):
                    v += 4 * kern(a + (2 * i - 1) * dx) + 2 * kern(a + 2 * i * dx)
                v += 4 * kern(b - dx)
                return v * dx / 3
            x0, xe = 0.0, (s + 1) / 3
            I = simpson(x0, xe)
            dx, tol, part = (xe - x0) / 10, 1e-8, 1e-2
            while True:
                x0, xe = xe, xe + 10 * dx
                Ip = simpson(x0, xe)
                I += Ip
                c = abs(Ip / I)
                if c < tol:
                    break
                if c < part:
                    dx *= 2; part *= 0.1
                if not part > tol:
                    break
            return I * np.sqrt(2 * KB * T / (np.pi * mu))

        Tom = {"hfd_he": np.array([15.0, 80.0, 300.0]), "mtt_ne": np.array([60.0, 300.0]), "pat_h2": np.array([50.0, 300.0])}[tag]
        if not with_omega:
            Tom = Tom[:0]
        lr = np.array([(1, 1), (1, 2), (2, 2), (2, 3), (3, 3)], np.int32)
        om, om_native = np.zeros((Tom.size, len(lr))), np.full((Tom.size, len(lr)), np.nan)
        for t in range(Tom.size):
            for k, (a, b) in enumerate(lr):
                om[t, k] = omega_exact(a, b, Tom[t])
                if (b + 1) / 30 * KB * Tom[t] / eps_J > 0.02:          # node spacing of the first block in reduced energy: no shared buckets
                    hf = L.kgref_generic_create(kind, comp, 0, err, 512)
                    L.kgref_set_quantum_active(hf, 1)
                    assert L.kgref_omega_model(hf, 0, 0, int(a), int(b), float(Tom[t]), C.byref(val)) == 0
                    om_native[t, k] = val.value
            print(tag, "omega at", Tom[t], om[t], om_native[t] / om[t] - 1, flush=True)
        out[f"{tag}_wts"] = np.array([w2[0], w2[1]]); out[f"{tag}_omega_native"] = om_native
        out.update({f"{tag}_l": l, f"{tag}_E": E, f"{tag}_numerov": num, f"{tag}_lj": lj, f"{tag}_Ej": Ej, f"{tag}_jkwb": jk,
                    f"{tag}_nq": nq, f"{tag}_Eq": Eq, f"{tag}_Q": Q, f"{tag}_Tom": Tom, f"{tag}_lr": lr, f"{tag}_omega": om})
        np.savez(os.path.join(GOLD, fname), **out)
    print("quantum fixture written")


if __name__ == "__main__" and "quantum" in sys.argv[1:]:
    quantum_reference()
if __name__ == "__main__" and "quantum_quick" in sys.argv[1:]:
    quantum_reference("quantum_reference_quick.npz", only=("mtt_ne", "pat_h2"), with_omega=False)
if __name__ == "__main__" and "quantum_h2" in sys.argv[1:]:
    quantum_reference("quantum_reference_h2.npz", only=("pat_h2",))


def quantum_transport(fname="quantum_transport_reference.npz"):
    """Dilute-gas viscosity and thermal conductivity (N = 2) of the reference's HFD-B2 helium with quantum_active = true (the pykingas
    default, pykingas/multiparam.py:126-129): every collision integral from Quantum::quantum_omega."""
    import ctypes as C
    from oracle import ref
    L = ref.lib()
    L.kgref_generic_create.restype = C.c_void_p
    L.kgref_generic_create.argtypes = [C.c_int, C.c_char_p, C.c_int, C.c_char_p, C.c_int]
    L.kgref_set_quantum_active.argtypes = [C.c_void_p, C.c_int]
    L.kgref_viscosity.argtypes = [C.c_void_p, C.c_double, C.c_double, C.POINTER(C.c_double), C.c_int, C.c_int, C.POINTER(C.c_double)]
    L.kgref_thermal_conductivity.argtypes = L.kgref_viscosity.argtypes
    L.kgref_set_eos_tables_nc.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_double), C.c_double, C.c_double, C.POINTER(C.c_double)]
    err = C.create_string_buffer(512)
    T = np.array([20.0, 77.0, 300.0])
    eta, lam = np.zeros(T.size), np.zeros(T.size)
    x = (C.c_double * 2)(0.5, 0.5)
    cp = (C.c_double * 1)(2.5 * 8.31446261815324)
    val = C.c_double()
    for t in range(T.size):
        h = L.kgref_generic_create(4, b"HE", 0, err, 512)          # fresh object per temperature: nothing shared through the bucketed caches
        L.kgref_set_quantum_active(h, 1)
        L.kgref_set_eos_tables_nc(h, 1, None, cp, 1e5, 0.024, None)
        assert L.kgref_viscosity(h, float(T[t]), 0.024, x, 2, 2, C.byref(val)) == 0
        eta[t] = val.value
        assert L.kgref_thermal_conductivity(h, float(T[t]), 0.024, x, 2, 2, C.byref(val)) == 0
        lam[t] = val.value
        print("He quantum", T[t], eta[t], lam[t], flush=True)
    np.savez(os.path.join(GOLD, fname), T=T, viscosity=eta, conductivity=lam, cp_ideal=cp[0])


if __name__ == "__main__" and "quantum_eta" in sys.argv[1:]:
    quantum_transport()
