"""Compact fluid parameter table (Mie potentials) for the components the hot path is exercised on.

Values are the published Mie parameters (sigma [m], eps/k [K], lambda_a, lambda_r) and molar masses [g/mol] that
the reference ships in its JSON fluid database (reference: fluids/<ID>.json, key "Mie"/"default", read by
cpp/KineticGas.cpp:65-88 and pykingas/py_KineticGas.py:48). Mie-FH<n> entries are the Feynman-Hibbs parameter
sets (key "Mie-FH<n>"/"default").
"""
BOLTZMANN = 1.38064852e-23           # reference: cpp/global_params.h:4
GAS_CONSTANT = 8.31446261815324      # cpp/global_params.h:5
AVOGADRO = GAS_CONSTANT / BOLTZMANN  # cpp/global_params.h:6
PI_REF = 3.14159265359               # cpp/global_params.h:7 (truncated on purpose: parity)
# pykingas forms the constructor arguments of the C++ core with scipy.constants (CODATA 2018): eps = eps_div_k * Boltzmann
# (pykingas/MieType.py:214), m = Mw 1e-3 / Avogadro (pykingas/py_KineticGas.py:61) -- while the core itself uses the 2014 values
# above. The host mirror does the same, so that its numbers are pykingas' numbers (upstream's golden CSVs to 1e-13).
BOLTZMANN_PY = 1.380649e-23
AVOGADRO_PY = 6.02214076e23
PLANCK = 6.62607015e-34
HBAR = PLANCK / (2.0 * PI_REF)

#        Mw        sigma       eps/k    la    lr
MIE = {
    "AR":   (39.948,   3.404e-10,  117.84, 6.0,  12.085),
    "KR":   (83.798,   3.645e-10,  176.1,  6.0,  14.0),
    "XE":   (131.3,    3.9612e-10, 229.47, 6.0,  12.275),
    "NE":   (20.183,   2.8019e-10, 29.875, 6.0,  9.6977),
    "HE":   (4.003,    2.8525e-10, 4.5764, 6.0,  8.3471),
    "H2":   (2.016,    3.2574e-10, 17.931, 6.0,  8.0),
    "O2":   (31.999,   3.46e-10,   118.0,  6.0,  12.0),
    "N2":   (28.013,   3.609e-10,  105.79, 6.0,  14.08),
    "CO2":  (44.01,    3.741e-10,  353.51, 6.66, 23.0),
    "C1":   (16.0425,  3.7412e-10, 153.36, 6.0,  12.65),
    "C2":   (30.07,    3.7257e-10, 206.12, 6.0,  12.4),
    "C3":   (44.097,   3.9056e-10, 239.89, 6.0,  13.006),
    "NC10": (142.286,  4.589e-10,  400.79, 6.0,  18.885),
    "LJF":  (1.0,      3e-10,      30.0,   6.0,  12.0),
}
#  (fluid, FH order) ->  sigma, eps/k, la, lr
MIE_FH = {
    ("HE", 1): (2.7443e-10, 5.4195, 6.0, 9.0),
    ("HE", 2): (2.549e-10, 10.952, 6.0, 13.0),
    ("H2", 1): (3.0243e-10, 26.706, 6.0, 9.0),
    ("H2", 2): (2.9195e-10, 55.729, 6.0, 20.0),
    ("NE", 1): (2.7778e-10, 37.501, 6.0, 13.0),
    ("NE", 2): (2.776e-10, 37.716, 6.0, 13.0),
    ("LJF", 1): (3e-10, 30.0, 6.0, 12.0),
    ("LJF", 2): (3e-10, 30.0, 6.0, 12.0),
}

# Ab-initio pair potentials (reference: fluids/<ID>.json keys "ModTangToennis", "HFD-B2", "Patowski", entry "default";
# read by cpp/multiparam.cpp:5-8,91-107,184-205). Energies in K, lengths in the unit named by L_unit / rm / Angstrom.
TANG_TOENNIES = {
    "AR": {'A_div_k': 46133014.6, 'b': 40.2517211, 'A_tilde_div_k': 936167.467, 'a_tilde': 21.5969557, 'a1': -29.833763, 'a2': -9.71208881, 'am1': 0.0275206827, 'am2': -0.010148905, 'L_unit': 1e-09, 'short_range_lim': 0.1504728, 'C': [0.442812017, 0.0326707684, 0.00245656537, 0.000188246247, 1.47012192e-05, 1.17006343e-06], 'eps_div_k': 143.123, 'sigma': 3.35741e-10},
    "NE": {'A_div_k': 18398301.8, 'b': 6.5723746, 'A_tilde_div_k': 19929886.9, 'a_tilde': 3.79784233, 'a1': -4.25209106, 'a2': -0.0243585519, 'am1': 2.1876006, 'am2': -0.929771828, 'L_unit': 1e-10, 'short_range_lim': 1.2, 'C': [43908.7458, 169024.069, 806233.71, 4765277.904, 34900399.11, 316728711.3], 'eps_div_k': 41.99815, 'sigma': 2.762e-10},
    "HE": {'A_div_k': 2833791.99, 'b': 2.00788607, 'A_tilde_div_k': 486211.48, 'a_tilde': 0.40944977284199885, 'a1': -1.986231822, 'a2': -0.0503428424, 'am1': -0.3514929118, 'am2': 0.1101468439, 'L_unit': 5.2917721e-11, 'short_range_lim': 0.5, 'C': [461339.753, 4458060.72, 58005040.2, 1031004230.0, 24137814200.0, 718387329000.0], 'eps_div_k': 11.064023706224583, 'sigma': 2.63903617e-10},
}
HFD_B2 = {
    "HE": {'A': 192215.29, 'alpha': 10.73520708, 'c6': 1.34920045, 'c8': 0.41365922, 'c10': 0.17078164, 'beta_star': -1.89296514, 'D': 1.4132, 'rm': 0.297, 'eps_div_k': 10.94, 'sigma': 2.64205e-10},
}
PATKOWSKI = {
    "H2": {'Rc': 1.5875316000000002, 'Ac': 13163.006387928035, 'Bc': -2.9743503851130892, 'Cex1': 15.8362221077622, 'Cex2': -3.277863589578727, 'Csp1': -0.27517157146222365, 'Csp2': 0.6868560829347136, 'Csp3': -0.19398157814701056, 'Csp4': 0.014014872862309222, 'delta': 3.277863589578727, 'C6': -83611.45487725046, 'C8': -414751.48191027937, 'C10': -2555564.249525306, 'eps_div_k': 35.45197264678177, 'sigma': 3.04661415419141e-10},
}


def mie_parameters(comps):
    """Return (ids, Mw[g/mol], sigma, eps_div_k, la, lr) lists for a comma separated component string."""
    ids = [c.strip() for c in comps.split(",")]
    rows = [MIE[c] for c in ids]
    return ids, [r[0] for r in rows], [r[1] for r in rows], [r[2] for r in rows], [r[3] for r in rows], [r[4] for r in rows]
