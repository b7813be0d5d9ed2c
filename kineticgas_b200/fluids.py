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


def mie_parameters(comps):
    """Return (ids, Mw[g/mol], sigma, eps_div_k, la, lr) lists for a comma separated component string."""
    ids = [c.strip() for c in comps.split(",")]
    rows = [MIE[c] for c in ids]
    return ids, [r[0] for r in rows], [r[1] for r in rows], [r[2] for r in rows], [r[3] for r in rows], [r[4] for r in rows]
