#!/usr/bin/env python
"""Tuning probe (GPU): first-tier acceptance threshold of the strided theta sums (KG_EM_ACCEPT16) against step time of configs[1],
tier statistics, and the Omega error against the reference-built regime fixture.   python tools/tune_accept.py [thr ...]"""
import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def child():
    import numpy as np
    from kineticgas_b200 import DeviceModel, mie_desc
    from kineticgas_b200.fluids import BOLTZMANN, AVOGADRO
    m = DeviceModel([mie_desc(3e-10, 30.0 * BOLTZMANN, 6.0, 12.0)], [1e-3 / AVOGADRO], True)
    T = 30.0 * np.geomspace(0.3, 100.0, 4096)
    lr = [(l, r) for l in range(1, 5) for r in range(l, 5)]
    p = m.omega_plan(T, [a for a, _ in lr], [b for _, b in lr])
    for _ in range(2):
        p.run()
    t0 = time.perf_counter()
    for _ in range(3):
        p.run()
    ms = (time.perf_counter() - t0) / 3 * 1e3
    c, q = p.counters(), p.quadrature_stats()
    g = np.load(os.path.join(ROOT, "tests", "golden", "regimes_reference.npz"))
    Tg = np.concatenate([g["grid_T"], g["edges_T"]])
    ref = np.concatenate([g["grid_omega"], g["edges_omega"]])
    om = m.omega_batch(Tg, g["lr"][:, 0], g["lr"][:, 1])[:, 0, :]
    err = np.abs(om / ref - 1)
    cold = Tg < 20
    print(json.dumps({"ms_step": ms, "ms_theta": c["ms_theta"], "ms_mesh": c["ms_mesh"], "tier2_frac": q["stride8_integrals"] / q["theta_integrals"],
                      "redo_frac": q["redo_integrals"] / q["theta_integrals"], "exec_frac": q["executed_nodes"] / q["reference_nodes"],
                      "ref_nodes": q["reference_nodes"], "err_max_warm": float(err[~cold].max()), "err_max_cold": float(err[cold].max()),
                      "err_p99_cold": float(np.quantile(err[cold], 0.99))}))


if __name__ == "__main__":
    if os.environ.get("KG_TUNE_CHILD"):
        child()
    else:
        for thr in (sys.argv[1:] or ["1e-9", "1e-8", "1e-7", "1e-6", "1e-5"]):
            env = dict(os.environ, KG_TUNE_CHILD="1", KG_EM_ACCEPT16=thr)
            out = subprocess.run([sys.executable, __file__], env=env, capture_output=True, text=True)
            print(thr, out.stdout.strip().splitlines()[-1] if out.stdout.strip() else out.stderr[-500:], flush=True)
