#!/usr/bin/env python
"""Serial depth of the quantum cross-section kernel: wall time of ONE energy (one warp) against a batch, HFD-B2 helium."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from kineticgas_b200 import HFD_B2
kin = HFD_B2("HE")
d = kin.dev
d.quantum_cross_sections(0, [1, 2, 3], [1.0], 1.0, 0.0)
for E in (0.01, 0.1, 1.0, 10.0, 50.0, 100.0, 119.0, 121.0, 300.0, 1000.0, 3000.0, 10000.0):
    t0 = time.perf_counter()
    q = d.quantum_cross_sections(0, [1, 2, 3], [E], 1.0, 0.0)
    t1 = time.perf_counter()
    print("E* %8.2f  one warp %8.2f ms   Q1 %.6e" % (E, 1e3 * (t1 - t0), float(np.ravel(q)[0])))
for E in (1.0, 100.0, 1000.0):
    Es = E * (1 + 1e-3 * np.arange(2368))
    t0 = time.perf_counter()
    d.quantum_cross_sections(0, [1, 2, 3], Es, 1.0, 0.0)
    print("E* %8.2f  2368 warps (one wave) %8.2f ms" % (E, 1e3 * (time.perf_counter() - t0)))
