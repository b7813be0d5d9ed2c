#!/usr/bin/env python
"""Profiling driver: quantum Omega(l,r)(T) of HFD-B2 helium, 64 temperatures x 5 (l,r) (the bench leg), once."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from kineticgas_b200 import HFD_B2
kin = HFD_B2("HE")
T = np.geomspace(10.0, 1000.0, 64)
t0 = time.perf_counter()
om = kin.dev.quantum_omega(0, T, [1, 1, 2, 2, 3], [1, 2, 2, 3, 3], 1.0, 0.0)
print("seconds", time.perf_counter() - t0, float(om[0, 0]))
