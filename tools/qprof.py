#!/usr/bin/env python
"""Profiling driver: one wave (2368 warps) of Numerov-heavy cross sections, HFD-B2 helium at E* ~ 100."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from kineticgas_b200 import HFD_B2
d = HFD_B2("HE").dev
E = float(sys.argv[1]) if len(sys.argv) > 1 else 100.0
d.quantum_cross_sections(0, [1, 2, 3], E * (1 + 1e-3 * np.arange(2368)), 1.0, 0.0)
