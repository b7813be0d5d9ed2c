#!/usr/bin/env python
"""Profiling driver: device tables (kg_tcache_create) of configs[4] -- 100 temperatures, 15 pairs of the 5-component RET-Mie mixture, N = 2 --
built three times from empty caches. Prints the stage timings of the last build and host-clock times of each."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import bench
from kineticgas_b200 import MieKinGas
kin = MieKinGas(bench.FC_COMPS, N=bench.FC_N, is_idealgas=False)
T, Vm, x = bench.fc_grid(1, 0)
for _ in range(3):
    kin._caches = {}
    t0 = time.perf_counter()
    cache = kin.tcache(T, bench.FC_N)
    print("tcache ms", 1e3 * (time.perf_counter() - t0), kin.dev.last_timings())
