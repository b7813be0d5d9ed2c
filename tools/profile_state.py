#!/usr/bin/env python
"""Profiling driver: configs[4] state points (5-component RET-Mie, N = 2, 10^5 points) through kg_transport_batch, three passes."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import bench
from kineticgas_b200 import MieKinGas, pinned_copy
kin = MieKinGas(bench.FC_COMPS, N=bench.FC_N, is_idealgas=False)
T, Vm, x = bench.fc_grid(1, 0)
eos = {"dmudn": pinned_copy(bench.sp_eos(T, Vm, x)), "cp_ideal": pinned_copy(np.full((T.size, 5), 4.5 * bench.R_GAS))}
cache = kin.tcache(T, bench.FC_N)
for _ in range(3):
    res = kin.transport_batch(T, Vm, x, N=bench.FC_N, eos_inputs=eos, cache=cache)
print(kin.dev.last_timings(), float(res["viscosity"][0]))
