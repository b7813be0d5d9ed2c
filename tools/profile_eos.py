#!/usr/bin/env python
"""Profiling driver: device SAFT-VR Mie EoS inputs (dmudn + Cp_ideal) of configs[4]'s 10^5 state points, three passes; Vm(T, 1 bar) once."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import bench
from kineticgas_b200 import DeviceSaftVRMie
T, Vm, x = bench.fc_grid(1, 0)
eos = DeviceSaftVRMie(bench.FC_COMPS)
for _ in range(3):
    r = eos.tv_batch(T, Vm, x, dmudn=True, cp_ideal=True)
    print("kernel ms", eos.last_kernel_ms())
v = eos.specific_volume_batch(T, np.full(T.size, 1e5), x)
print("vol kernel ms", eos.last_kernel_ms(), float(r["dmudn"][0, 0, 0]), float(v[0]))
