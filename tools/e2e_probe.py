"""Where the host-buffer (e2e) call of the configs[1] workload spends its time: plan creation, kernels, fetch, destroy."""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from kineticgas_b200.model import DeviceModel, mie_desc

Mw, s, e, la, lr, KB, NA = bench.ljf()
model = DeviceModel([mie_desc(s, e * KB, la, lr)], [Mw * 1e-3 / NA], True, device=0)
T = bench.grid(1, 0)
ls = np.array([l for l, r in bench.LR], np.int32)
rs = np.array([r for l, r in bench.LR], np.int32)
for it in range(6):
    torch.cuda.synchronize()
    t0 = time.perf_counter(); p = model.omega_plan(T, ls, rs); torch.cuda.synchronize()
    t1 = time.perf_counter(); p.run(sync=True)
    t2 = time.perf_counter(); om = p.fetch()
    t3 = time.perf_counter(); p.close(); torch.cuda.synchronize()
    t4 = time.perf_counter()
    print(f"create {1e3*(t1-t0):.2f}  run {1e3*(t2-t1):.2f}  fetch {1e3*(t3-t2):.2f}  close {1e3*(t4-t3):.2f}  total {1e3*(t4-t0):.2f} ms", flush=True)
