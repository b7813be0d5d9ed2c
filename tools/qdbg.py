import numpy as np, os, sys, time
sys.path.insert(0, os.getcwd())
G = dict(np.load("tests/golden/quantum_reference.npz"))
from kineticgas_b200 import HFD_B2
kin = HFD_B2("HE")
tag = "hfd_he"
lr = G[tag + "_lr"]
for rep in range(2):
    t0 = time.perf_counter()
    om = kin.dev.quantum_omega(0, G[tag + "_Tom"], lr[:, 0], lr[:, 1], 1.0, 0.0)
    print("3 T x 5 (l,r): %.1f ms" % (1e3 * (time.perf_counter() - t0)), np.abs(om / G[tag + "_omega"] - 1).max())
T = np.geomspace(10.0, 1000.0, 64)
t0 = time.perf_counter()
om = kin.dev.quantum_omega(0, T, lr[:, 0], lr[:, 1], 1.0, 0.0)
dt = time.perf_counter() - t0
print("64 T x 5 (l,r) = 320 Omega: %.1f ms -> %.0f Omega/s" % (1e3 * dt, 320 / dt))
t0 = time.perf_counter()
eta = [kin.viscosity(t, 0.024, [1.0], N=2) for t in (20.0, 300.0)]
print("quantum viscosity He 20 K, 300 K:", eta, "%.1f ms" % (1e3 * (time.perf_counter() - t0)))
kin.set_quantum_active(False)
print("classical:", [kin.viscosity(t, 0.024, [1.0], N=2) for t in (20.0, 300.0)])
