#!/usr/bin/env python
"""CPU probe behind the stride choice of theta_integral_em (kineticgas_b200/csrc/kg_device.cuh): for random LJ 12-6 trajectories
(T*, g, b) it forms the reference's tanh-sinh terms (cpp/Integration/Integration.cpp:417-437, h = 1e-3), the exact partial sum up
to a coarse node b, and the Euler-Maclaurin estimates of stride 8 / 16 / 32 with 7-point end derivatives, and prints
  * the true relative error of each stride,
  * what the stride-S / stride-2S consistency check accepts at a given threshold and the worst true error it lets through.
numpy only; not part of the product path.   python tools/em_stride_probe.py [n_trajectories] [seed]"""
import sys
from math import factorial
import numpy as np
from scipy.optimize import brentq

PI, H, KMAX = 3.14159265359, 1e-3, 3328
k = np.arange(1, KMAX + 1)
u = np.tanh(PI * np.sinh(k * H) / 2)
w = (PI / 2) * H * np.cosh(k * H) / np.cosh(PI * np.sinh(k * H) / 2) ** 2
w[0] = (PI / 2) * H * np.cosh(H) / np.cosh(PI * np.sinh(H * 1e-3) / 2) ** 2        # the first-weight quirk, Integration.cpp:422


def get_R(Tst, g, b):
    f = lambda r: 4 * (r ** -12 - r ** -6) / (Tst * g * g) + (b / r) ** 2 - 1
    r0 = max(b, 1.0)
    while f(r0) < 0:
        r0 *= 0.95
    hi = r0 / 0.95
    return r0 if f(hi) > 0 else brentq(f, r0, hi, xtol=1e-15)


def terms(Tst, g, b):
    R = get_R(Tst, g, b)
    U = 4 * ((u / R) ** 12 - (u / R) ** 6) / (Tst * g * g)
    X = (R / b) ** 2 / u ** 2 * (1 - U) - 1
    with np.errstate(all="ignore"):
        return (w / u) / np.sqrt(X)


def ref_sum(t):
    I = t[0]
    for j in range(1, len(t)):
        if not np.isfinite(t[j]) or not (abs(t[j]) > abs(1e-5 * I)):
            break
        I += t[j]
    return I, j + 1


def fd(f, n):
    m = len(f)
    A = np.vander(np.arange(m), m, increasing=True).T
    rhs = np.zeros(m)
    rhs[n] = factorial(n)
    return np.linalg.solve(A, rhs) @ f


def em(t, S, b):
    idx = np.arange(2, b + 1, S)
    gs = t[idx - 1]
    T = S * (gs.sum() - 0.5 * (gs[0] + gs[-1]))
    fa, fb = t[1:8], t[b - 1:b + 6]
    B = {1: 1 / 12, 3: -1 / 720, 5: 1 / 30240}
    return T + 0.5 * (gs[0] + gs[-1]) + sum(B[n] * (1 - S ** (n + 1)) * (fd(fb, n) - fd(fa, n)) for n in (1, 3, 5))


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 2000
    rng = np.random.default_rng(int(sys.argv[2]) if len(sys.argv) > 2 else 2)
    rows = []
    for _ in range(n):
        Tst = np.exp(rng.uniform(np.log(0.3), np.log(3)))
        g, b = rng.uniform(0.02, 3.5), rng.uniform(0.02, 3.0)
        t = terms(Tst, g, b)
        I, K = ref_sum(t)
        if K < 200:
            continue
        bb = 2 + 32 * ((K - 12) // 32)
        exact = t[1:bb].sum()
        e8, e16, e32 = em(t, 8, bb), em(t, 16, bb), em(t, 32, bb)
        rows.append([abs(e8 - exact) / I, abs(e16 - exact) / I, abs(e32 - exact) / I, abs(e8 - e16) / I, abs(e16 - e32) / I])
    r = np.array(rows)
    print("trajectories", len(r))
    for c, name in enumerate(["stride 8", "stride 16", "stride 32"]):
        e = r[:, c]
        print(f"{name}: median {np.median(e):.1e}  p99 {np.quantile(e, .99):.1e}  max {e.max():.1e}")
    for thr in (1e-10, 1e-9):
        acc = r[:, 3] <= thr
        print(f"stride 8 accepted by |i8 - i16| <= {thr:.0e}: {acc.mean():.3f}, worst true error {r[acc, 0].max():.1e}")
    for thr in (1e-9, 1e-8, 1e-7):
        acc = r[:, 4] <= thr
        print(f"stride 16 accepted by |i16 - i32| <= {thr:.0e}: {acc.mean():.3f}, worst true error {r[acc, 1].max():.1e}")


if __name__ == "__main__":
    main()
