"""ctypes access to oracle/libkgoracle.so (the plain-C restatement, oracle/kg_oracle.c).
TEST INFRASTRUCTURE ONLY -- importable from tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs; never from kineticgas_b200/."""
import ctypes as C
import os
import subprocess
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None
dp = C.POINTER(C.c_double)
KGO_MAX_TERMS = 6


class Pair(C.Structure):
    _fields_ = [("sigma", C.c_double), ("eps", C.c_double), ("nterms", C.c_int),
                ("C", C.c_double * KGO_MAX_TERMS), ("lam", C.c_double * KGO_MAX_TERMS)]


class Lattice(C.Structure):
    _fields_ = [("end_g", C.c_double), ("end_b", C.c_double), ("dg", C.c_double), ("db", C.c_double),
                ("rlg", C.c_int), ("rlb", C.c_int), ("lim", C.c_double)]


def build():
    subprocess.check_call(["make", "-s", "-C", _HERE, "libkgoracle.so"])


def lib():
    global _LIB
    if _LIB is None:
        so = os.path.join(_HERE, "libkgoracle.so")
        if not os.path.exists(so):
            build()
        L = C.CDLL(so)
        L.kgo_mie_pair.argtypes = [C.c_double] * 4 + [C.POINTER(Pair)]
        L.kgo_lattice_for_T.argtypes = [C.c_double, C.POINTER(Lattice)]
        for n in ("kgo_potential", "kgo_potential_r"):
            getattr(L, n).restype = C.c_double
            getattr(L, n).argtypes = [C.POINTER(Pair), C.c_double]
        L.kgo_tanh_sinh_test.restype = C.c_double
        L.kgo_tanh_sinh_test.argtypes = [C.c_int, C.c_double, C.c_double, C.c_double]
        for n in ("kgo_get_R", "kgo_theta", "kgo_chi"):
            getattr(L, n).restype = C.c_double
            getattr(L, n).argtypes = [C.POINTER(Pair), C.c_double, C.c_double, C.c_double, C.POINTER(C.c_int)]
        L.kgo_w_integrand.restype = C.c_double
        L.kgo_w_integrand.argtypes = [C.POINTER(Pair), C.c_double, C.c_double, C.c_double, C.c_int, C.c_int, C.POINTER(C.c_int)]
        L.kgo_integrate2d_table.restype = C.c_double
        L.kgo_integrate2d_table.argtypes = [dp, C.c_int, C.c_int] + [C.c_double] * 6 + [C.c_int, C.c_int, C.c_double, C.POINTER(C.c_int)]
        L.kgo_w_integral.restype = C.c_double
        L.kgo_w_integral.argtypes = [C.POINTER(Pair), C.c_double, C.c_int, C.c_int, C.POINTER(Lattice), C.POINTER(C.c_int),
                                     C.POINTER(C.c_long), C.POINTER(C.c_long), C.POINTER(C.c_long)]
        L.kgo_tl_ewca.restype = C.c_double
        L.kgo_tl_ewca.argtypes = [C.POINTER(Pair), C.c_double, C.c_int, C.POINTER(C.c_int)]
        L.kgo_get_bmid.restype = C.c_double
        L.kgo_get_bmid.argtypes = [C.POINTER(Pair), C.c_double, C.c_double, C.POINTER(C.c_int)]
        L.kgo_get_b_max_g.restype = C.c_double
        L.kgo_get_b_max_g.argtypes = [C.POINTER(Pair), C.c_double, C.c_double, C.c_double, C.POINTER(C.c_int)]
        L.kgo_omega_prefactor.restype = C.c_double
        L.kgo_omega_prefactor.argtypes = [C.c_double, C.c_double, C.c_double, C.c_double, C.c_int]
        _LIB = L
    return _LIB


def mie_pair(sigma, eps, la, lr):
    p = Pair()
    lib().kgo_mie_pair(sigma, eps, la, lr, C.byref(p))
    return p


def lattice_for_T(T):
    l = Lattice()
    lib().kgo_lattice_for_T(T, C.byref(l))
    return l


def chi(pair, T, g, b):
    st = C.c_int(0)
    return np.array([lib().kgo_chi(C.byref(pair), T, gi, bi, C.byref(st)) for gi, bi in zip(np.atleast_1d(g), np.atleast_1d(b))])


def get_R(pair, T, g, b):
    st = C.c_int(0)
    return np.array([lib().kgo_get_R(C.byref(pair), T, gi, bi, C.byref(st)) for gi, bi in zip(np.atleast_1d(g), np.atleast_1d(b))])


def w_integral(pair, T, l, r, lattice=None, counts=False):
    lat = lattice or lattice_for_T(T)
    st = C.c_int(0)
    nchi, nt, nl = C.c_long(), C.c_long(), C.c_long()
    v = lib().kgo_w_integral(C.byref(pair), T, l, r, C.byref(lat), C.byref(st), C.byref(nchi), C.byref(nt), C.byref(nl))
    if st.value:
        raise RuntimeError("get_R reached max iter!")
    return (v, nchi.value, nt.value, nl.value) if counts else v


def integrate2d_table(tab, ox, oy, ex, ey, dx, dy, rlx, rly, lim):
    tab = np.ascontiguousarray(tab, float)
    n = C.c_int()
    v = lib().kgo_integrate2d_table(tab.ctypes.data_as(dp), tab.shape[0], tab.shape[1], ox, oy, ex, ey, dx, dy, rlx, rly, lim, C.byref(n))
    return v, n.value


def tl_ewca(pair, T, prop):
    """prop 0 = MTL, 1 = ETL; metres"""
    st = C.c_int(0)
    v = lib().kgo_tl_ewca(C.byref(pair), T, prop, C.byref(st))
    if st.value & 8:
        raise RuntimeError("Unable to converge bmid!")
    return v
