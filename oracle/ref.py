"""ctypes access to oracle/_ref/libkgref.so (the unmodified reference, see oracle/Makefile, oracle/ref_driver.cpp).
TEST INFRASTRUCTURE ONLY -- may be imported from tests/, bench.py's cpu_baseline / --impl reference leg and
oracle/make_golden.py, never from the product package."""
import ctypes as C
import os
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None
BOLTZMANN = 1.38064852e-23                  # cpp/global_params.h:4 (CODATA 2014), what the C++ core computes with
AVOGADRO = 8.31446261815324 / BOLTZMANN     # cpp/global_params.h:6
# pykingas builds its constructor arguments with scipy.constants (CODATA 2018, exact): eps = eps_div_k * Boltzmann
# (pykingas/MieType.py:214), m = Mw 1e-3 / Avogadro (pykingas/py_KineticGas.py:61). The C++ core then forms T* = kT / eps with ITS
# Boltzmann constant, so T* is off by 3.5e-7 relative -- which every number pykingas returns (and upstream's golden CSVs) carries.
BOLTZMANN_PY = 1.380649e-23
AVOGADRO_PY = 6.02214076e23

dp = C.POINTER(C.c_double)
ip = C.POINTER(C.c_int)


def available():
    return os.path.exists(os.path.join(_HERE, "_ref", "libkgref.so"))


def _declare(L):
    L.kgref_mie_create.restype = C.c_void_p
    L.kgref_mie_create.argtypes = [C.c_int, dp, dp, dp, dp, dp, C.c_int, C.c_int]
    L.kgref_last_error.restype = C.c_char_p
    L.kgref_last_error.argtypes = [C.c_void_p]
    for n in ("kgref_destroy", "kgref_clear_cache"):
        getattr(L, n).argtypes = [C.c_void_p]
    L.kgref_set_omega_correlation_active.argtypes = [C.c_void_p, C.c_int]
    L.kgref_set_tl_model.argtypes = [C.c_void_p, C.c_int]
    L.kgref_set_eos_tables_nc.argtypes = [C.c_void_p, C.c_int, dp, dp, C.c_double, C.c_double, dp]
    for n in ("kgref_potential", "kgref_potential_r"):
        getattr(L, n).restype = C.c_double
        getattr(L, n).argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_double]
    for n in ("kgref_chi", "kgref_theta", "kgref_get_R"):
        getattr(L, n).restype = C.c_double
        getattr(L, n).argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_double, C.c_double, C.c_double]
    for n in ("kgref_chi_batch", "kgref_get_R_batch"):
        getattr(L, n).argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_double, C.c_int, dp, dp, dp]
    L.kgref_w_integrand.restype = C.c_double
    L.kgref_w_integrand.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_double, C.c_double, C.c_double, C.c_int, C.c_int]
    for n in ("kgref_omega", "kgref_omega_model"):
        getattr(L, n).argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_double, dp]
    L.kgref_w_integral_tester.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_double, C.c_int, C.c_int,
                                          C.c_double, C.c_double, C.c_double, C.c_double, C.c_int, C.c_int, C.c_double, dp]
    L.kgref_w_integral_trace.argtypes = L.kgref_w_integral_tester.argtypes + [dp, C.c_int, ip]
    L.kgref_integrate2d_table.restype = C.c_double
    L.kgref_integrate2d_table.argtypes = [dp, C.c_int, C.c_int] + [C.c_double] * 6 + [C.c_int, C.c_int, C.c_double, ip]
    L.kgref_tanh_sinh_test.restype = C.c_double
    L.kgref_tanh_sinh_test.argtypes = [C.c_int, C.c_double, C.c_double, C.c_double]
    for n in ("kgref_mtl", "kgref_etl", "kgref_rdf"):
        getattr(L, n).argtypes = [C.c_void_p, C.c_double, C.c_double, dp, dp]
    L.kgref_bh_diameters.argtypes = [C.c_void_p, C.c_double, dp]
    L.kgref_K_factors.argtypes = [C.c_void_p, C.c_double, C.c_double, dp, dp, dp]
    L.kgref_matrix_vector.argtypes = [C.c_void_p, C.c_int, C.c_double, C.c_double, dp, C.c_int, dp, dp, ip]
    for n in ("kgref_viscosity", "kgref_thermal_conductivity", "kgref_thermal_conductivity_contribs"):
        getattr(L, n).argtypes = [C.c_void_p, C.c_double, C.c_double, dp, C.c_int, C.c_int, dp]
    L.kgref_interdiffusion.argtypes = [C.c_void_p, C.c_double, C.c_double, dp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                       C.c_int, dp, ip, ip]
    L.kgref_thermal_diffusion_coeff.argtypes = [C.c_void_p, C.c_double, C.c_double, dp, C.c_int, C.c_int, C.c_int, C.c_int,
                                                C.c_int, dp, ip]
    L.kgref_thermal_diffusion_ratio.argtypes = [C.c_void_p, C.c_double, C.c_double, dp, C.c_int, C.c_int, dp, ip]
    L.kgref_soret_coefficient.argtypes = [C.c_void_p, C.c_double, C.c_double, dp, C.c_int, C.c_int, C.c_int, dp, ip]
    L.kgref_qmie_create.restype = C.c_void_p
    L.kgref_qmie_create.argtypes = [C.c_int, dp, dp, dp, dp, dp, C.c_int, C.c_int, C.c_int]
    L.kgref_set_internals.argtypes = [C.c_void_p, C.c_double, C.c_double]
    L.kgref_c_eff.argtypes = [C.c_void_p, C.c_int, C.c_int, dp, ip]
    L.kgref_generic_create.restype = C.c_void_p
    L.kgref_generic_create.argtypes = [C.c_int, C.c_char_p, C.c_int, C.c_char_p, C.c_int]
    L.kgref_generic_scales.argtypes = [C.c_void_p, dp]
    L.kgref_potential_rr.restype = C.c_double
    L.kgref_potential_rr.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_double]
    L.kgref_cross_section.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_double, dp]
    L.kgref_second_virial.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_double, dp]
    return L


def lib():
    global _LIB
    if _LIB is None:
        _LIB = _declare(C.CDLL(os.path.join(_HERE, "_ref", "libkgref.so")))
    return _LIB


# ---- integration/_build/libkgshim.so: the same unmodified reference objects + GpuMieKinGas (integration/gpu_miekingas.cpp),
# i.e. the reference with its precompute_* / transfer-length entry routed to libkgcuda.so. Holds BOTH arms of the drop-in test.
_SHIM = None
SHIM_PATH = os.path.join(os.path.dirname(_HERE), "integration", "_build", "libkgshim.so")


def shim_available():
    return os.path.exists(SHIM_PATH)


def shim_lib():
    global _SHIM
    if _SHIM is None:
        L = _declare(C.CDLL(SHIM_PATH))
        L.kgshim_mie_create.restype = C.c_void_p
        L.kgshim_mie_create.argtypes = [C.c_int, dp, dp, dp, dp, dp, C.c_int, C.c_int, C.c_int, C.c_char_p, C.c_int]
        L.kgshim_counters.argtypes = [C.c_void_p, C.POINTER(C.c_long), C.c_int]
        L.kgshim_omega_table.argtypes = [C.c_void_p, C.c_int, dp, C.c_int, ip, ip, dp]
        _SHIM = L
    return _SHIM


def _p(a):
    return a.ctypes.data_as(dp)


def lattice_regime(T):
    """The T-dependent lattice switch of Spherical::w_integral (cpp/Spherical.cpp:60-84)."""
    end = (8.0, 5.0)
    dg, db, rlg, rlb = 0.5, 0.03125, 4, 16
    if T < 70:
        dg /= 4.0
        end = (6.0, 3.5)
    if T < 50:
        rlb //= 2
    if T < 35:
        dg /= 2
        rlg *= 2
        db /= 2
        end = (3.5, 3.5)
    if T < 20:
        dg /= 2
        rlg *= 2
        db /= 2
        rlb *= 2
    return end[0], end[1], dg, db, rlg, rlb, 1e-5


class RefMie:
    """The reference's MieKinGas built through its vector constructor (cpp/MieKinGas.cpp:14-17), parameters mixed
    the way pykingas/MieType.py:201-260 does (kij = lij = 0)."""

    def __init__(self, Mw, sigma, eps_div_k, la, lr, is_idealgas=False, py_constants=True):
        single = len(Mw) == 1
        if single:
            Mw, sigma, eps_div_k, la, lr = [list(v) * 2 for v in (Mw, sigma, eps_div_k, la, lr)]
        self.nc = len(Mw)
        self.single = single
        kB, NA = (BOLTZMANN_PY, AVOGADRO_PY) if py_constants else (BOLTZMANN, AVOGADRO)
        self.m = np.array(Mw, float) * 1e-3 / NA
        s = np.array(sigma, float)
        e = np.array(eps_div_k, float) * kB
        self.sigma = 0.5 * (s[:, None] + s[None, :])
        self.eps = np.sqrt(e[None, :] * e[:, None])
        la_, lr_ = np.array(la, float), np.array(lr, float)
        self.la = 3 + np.sqrt((la_[:, None] - 3) * (la_[None, :] - 3))
        self.lr = 3 + np.sqrt((lr_[:, None] - 3) * (lr_[None, :] - 3))
        for a in ("sigma", "eps", "la", "lr"):
            setattr(self, a, np.ascontiguousarray(getattr(self, a)))
        self.L = getattr(self, "_lib_override", None) or lib()
        if getattr(self, "_gpu", False):
            err = C.create_string_buffer(512)
            self.h = self.L.kgshim_mie_create(self.nc, _p(self.m), _p(self.sigma), _p(self.eps), _p(self.la), _p(self.lr),
                                              int(is_idealgas), int(single), int(self._device), err, 512)
            if not self.h:
                raise RuntimeError(err.value.decode())
        else:
            self.h = self.L.kgref_mie_create(self.nc, _p(self.m), _p(self.sigma), _p(self.eps), _p(self.la), _p(self.lr),
                                             int(is_idealgas), int(single))

    def __del__(self):
        try:
            self.L.kgref_destroy(self.h)
        except Exception:
            pass

    def _chk(self, rc):
        if rc != 0:
            raise RuntimeError(self.L.kgref_last_error(self.h).decode())

    def set_omega_correlation_active(self, a):
        self.L.kgref_set_omega_correlation_active(self.h, int(a))

    def clear_cache(self):
        self.L.kgref_clear_cache(self.h)

    def chi(self, i, j, T, g, b):
        g = np.ascontiguousarray(np.atleast_1d(g), float)
        b = np.ascontiguousarray(np.atleast_1d(b), float)
        out = np.empty_like(g)
        self._chk(self.L.kgref_chi_batch(self.h, i, j, T, g.size, _p(g), _p(b), _p(out)))
        return out

    def get_R(self, i, j, T, g, b):
        g = np.ascontiguousarray(np.atleast_1d(g), float)
        b = np.ascontiguousarray(np.atleast_1d(b), float)
        out = np.empty_like(g)
        self._chk(self.L.kgref_get_R_batch(self.h, i, j, T, g.size, _p(g), _p(b), _p(out)))
        return out

    def omega(self, i, j, l, r, T, model=False):
        out = C.c_double()
        f = self.L.kgref_omega_model if model else self.L.kgref_omega
        self._chk(f(self.h, i, j, l, r, T, C.byref(out)))
        return out.value

    def cross_section(self, i, j, l, E):
        """Spherical::cross_section (cpp/Spherical.cpp:104-127): Q^(l)(E), E in units of eps_ij."""
        out = C.c_double()
        self._chk(self.L.kgref_cross_section(self.h, i, j, l, float(E), C.byref(out)))
        return out.value

    def second_virial(self, i, j, T):
        out = C.c_double()
        self._chk(self.L.kgref_second_virial(self.h, i, j, float(T), C.byref(out)))
        return out.value

    def w_integral(self, i, j, T, l, r, regime=None):
        out = C.c_double()
        reg = regime or lattice_regime(T)
        self._chk(self.L.kgref_w_integral_tester(self.h, i, j, T, l, r, *reg, C.byref(out)))
        return out.value

    def w_integral_trace(self, i, j, T, l, r, regime=None, max_pts=200000):
        out = C.c_double()
        n = C.c_int()
        pts = np.zeros((max_pts, 3))
        reg = regime or lattice_regime(T)
        self._chk(self.L.kgref_w_integral_trace(self.h, i, j, T, l, r, *reg, C.byref(out), _p(pts), max_pts, C.byref(n)))
        return out.value, pts[:n.value]

    def _mat(self, f, rho, T, x):
        x = np.ascontiguousarray(x, float)
        out = np.empty((self.nc, self.nc))
        self._chk(f(self.h, rho, T, _p(x), _p(out)))
        return out

    def mtl(self, rho, T, x):
        return self._mat(self.L.kgref_mtl, rho, T, x)

    def etl(self, rho, T, x):
        return self._mat(self.L.kgref_etl, rho, T, x)

    def rdf(self, rho, T, x):
        return self._mat(self.L.kgref_rdf, rho, T, x)

    def K_factors(self, rho, T, x):
        x = np.ascontiguousarray(x, float)
        K, Kp = np.empty(self.nc), np.empty(self.nc)
        self._chk(self.L.kgref_K_factors(self.h, rho, T, _p(x), _p(K), _p(Kp)))
        return K, Kp

    def matrix_vector(self, which, rho, T, x, N):
        x = np.ascontiguousarray(x, float)
        dim = C.c_int()
        self._chk(self.L.kgref_matrix_vector(self.h, which, rho, T, _p(x), N, None, None, C.byref(dim)))
        d = dim.value
        M, v = np.empty((d, d)), np.empty(d)
        self._chk(self.L.kgref_matrix_vector(self.h, which, rho, T, _p(x), N, _p(M), _p(v), C.byref(dim)))
        return M, v

    def set_eos(self, dmudn=None, cp_id=None, p=0.0, vm=0.0, dvdn=None):
        nc_eos = 1 if self.single else self.nc
        a = [None if v is None else np.ascontiguousarray(v, float) for v in (dmudn, cp_id, dvdn)]
        self._eos_keep = a
        self.L.kgref_set_eos_tables_nc(self.h, nc_eos, None if a[0] is None else _p(a[0]), None if a[1] is None else _p(a[1]),
                                       p, vm, None if a[2] is None else _p(a[2]))

    def _scalar(self, f, T, Vm, x, N):
        x = np.ascontiguousarray(x, float)
        out = C.c_double()
        self._chk(f(self.h, T, Vm, _p(x), x.size, N, C.byref(out)))
        return out.value

    def viscosity(self, T, Vm, x, N=2):
        return self._scalar(self.L.kgref_viscosity, T, Vm, x, N)

    def thermal_conductivity(self, T, Vm, x, N=2):
        return self._scalar(self.L.kgref_thermal_conductivity, T, Vm, x, N)

    def thermal_conductivity_contribs(self, T, Vm, x, N=2):
        x = np.ascontiguousarray(x, float)
        out = np.empty(3)
        self._chk(self.L.kgref_thermal_conductivity_contribs(self.h, T, Vm, _p(x), x.size, N, _p(out)))
        return out

    def interdiffusion(self, T, Vm, x, N=2, frame=0, dep=-1, solvent=-1, do_compress=True):
        x = np.ascontiguousarray(x, float)
        out = np.empty(self.nc * self.nc)
        r, c = C.c_int(), C.c_int()
        self._chk(self.L.kgref_interdiffusion(self.h, T, Vm, _p(x), x.size, N, frame, dep, solvent, int(do_compress), _p(out),
                                              C.byref(r), C.byref(c)))
        return out[:r.value * c.value].reshape(r.value, c.value).copy()

    def thermal_diffusion_coeff(self, T, Vm, x, N=2, frame=0, dep=-1, solvent=-1):
        x = np.ascontiguousarray(x, float)
        out = np.empty(self.nc)
        n = C.c_int()
        self._chk(self.L.kgref_thermal_diffusion_coeff(self.h, T, Vm, _p(x), x.size, N, frame, dep, solvent, _p(out), C.byref(n)))
        return out[:n.value].copy()


def _ref_thermal_diffusion_ratio(self, T, Vm, x, N=2):
    x = np.ascontiguousarray(x, float)
    out = np.empty(self.nc)
    n = C.c_int()
    self._chk(self.L.kgref_thermal_diffusion_ratio(self.h, T, Vm, _p(x), x.size, N, _p(out), C.byref(n)))
    return out[:n.value].copy()




def _ref_soret_coefficient(self, T, Vm, x, N=2, dep=-1):
    x = np.ascontiguousarray(x, float)
    out = np.empty(self.nc)
    n = C.c_int()
    self._chk(self.L.kgref_soret_coefficient(self.h, T, Vm, _p(x), x.size, N, dep, _p(out), C.byref(n)))
    return out[:n.value].copy()


RefMie.thermal_diffusion_ratio = _ref_thermal_diffusion_ratio
RefMie.soret_coefficient = _ref_soret_coefficient


class RefQuantumMie(RefMie):
    """The reference's QuantumMie through its vector constructor (cpp/QuantumMie.cpp:5-49); parameters mixed like
    pykingas/MieType.py (kij = lij = 0). Every call that needs the T-dependent coefficients first calls set_internals,
    as the reference's public property methods do."""

    def __init__(self, Mw, sigma, eps_div_k, la, lr, fh_order, is_idealgas=True, py_constants=True):
        single = len(Mw) == 1
        if single:
            Mw, sigma, eps_div_k, la, lr = [list(v) * 2 for v in (Mw, sigma, eps_div_k, la, lr)]
        self.nc = len(Mw)
        self.single = single
        kB, NA = (BOLTZMANN_PY, AVOGADRO_PY) if py_constants else (BOLTZMANN, AVOGADRO)
        self.m = np.array(Mw, float) * 1e-3 / NA
        s = np.array(sigma, float)
        e = np.array(eps_div_k, float) * kB
        self.sigma = np.ascontiguousarray(0.5 * (s[:, None] + s[None, :]))
        self.eps = np.ascontiguousarray(np.sqrt(e[None, :] * e[:, None]))
        la_, lr_ = np.array(la, float), np.array(lr, float)
        self.la = np.ascontiguousarray(3 + np.sqrt((la_[:, None] - 3) * (la_[None, :] - 3)))
        self.lr = np.ascontiguousarray(3 + np.sqrt((lr_[:, None] - 3) * (lr_[None, :] - 3)))
        self.L = lib()
        self.h = self.L.kgref_qmie_create(self.nc, _p(self.m), _p(self.sigma), _p(self.eps), _p(self.la), _p(self.lr), int(fh_order),
                                          int(is_idealgas), int(single))

    def set_internals(self, T, rho=1e20):
        self._chk(self.L.kgref_set_internals(self.h, rho, T))

    def c_eff(self, i, j):
        out = np.zeros(6)
        n = C.c_int()
        self._chk(self.L.kgref_c_eff(self.h, i, j, _p(out), C.byref(n)))
        return out[:n.value]

    def omega(self, i, j, l, r, T, model=False):
        self.set_internals(T)
        return super().omega(i, j, l, r, T, model=model)

    def chi(self, i, j, T, g, b):
        self.set_internals(T)
        return super().chi(i, j, T, g, b)


class RefGeneric(RefMie):
    """The reference's non-power-law models: kind 1 LJSpline, 2 LJTS, 3 ModTangToennis, 4 HFD_B2, 5 Patowski, with
    fh_order > 0 selecting the FH_Corrected class (cpp/multiparam.h:36-131). Single-component, classical omega."""

    def __init__(self, kind, comps="", fh_order=0):
        self.L = lib()
        err = C.create_string_buffer(512)
        self.h = self.L.kgref_generic_create(kind, comps.encode(), fh_order, err, 512)
        if not self.h:
            raise RuntimeError(err.value.decode())
        self.nc, self.single, self.kind, self.fh_order = 2, True, kind, fh_order      # single-component: two identical species inside
        sc = np.zeros(4)
        self._chk(self.L.kgref_generic_scales(self.h, _p(sc)))
        self.sigma0, self.eps0, self.m0, self.fh_D = sc

    def set_internals(self, T):
        self._chk(self.L.kgref_set_internals(self.h, 0.0, T))

    def potential(self, r, T=None):
        if T is not None:
            self.set_internals(T)
        return np.array([self.L.kgref_potential(self.h, 0, 0, float(x)) for x in np.atleast_1d(r)])

    def potential_r(self, r, T=None):
        if T is not None:
            self.set_internals(T)
        return np.array([self.L.kgref_potential_r(self.h, 0, 0, float(x)) for x in np.atleast_1d(r)])

    def omega(self, i, j, l, r, T, model=False):
        self.set_internals(T)
        return super().omega(i, j, l, r, T, model=model)

    def chi(self, i, j, T, g, b):
        self.set_internals(T)
        return super().chi(i, j, T, g, b)

    def get_R(self, i, j, T, g, b):
        self.set_internals(T)
        return super().get_R(i, j, T, g, b)

    def w_integral(self, i, j, T, l, r, regime=None):
        self.set_internals(T)
        return super().w_integral(i, j, T, l, r, regime)


class ShimStockMie(RefMie):
    """The stock MieKinGas of libkgshim.so (same objects as oracle/_ref, second arm of the drop-in comparison)."""

    def __init__(self, Mw, sigma, eps_div_k, la, lr, is_idealgas=False):
        self._lib_override = shim_lib()
        super().__init__(Mw, sigma, eps_div_k, la, lr, is_idealgas)


class ShimGpuMie(RefMie):
    """GpuMieKinGas (integration/gpu_miekingas.cpp): the reference's own property code, Omega / transfer lengths fed from the device."""

    def __init__(self, Mw, sigma, eps_div_k, la, lr, is_idealgas=False, device=0):
        self._lib_override = shim_lib()
        self._gpu, self._device = True, device
        super().__init__(Mw, sigma, eps_div_k, la, lr, is_idealgas)

    def counters(self, reset=False):
        out = (C.c_long * 3)()
        if self.L.kgshim_counters(self.h, out, int(reset)) != 0:
            raise RuntimeError("not a GpuMieKinGas handle")
        return {"cpu_theta_integrand_calls": out[0], "device_feeds": out[1], "omega_values_stored": out[2]}

    def set_gpu_eos(self, eos):
        """Install the device SAFT-VR Mie EoS (integration/gpu_saft_eos.h) through the reference's set_eos; `eos` is a
        kineticgas_b200.DeviceSaftVRMie whose parameters are handed over (the C++ side creates its own handle)."""
        sig, eps, la, lr = (np.ascontiguousarray(a, float) for a in eos.pair_matrices())
        m, cpt = np.ascontiguousarray(eos.m, float), np.ascontiguousarray(eos.cp_type, np.int32)
        cpp, mw = np.ascontiguousarray(eos.cp_par, float), np.ascontiguousarray(eos.mw, float)
        err = C.create_string_buffer(512)
        self.L.kgshim_set_gpu_eos.argtypes = [C.c_void_p, C.c_int, dp, dp, dp, dp, dp, ip, dp, dp, C.c_char_p, C.c_int]
        if self.L.kgshim_set_gpu_eos(self.h, eos.nc, _p(m), _p(sig), _p(eps), _p(la), _p(lr), cpt.ctypes.data_as(ip), _p(cpp), _p(mw), err, 512):
            raise RuntimeError(err.value.decode())

    def omega_table(self, T, l, r):
        T = np.ascontiguousarray(np.atleast_1d(T), float)
        l = np.ascontiguousarray(l, np.int32)
        r = np.ascontiguousarray(r, np.int32)
        npairs = 1 if self.single else self.nc * (self.nc + 1) // 2
        out = np.empty((T.size, npairs, l.size))
        self._chk(self.L.kgshim_omega_table(self.h, T.size, _p(T), l.size, l.ctypes.data_as(ip), r.ctypes.data_as(ip), _p(out)))
        return out
