"""Host-side mirror of the reference's model interface for the hot path (pykingas/MieKinGas.py, pykingas/MieType.py,
pykingas/py_KineticGas.py:180-715), backed by the C-ABI library (include/kgcuda.h).

Same method names, argument meaning and error behaviour as pykingas; every (T, Vm, x) argument additionally accepts
arrays -- the whole batch goes to the GPU in one call (kg_transport_batch). EoS-derived inputs come either from a
thermopack-like object (`set_eos`, the calls of cpp/eos_interface.h:143-181: chemical_potential_tv, idealenthalpysingle,
specific_volume) evaluated on the host per state point, or directly as arrays (`eos_inputs=`), which is what sweeps use.
"""
import ctypes as C
import warnings
import numpy as np

from . import _lib
from ._lib import dp, ip
from .model import DeviceModel, mie_desc, raise_on_status
from .fluids import MIE, BOLTZMANN, AVOGADRO, BOLTZMANN_PY, AVOGADRO_PY

FLT_EPS = 1e-12
STATUS_COMPOSITION = 128      # include/kgcuda.h KG_STATUS_COMPOSITION
FRAMES = {"CoM": 0, "CoN": 1, "CoV": 2, "solvent": 3, "zarate": 4, "zarate_x": 5, "zarate_w": 6}      # cpp/utils.h:37-45, KineticGas::frame_of_reference_map
PROP_VISCOSITY, PROP_CONDUCTIVITY, PROP_DIFFUSION, PROP_THERMAL_DIFFUSION = 1, 2, 4, 8


class TransportOpts(C.Structure):
    _fields_ = [("N", C.c_int32), ("is_idealgas", C.c_int32), ("omega_correlation_active", C.c_int32), ("frame_of_reference", C.c_int32),
                ("dependent_idx", C.c_int32), ("solvent_idx", C.c_int32), ("do_compress", C.c_int32), ("properties", C.c_uint32)]


class EosInputs(C.Structure):
    _fields_ = [("dmudn", dp), ("cp_ideal", dp), ("pressure", dp), ("dvdn", dp)]


class TransportOut(C.Structure):
    _fields_ = [(k, dp) for k in ("viscosity", "conductivity", "conductivity_contribs", "D", "DT", "rdf", "K", "K_prime",
                                  "cond_matrix", "cond_vector", "visc_matrix", "visc_vector", "thermal_diffusion_ratio", "soret_zarate")]


def _ptr(a):
    if a is None:
        return None
    if hasattr(a, "address"):                       # eos.DevicePtr: a device-resident array from kg_eos_tv_batch_dev
        return C.cast(C.c_void_p(a.address), dp)
    return a.ctypes.data_as(dp)


class TCache:
    """Device-resident per-temperature tables (kg_tcache): the GPU-side KineticGasCache."""

    def __init__(self, kin, T, N, properties, coeff_T=None):
        # holds the DeviceModel (not the MieKinGas object): no reference cycle, and the kg_model outlives its caches
        self.dev = kin.dev
        self.T = np.ascontiguousarray(np.atleast_1d(T), float)
        self.opts = kin._opts(N, properties)
        h = C.c_void_p()
        L = self.dev.L
        self.coeff_T = None if coeff_T is None else np.ascontiguousarray(coeff_T, float)
        self.dev._chk(L.kg_tcache_create(self.dev.h, C.byref(self.opts), self.T.size, _ptr(self.T), _ptr(self.coeff_T), C.byref(h)), "kg_tcache_create")
        self.h = h

    def lr(self):
        L = self.dev.L
        n = L.kg_tcache_lr(self.h, None, None)
        l, r = np.zeros(n, np.int32), np.zeros(n, np.int32)
        L.kg_tcache_lr(self.h, l.ctypes.data_as(ip), r.ctypes.data_as(ip))
        return l, r

    def fetch(self):
        """(omega [nT][npairs][nlr], mtl, etl [nT][npairs], is_correlation) on the host: what feeds cache.omega.store."""
        l, _ = self.lr()
        shp = (self.T.size, self.dev.npairs, l.size)
        om, ic = np.zeros(shp), np.zeros(shp, np.uint8)
        mtl, etl = np.zeros(shp[:2]), np.zeros(shp[:2])
        self.dev._chk(self.dev.L.kg_tcache_fetch(self.h, _ptr(om), _ptr(mtl), _ptr(etl), ic.ctypes.data_as(C.POINTER(C.c_uint8))),
                          "kg_tcache_fetch")
        return om, mtl, etl, ic

    def close(self):
        if getattr(self, "h", None):
            self.dev.L.kg_tcache_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class MieKinGas:
    """pykingas.MieKinGas on the GPU. Parameters from the compact table in fluids.py unless given explicitly."""

    def __init__(self, comps, mole_weights=None, sigma=None, eps_div_k=None, la=None, lr=None, lij=0, kij=0,
                 N=4, is_idealgas=False, use_eos=None, device=0):
        self.comps = comps.split(",")
        self.ncomps = len(self.comps)
        self._is_singlecomp = self.ncomps == 1
        if self._is_singlecomp:
            self.comps = [self.comps[0]] * 2
            self.ncomps = 2
        self.default_N = N
        self.is_idealgas = is_idealgas
        self.lij, self.kij = lij, kij

        def pick(given, col):
            db = [MIE[c][col] for c in self.comps]
            if given is None:
                return np.array(db, float)
            given = list(given) * (2 if (self._is_singlecomp and len(given) == 1) else 1)
            return np.array([db[i] if given[i] is None else given[i] for i in range(self.ncomps)], float)

        Mw, sig, epsk, la_, lr_ = (pick(v, k) for k, v in enumerate((mole_weights, sigma, eps_div_k, la, lr)))
        self.mole_weights = Mw * 1e-3 / AVOGADRO_PY            # pykingas/py_KineticGas.py:61 (scipy's Avogadro)
        self.m = self.mole_weights.copy()
        nc = self.ncomps
        off = np.ones((nc, nc)) - np.identity(nc)
        # mixing rules: pykingas/MieType.py:201-249
        self.sigma_ij = (np.ones((nc, nc)) - lij * off) * 0.5 * (sig[None, :] + sig[:, None])
        eps = epsk * BOLTZMANN_PY                               # pykingas/MieType.py:214 (scipy's Boltzmann)
        self.epsilon_ij = (np.ones((nc, nc)) - kij * off) * np.sqrt(eps[None, :] * eps[:, None])
        self.la = 3 + (1 - lij) * np.sqrt((la_[None, :] - 3) * (la_[:, None] - 3))
        self.lr = 3 + (1 - lij) * np.sqrt((lr_[None, :] - 3) * (lr_[:, None] - 3))
        self.sigma, self.epsilon = self.sigma_ij, self.epsilon_ij
        self.eos = use_eos
        if isinstance(use_eos, str) and use_eos == "device":
            # the EoS pykingas would construct (saftvrmie(comps), pykingas/MieKinGas.py:55-60), evaluated on the GPU; with explicitly given
            # potential parameters the EoS takes the RET model's, segment numbers kept (pykingas/MieKinGas.py:46-65, MieType.py:167-191)
            from .eos import DeviceSaftVRMie
            self.eos = DeviceSaftVRMie(comps, device=device)
            if any(v is not None for v in (mole_weights, sigma, eps_div_k, la, lr)):
                for i in range(self.eos.nc):
                    self.eos.set_pure_fluid_param(i + 1, self.eos.m[i], self.sigma_ij[i, i], self.epsilon_ij[i, i] / BOLTZMANN_PY, self.la[i, i], self.lr[i, i])
        self.omega_correlation_active = True
        self._device = device
        self._build_device_model()

    # ------------------------------------------------------------------------------------------------ internals
    def _build_device_model(self):
        ncd = 1 if self._is_singlecomp else self.ncomps
        descs = [mie_desc(self.sigma_ij[i, j], self.epsilon_ij[i, j], self.la[i, j], self.lr[i, j]) for i in range(ncd) for j in range(i, ncd)]
        self.dev = DeviceModel(descs, self.m[:ncd], self._is_singlecomp, device=self._device)
        self._caches = {}

    def _opts(self, N, properties, frame=1, dependent_idx=-1, solvent_idx=-1, do_compress=True):
        o = TransportOpts()
        o.N, o.is_idealgas, o.omega_correlation_active = int(N), int(self.is_idealgas), int(self.omega_correlation_active)
        o.frame_of_reference, o.dependent_idx, o.solvent_idx, o.do_compress, o.properties = frame, dependent_idx, solvent_idx, int(do_compress), properties
        return o

    def set_eos(self, eos):
        self.eos = eos

    def _coeff_T(self, T):
        """Per-temperature effective coefficients [nT][npairs][6] of T-dependent potentials; None for plain Mie."""
        return None

    def omega_table(self, T, l, r):
        """Additive batched entry point: Omega^(l,r)_ij by numerical quadrature (Spherical::omega, no
        correlation) for every T, pair and (l,r): [nT][npairs][nlr], SI units."""
        T = np.atleast_1d(np.asarray(T, float))
        return self.dev.omega_batch(T, l, r, coeff_T=self._coeff_T(T))

    def set_omega_correlation_active(self, active):
        """cpp/MieKinGas.cpp:150-155 -- also drops the device tables, like cache.omega.clear()."""
        if active != self.omega_correlation_active:
            self._caches = {}
        self.omega_correlation_active = bool(active)

    def validate_and_sanitize_composition(self, x, check_sum=True):
        """pykingas/py_KineticGas.py:119-141, vectorised over a leading batch axis. check_sum=False: the state kernel makes the unity check
        per state point (KG_STATUS_COMPOSITION) -- a numpy reduction over the short last axis of 10^5 compositions costs 1-2.5 ms of host
        time, a quarter of the whole batch call."""
        x = np.asarray(x, float)
        if check_sum and np.any(np.abs(x.sum(axis=-1) - 1) > FLT_EPS):
            warnings.warn("Mole fractions do not sum to unity", RuntimeWarning, stacklevel=3)
        elif self._is_singlecomp:
            return np.broadcast_to(np.array([0.5, 0.5]), x.shape[:-1] + (2,))
        elif x.shape[-1] != self.ncomps:
            raise IndexError(f"{x.shape[-1]} mole fractions were supplied for a {self.ncomps} component mixture!")
        return x

    CACHE_LIMIT = 8          # device table sets kept per model (least recently used goes first)

    def tcache(self, T, N=None, properties=PROP_VISCOSITY | PROP_CONDUCTIVITY | PROP_DIFFUSION | PROP_THERMAL_DIFFUSION):
        """Build (or reuse) the device tables for the distinct temperatures in T. Like the reference's KineticGasCache the tables are
        shared by all properties: a set is keyed on (N, correlation switch, ideal-gas switch, temperatures), small sets are built
        for every property at once (viscosity followed by thermal_conductivity at the same T integrates nothing twice), a set that
        lacks a requested property is rebuilt with the union. At most CACHE_LIMIT sets are kept."""
        N = self.default_N if N is None else N
        Tu = np.unique(np.asarray(T, float))
        key = (N, self.omega_correlation_active, bool(self.is_idealgas), Tu.tobytes())
        c = self._caches.get(key)
        if c is not None and (properties & ~c.opts.properties) == 0:
            self._caches[key] = self._caches.pop(key)          # most recently used last
            return c
        props = properties | (c.opts.properties if c is not None else 0)
        if Tu.size <= 16:
            props |= PROP_VISCOSITY | PROP_CONDUCTIVITY | PROP_DIFFUSION | PROP_THERMAL_DIFFUSION
        if c is not None:
            self._caches.pop(key).close()
        self._caches[key] = TCache(self, Tu, N, props, self._coeff_T(Tu))
        while len(self._caches) > self.CACHE_LIMIT:
            self._caches.pop(next(iter(self._caches))).close()
        return self._caches[key]

    def _eos_arrays(self, T, Vm, x, need_dmudn, need_cp, need_dvdn, eos_inputs):
        """EoS inputs as arrays: given by the caller, or evaluated per state point on the host EoS object through the same
        calls the reference's PyWrapper makes (cpp/eos_interface.h:143-181)."""
        n = T.size
        nce = 1 if self._is_singlecomp else self.ncomps
        out = dict(eos_inputs or {})
        xe = np.ones((n, 1)) if self._is_singlecomp else x
        if hasattr(self.eos, "tv_batch"):
            # device EoS (kineticgas_b200.eos.DeviceSaftVRMie): every missing input of the whole batch in one GPU pass
            miss = {"dmudn": need_dmudn and "dmudn" not in out, "cp_ideal": need_cp and "cp_ideal" not in out, "dvdn": need_dvdn and "dvdn" not in out}
            if any(miss.values()):
                call = self.eos.tv_batch_dev if (n >= 4096 and hasattr(self.eos, "tv_batch_dev")) else self.eos.tv_batch
                out.update(call(T, Vm, xe, dmudn=miss["dmudn"], cp_ideal=miss["cp_ideal"], dvdn=miss["dvdn"]))
        if need_dmudn and "dmudn" not in out:
            if self.eos is None:
                raise RuntimeError("EoS is not set")
            out["dmudn"] = np.array([np.asarray(self.eos.chemical_potential_tv(T[k], Vm[k], list(xe[k]), None, None, True)[1]) for k in range(n)])
        if need_cp and "cp_ideal" not in out:
            if self.eos is None:
                raise RuntimeError("EoS is not set")
            out["cp_ideal"] = np.array([[self.eos.idealenthalpysingle(T[k], c + 1, True)[1] for c in range(nce)] for k in range(n)])
        if need_dvdn and "dvdn" not in out:
            if self.eos is None:
                raise RuntimeError("EoS is not set")
            rows = []
            for k in range(n):
                p, = self.eos.pressure_tv(T[k], Vm[k], list(xe[k]))
                rows.append(np.asarray(self.eos.specific_volume(T[k], p, list(xe[k]), self.eos.VAPPH, None, None, True)[1]))
            out["dvdn"] = np.array(rows)
        return {k: (v if hasattr(v, "address") else np.ascontiguousarray(np.asarray(v, float).reshape(n, -1))) for k, v in out.items()}

    def _run(self, T, Vm, x, N, properties, want, frame=1, dependent_idx=-1, solvent_idx=-1, do_compress=True, eos_inputs=None, cache=None):
        """One kg_transport_batch call. want: names of kg_transport_out fields to return. Returns dict of arrays + 'scalar' flag."""
        N = self.default_N if N is None else N
        T_in = np.asarray(T, float)
        scalar = T_in.ndim == 0 and np.ndim(Vm) == 0 and np.ndim(x) <= 1
        device_sum_check = (not self._is_singlecomp) and np.ndim(x) == 2 and np.shape(x)[0] >= 32768
        x = self.validate_and_sanitize_composition(x, check_sum=not device_sum_check)
        n = int(np.broadcast_shapes(T_in.shape, np.shape(Vm), x.shape[:-1])[0]) if not scalar else 1
        Tn = np.ascontiguousarray(np.broadcast_to(T_in, (n,)), float)
        Vn = np.ascontiguousarray(np.broadcast_to(np.asarray(Vm, float), (n,)), float)
        xn = np.ascontiguousarray(np.broadcast_to(x, (n, self.ncomps)), float)
        nc = self.ncomps
        nn = N * nc
        nD = nc - 1 if (do_compress or frame >= FRAMES["zarate"]) else nc
        shapes = {"viscosity": (n,), "conductivity": (n,), "conductivity_contribs": (n, 3), "D": (n, nD, nD), "DT": (n, nc),
                  "rdf": (n, nc, nc), "K": (n, nc), "K_prime": (n, nc), "cond_matrix": (n, nn, nn), "cond_vector": (n, nn),
                  "visc_matrix": (n, nn, nn), "visc_vector": (n, nn), "thermal_diffusion_ratio": (n, nc), "soret_zarate": (n, nc)}
        # large batches: results land in page-locked arrays (asynchronous D2H, overlapped chunk by chunk with the kernels)
        bufs = {k: (_lib.pinned_empty(shapes[k]) if n >= 32768 else np.zeros(shapes[k])) for k in want}
        out = TransportOut()
        for k, b in bufs.items():
            setattr(out, k, _ptr(b))
        opts = self._opts(N, properties, frame, dependent_idx, solvent_idx, do_compress)
        want_D = ("D" in want) or ("DT" in want) or ("thermal_diffusion_ratio" in want) or ("soret_zarate" in want)
        eos_arr = self._eos_arrays(Tn, Vn, xn, want_D, ("conductivity" in want) or ("conductivity_contribs" in want),
                                   frame == FRAMES["CoV"] and want_D, eos_inputs)
        eos = EosInputs()
        for k in ("dmudn", "cp_ideal", "pressure", "dvdn"):
            if k in eos_arr:
                setattr(eos, k, _ptr(eos_arr[k]))
        st = np.zeros(n, np.int32)
        if cache is None:
            cache = self.tcache(Tn, N, properties)
        xarg = None if self._is_singlecomp else _ptr(xn)
        self.dev._chk(self.dev.L.kg_transport_batch(self.dev.h, cache.h, n, _ptr(Tn), _ptr(Vn), xarg, C.byref(opts), C.byref(eos), C.byref(out),
                                                     st.ctypes.data_as(ip)), "kg_transport_batch")
        if st.max(initial=0) >= STATUS_COMPOSITION:             # warning bit of the state kernel
            if device_sum_check and (st & STATUS_COMPOSITION).any():
                warnings.warn("Mole fractions do not sum to unity", RuntimeWarning, stacklevel=3)
            st &= ~np.int32(STATUS_COMPOSITION)
        raise_on_status(st)
        bufs["scalar"] = scalar
        return bufs

    # ------------------------------------------------------------------------------------------------ TV-properties
    def viscosity(self, T, Vm, x, N=None, **kw):
        r = self._run(T, Vm, x, N, PROP_VISCOSITY, ["viscosity"], **kw)
        return float(r["viscosity"][0]) if r["scalar"] else r["viscosity"]

    def thermal_conductivity(self, T, Vm, x, N=None, idealgas=None, contributions="all", **kw):
        idealgas = self.is_idealgas if idealgas is None else idealgas
        if idealgas:
            Vm = np.full(np.shape(Vm), 1e12) if np.ndim(Vm) else 1e12
        r = self._run(T, Vm, x, N, PROP_CONDUCTIVITY, ["conductivity", "conductivity_contribs"], **kw)
        if contributions != "all":
            c = r["conductivity_contribs"]
            d = {k: (float(c[0, i]) if r["scalar"] else c[:, i]) for i, k in enumerate("tdi") if k in contributions}
            return d
        return float(r["conductivity"][0]) if r["scalar"] else r["conductivity"]

    def interdiffusion(self, T, Vm, x, N=None, use_independent=True, dependent_idx=-1, frame_of_reference="CoN", use_binary=True,
                       solvent_idx=-1, **kw):
        if frame_of_reference not in FRAMES:
            raise RuntimeError("Invalid frame of reference : " + frame_of_reference)
        r = self._run(T, Vm, x, N, PROP_DIFFUSION, ["D"], frame=FRAMES[frame_of_reference], dependent_idx=dependent_idx,
                      solvent_idx=solvent_idx, do_compress=use_independent, **kw)
        D = r["D"]
        if use_binary and self.ncomps == 2:
            D = D[:, 0, 0]
        return D[0] if r["scalar"] else D

    def thermal_diffusion_coeff(self, T, Vm, x, N=None, use_independent=False, dependent_idx=-1, frame_of_reference="CoN", solvent_idx=-1, **kw):
        N = self.default_N if N is None else N
        if N < 2:
            warnings.warn(f"Thermal diffusion is a 2nd order phenomena, cannot be computed for N < 2 (got N = {N})", RuntimeWarning, stacklevel=2)
            return np.full(self.ncomps, np.nan)
        if frame_of_reference not in FRAMES:
            raise RuntimeError("Invalid frame of reference : " + frame_of_reference)
        r = self._run(T, Vm, x, N, PROP_THERMAL_DIFFUSION, ["DT"], frame=FRAMES[frame_of_reference], dependent_idx=dependent_idx,
                      solvent_idx=solvent_idx, **kw)
        DT = r["DT"]
        if frame_of_reference == "zarate":          # Ncomps - 1 independent coefficients (cpp/KineticGas_properties.cpp:262-277)
            DT = DT[:, :self.ncomps - 1]
        return DT[0] if r["scalar"] else DT

    def thermal_diffusion_ratio(self, T, Vm, x, N=None, **kw):
        """k_T,i (pykingas/py_KineticGas.py:279-315 -> KineticGas::thermal_diffusion_ratio, cpp/KineticGas_properties.cpp:282-312)."""
        N = self.default_N if N is None else N
        if N < 2:
            warnings.warn(f"Thermal diffusion is a 2nd order phenomena, cannot be computed for N < 2 (got N = {N})", RuntimeWarning, stacklevel=2)
            return np.full(self.ncomps, np.nan)
        r = self._run(T, Vm, x, N, PROP_THERMAL_DIFFUSION, ["thermal_diffusion_ratio"], frame=FRAMES["CoM"], do_compress=False, **kw)
        return r["thermal_diffusion_ratio"][0] if r["scalar"] else r["thermal_diffusion_ratio"]

    def thermal_diffusion_factor(self, T, Vm, x, N=None, **kw):
        """alpha_ij = k_T,i - k_T,j (cpp/KineticGas_properties.cpp:314-323)."""
        kT = np.asarray(self.thermal_diffusion_ratio(T, Vm, x, N=N, **kw))
        return kT[..., :, None] - kT[..., None, :]

    def soret_coefficient(self, T, Vm, x, N=None, use_zarate=True, dependent_idx=-1, **kw):
        """pykingas/py_KineticGas.py:347-399. use_zarate (the reference's default): S_T = D^-1 D_T in the Zarate frame, Ncomps - 1
        entries, ideal gas only (KineticGas::soret_coefficient, cpp/KineticGas_properties.cpp:325-330); else alpha_T / T."""
        if use_zarate:
            N = self.default_N if N is None else N
            if N < 2:
                warnings.warn(f"Thermal diffusion is a 2nd order phenomena, cannot be computed for N < 2 (got N = {N})", RuntimeWarning, stacklevel=2)
                return np.full((self.ncomps, self.ncomps), np.nan)
            if not self.is_idealgas:
                raise RuntimeError("Zarate relation for Soret coefficient only implemented for ideal gas!")
            r = self._run(T, Vm, x, N, PROP_DIFFUSION | PROP_THERMAL_DIFFUSION, ["soret_zarate"], frame=FRAMES["zarate"],
                          dependent_idx=dependent_idx, **kw)
            S = r["soret_zarate"][:, :self.ncomps - 1]
            return S[0] if r["scalar"] else S
        a = self.thermal_diffusion_factor(T, Vm, x, N=N, **kw)
        T = np.asarray(T, float)
        return a / (T[..., None, None] if T.ndim else T)

    def transport_batch(self, T, Vm, x, N=None, properties=PROP_VISCOSITY | PROP_CONDUCTIVITY | PROP_DIFFUSION | PROP_THERMAL_DIFFUSION,
                        frame_of_reference="CoN", dependent_idx=-1, solvent_idx=-1, use_independent=True, eos_inputs=None, cache=None):
        """All requested properties of a batch of state points in one device pass (additive batched entry point)."""
        want = []
        if properties & PROP_VISCOSITY:
            want.append("viscosity")
        if properties & PROP_CONDUCTIVITY:
            want += ["conductivity", "conductivity_contribs"]
        if properties & PROP_DIFFUSION:
            want.append("D")
        if properties & PROP_THERMAL_DIFFUSION:
            want.append("DT")
        r = self._run(T, Vm, x, N, properties, want, frame=FRAMES[frame_of_reference], dependent_idx=dependent_idx, solvent_idx=solvent_idx,
                      do_compress=use_independent, eos_inputs=eos_inputs, cache=cache)
        r.pop("scalar")
        return r

    # ------------------------------------------------------------------------------------------------ Tp-properties
    def _vm(self, T, p, x):
        if self.eos is None:
            raise RuntimeError("EoS is not set")
        if np.ndim(T) == 0 and np.ndim(p) == 0:
            Vm, = self.eos.specific_volume(T, p, x, self.eos.VAPPH)      # assuming vapour, pykingas/py_KineticGas.py:642-715
            return Vm
        T, p = np.broadcast_arrays(np.asarray(T, float), np.asarray(p, float))
        xx = np.broadcast_to(np.asarray(x, float), T.shape + (np.shape(x)[-1],))
        if hasattr(self.eos, "specific_volume_batch"):
            return self.eos.specific_volume_batch(T, p, np.ones(T.shape + (1,)) if self._is_singlecomp else xx)
        return np.array([self.eos.specific_volume(T[k], p[k], list(xx[k]), self.eos.VAPPH)[0] for k in range(T.size)])

    def viscosity_tp(self, T, p, x, N=None):
        return self.viscosity(T, self._vm(T, p, x), x, N=N)

    def thermal_conductivity_tp(self, T, p, x, N=None, contributions="all"):
        return self.thermal_conductivity(T, self._vm(T, p, x), x, N=N, contributions=contributions)

    def interdiffusion_tp(self, T, p, x, N=None, use_independent=True, dependent_idx=-1, frame_of_reference="CoN", use_binary=True, solvent_idx=-1):
        return self.interdiffusion(T, self._vm(T, p, x), x, N=N, use_independent=use_independent, dependent_idx=dependent_idx,
                                   frame_of_reference=frame_of_reference, use_binary=use_binary, solvent_idx=solvent_idx)

    def thermal_diffusion_coeff_tp(self, T, p, x, N=None, use_independent=False, dependent_idx=-1, frame_of_reference="CoN", solvent_idx=-1):
        return self.thermal_diffusion_coeff(T, self._vm(T, p, x), x, N=N, use_independent=use_independent, dependent_idx=dependent_idx,
                                            frame_of_reference=frame_of_reference, solvent_idx=solvent_idx)

    # ------------------------------------------------------------------------------------------------ internals exposed by the bindings
    def _rho_call(self, particle_density, T, x, N, properties, want):
        Vm = AVOGADRO / np.asarray(particle_density, float)
        r = self._run(T, Vm, x, N, properties, want)
        return r[want[0]][0] if r["scalar"] else r[want[0]]

    def get_rdf(self, particle_density, T, x):
        return self._rho_call(particle_density, T, x, 2, PROP_VISCOSITY, ["rdf"])

    def get_K_factors(self, particle_density, T, x):
        return self._rho_call(particle_density, T, x, 2, PROP_VISCOSITY, ["K"])

    def get_K_prime_factors(self, particle_density, T, x):
        return self._rho_call(particle_density, T, x, 2, PROP_VISCOSITY, ["K_prime"])

    def get_conductivity_matrix(self, particle_density, T, mole_fracs, N=None):
        return self._rho_call(particle_density, T, mole_fracs, N, PROP_DIFFUSION, ["cond_matrix"])

    def get_conductivity_vector(self, particle_density, T, mole_fracs, N=None):
        return self._rho_call(particle_density, T, mole_fracs, N, PROP_DIFFUSION, ["cond_vector"])

    def get_viscosity_matrix(self, particle_density, T, mole_fracs, N=None):
        return self._rho_call(particle_density, T, mole_fracs, N, PROP_VISCOSITY, ["visc_matrix"])

    def get_viscosity_vector(self, particle_density, T, mole_fracs, N=None):
        return self._rho_call(particle_density, T, mole_fracs, N, PROP_VISCOSITY, ["visc_vector"])

    def _pair_matrix(self, v):
        if self._is_singlecomp:
            return np.full((2, 2), v[0])
        out = np.zeros((self.ncomps, self.ncomps))
        k = 0
        for i in range(self.ncomps):
            for j in range(i, self.ncomps):
                out[i, j] = out[j, i] = v[k]
                k += 1
        return out

    def get_mtl(self, particle_density, T, x):
        return self._pair_matrix(self.dev.transfer_lengths([T], coeff_T=self._coeff_T([T]))[0][0])

    def get_etl(self, particle_density, T, x):
        return self._pair_matrix(self.dev.transfer_lengths([T], coeff_T=self._coeff_T([T]))[1][0])


class QuantumMie(MieKinGas):
    """pykingas.QuantumMie (Feynman-Hibbs corrected Mie, cpp/QuantumMie.cpp, cpp/ExtendedSutherland.cpp) on the GPU.

    u(r;T) = eps sum_k C_k beta^(bexp_k) (sigma/r)^lam_k with beta = eps/kT: 2, 4 or 6 power-law terms for FH order 0, 1, 2
    (cpp/QuantumMie.cpp:73-109). The device kernels take the terms' exponents from the model and the T-dependent
    effective coefficients C_eff(T) = C_k beta^bexp_k (cpp/ExtendedSutherland.cpp:182-200) per temperature. Collision
    integrals are always integrated (ExtSutherland has no correlation branch). is_idealgas=False runs the Revised Enskog
    assembly with the extended-Sutherland RDF at contact (ExtSutherland::saft_rdf, cpp/ExtendedSutherland.cpp:315-333) on the device."""

    def __init__(self, comps, mole_weights=None, sigma=None, eps_div_k=None, la=None, lr=None, lij=0, kij=0,
                 N=4, FH_orders=None, is_idealgas=False, use_eos=None, device=0):
        from .fluids import MIE_FH, HBAR
        ids = comps.split(",")
        if isinstance(FH_orders, int):
            FH_orders = [FH_orders] * len(ids)
        if FH_orders is None:
            FH_orders = [0] * len(ids)          # database default_fh_order is 0 for HE, H2 (fluids/HE.json:117, fluids/H2.json:128)
        if len(set(FH_orders)) != 1:
            raise RuntimeError("All FH orders must be equal!")        # cpp/QuantumMie.cpp:39-41
        self.FH_order = int(FH_orders[0])

        def db(col):
            return [(MIE_FH[(c, self.FH_order)][col] if self.FH_order > 0 else MIE[c][col + 1]) for c in ids]

        def merged(given, col):
            base = db(col)
            return base if given is None else [base[i] if given[i] is None else given[i] for i in range(len(ids))]

        self._HBAR = HBAR
        super().__init__(comps, mole_weights=mole_weights, sigma=merged(sigma, 0), eps_div_k=merged(eps_div_k, 1),
                         la=merged(la, 2), lr=merged(lr, 3), lij=lij, kij=kij, N=N, is_idealgas=is_idealgas, use_eos=use_eos, device=device)
        self.omega_correlation_active = False

    def _build_device_model(self):
        from .model import power_sum_desc
        nt = 2 * (self.FH_order + 1)
        ncd = 1 if self._is_singlecomp else self.ncomps
        self._C, self._bexp, descs = [], [], []
        for i in range(ncd):
            for j in range(i, ncd):
                la, lr, sig, eps = self.la[i, j], self.lr[i, j], self.sigma_ij[i, j], self.epsilon_ij[i, j]
                C0 = (lr / (lr - la)) * (lr / la) ** (la / (lr - la))
                q1 = lambda lam: lam * (lam - 1)                                   # cpp/QuantumMie.h:52-58
                q2 = lambda lam: 0.5 * (lam + 2) * (lam + 1) * q1(lam)
                mu_inv = 1.0 / self.m[i] + 1.0 / self.m[j]
                D1 = self._HBAR ** 2 * (mu_inv if self.FH_order > 0 else 0.0) / 24.0
                D2 = self._HBAR ** 2 * (mu_inv if self.FH_order > 1 else 0.0) / 24.0
                Ck = [C0, -C0, C0 * D1 * q1(lr) * sig ** -2 * eps ** -1, -C0 * D1 * q1(la) * sig ** -2 * eps ** -1,
                      C0 * D2 ** 2 * q2(lr) * sig ** -4 * eps ** -2, -C0 * D2 ** 2 * q2(la) * sig ** -4 * eps ** -2][:nt]
                lam = [lr, la, lr + 2.0, la + 2.0, lr + 4.0, la + 4.0][:nt]
                self._C.append(Ck)
                self._bexp.append([0, 0, 1, 1, 2, 2][:nt])
                descs.append(power_sum_desc(sig, eps, Ck, lam))
        self._pair_eps = [d.eps for d in descs]
        self.dev = DeviceModel(descs, self.m[:ncd], self._is_singlecomp, device=self._device)
        if self.FH_order > 0:
            self.dev.set_temperature_dependent(True)       # every call must carry the per-temperature coefficients (coeff_T)
        self._caches = {}

    def _coeff_T(self, T):
        T = np.atleast_1d(np.asarray(T, float))
        out = np.zeros((T.size, len(self._C), _lib.KG_MAX_TERMS))
        for p, (Ck, be, eps) in enumerate(zip(self._C, self._bexp, self._pair_eps)):
            beta = eps / (BOLTZMANN * T)
            for k, (c, b) in enumerate(zip(Ck, be)):
                out[:, p, k] = c * beta ** b if b else c
        return out

    def potential(self, i, j, r, T):
        """u_ij(r; T) [J] (ExtSutherland::potential(i, j, r, T, rho), cpp/ExtendedSutherland.cpp:121-130)."""
        p = self.dev.pair_index(0 if self._is_singlecomp else i, 0 if self._is_singlecomp else j)
        cf = self._coeff_T([T])[0, p]
        d = self.dev.descs[p]
        return d.eps * sum(cf[k] * (d.sigma / r) ** d.lam[k] for k in range(d.nterms))

    def get_sigma_eff(self, T):
        """Root of u(r;T) by the reference's Newton iteration (cpp/ExtendedSutherland.cpp:202-217), host side utility."""
        out = np.zeros((self.ncomps, self.ncomps))
        for i in range(self.ncomps):
            for j in range(i, self.ncomps):
                s = self.sigma_ij[i, j]
                while True:
                    h = s * 1e-7
                    u = self.potential(i, j, s, T)
                    ur = (self.potential(i, j, s + h, T) - self.potential(i, j, s - h, T)) / (2 * h)
                    s -= u / ur
                    if abs(u) / self.epsilon_ij[i, j] <= 1e-6:
                        break
                out[i, j] = out[j, i] = s
        return out

    def get_mtl(self, particle_density, T, x):
        return self._pair_matrix(self.dev.transfer_lengths([T], coeff_T=self._coeff_T([T]))[0][0])

    def get_etl(self, particle_density, T, x):
        return self._pair_matrix(self.dev.transfer_lengths([T], coeff_T=self._coeff_T([T]))[1][0])


class _GenericSpherical(MieKinGas):
    """Single-component models on the non-power-law potentials of include/kgcuda.h (kinds 1-5). The reference classes
    (cpp/LJSpline.h, cpp/LJTS.h, cpp/multiparam.h) are single-component; collision integrals always come from the
    quadrature (Quantum::omega forwards to Spherical::omega when quantum_is_active is false, cpp/Quantum.cpp:1261-1268).
    Dilute-gas (is_idealgas) properties run through the same device tables as MieKinGas."""
    KIND = None

    def _setup(self, comp_id, Mw, sigma, eps_div_k, params, FH_order, N, device, use_eos):
        from .fluids import HBAR
        self.comps = [comp_id, comp_id]
        self.ncomps, self._is_singlecomp = 2, True
        self.default_N, self.is_idealgas = N, True
        self.lij = self.kij = 0
        self.m = np.array([Mw, Mw], float) * 1e-3 / AVOGADRO
        self.mole_weights = self.m.copy()
        self.sigma_ij = np.full((2, 2), float(sigma))
        self.epsilon_ij = np.full((2, 2), float(eps_div_k) * BOLTZMANN)
        self.sigma, self.epsilon = self.sigma_ij, self.epsilon_ij
        self.eos = use_eos
        self.omega_correlation_active = False
        self.FH_order = int(FH_order)
        self._params = list(params)
        self._fh_D = HBAR ** 2 / (24.0 * (self.m[0] * self.m[0] / (self.m[0] + self.m[0])))    # cpp/extentions.h:308-314
        self._device = device
        self._build_device_model()

    def _build_device_model(self):
        from .model import generic_desc
        d = generic_desc(self.KIND, self.sigma_ij[0, 0], self.epsilon_ij[0, 0], self._params, self.FH_order,
                         self._fh_D if self.FH_order > 0 else 0.0)
        self.dev = DeviceModel([d], self.m[:1], True, device=self._device)
        self._caches = {}

    def set_FH_order(self, order):
        """FH_Corrected::set_FH_order (cpp/extentions.h:316-321): drops every cached table."""
        if int(order) != self.FH_order:
            self.FH_order = int(order)
            self._build_device_model()
            self.set_quantum_active(self._quantum_active)

    # ---- quantum-mechanical collision integrals (pykingas/Quantum.py:125-135, cpp/Quantum.cpp:1266-1268, 1520-1528)
    _quantum_active = False
    HALF_SPIN, ROT_GROUND_STATE = 0, 0            # fluids/*.json "Quantum" blocks of HE, NE, AR: spin-0 bosons in the rotational ground state

    def symmetry_weights(self):
        """Quantum::get_symmetry_weights (cpp/Quantum.cpp:208-223) of the like pair."""
        s, j = self.HALF_SPIN / 2.0, self.ROT_GROUND_STATE
        w = [((s + 1) * (j + 1) + s * j) / ((2 * s + 1) * (2 * j + 1)), ((s + 1) * j + s * (j + 1)) / ((2 * s + 1) * (2 * j + 1))]
        return w[::-1] if self.HALF_SPIN % 2 else w

    def set_quantum_active(self, active):
        """While active the collision integrals of every property come from partial-wave phase shifts (Quantum::quantum_omega) instead of
        classical trajectories. Not available for Feynman-Hibbs corrected potentials (the reference classes FH_* are classical)."""
        active = bool(active) and self.FH_order == 0
        ws, wa = self.symmetry_weights()
        a, b = np.array([ws], float), np.array([wa], float)
        if getattr(self, "dev", None) is not None:          # host-only descriptions (tests of the descriptor blocks) have no device model
            self.dev._chk(self.dev.L.kg_model_set_quantum(self.dev.h, int(active), _ptr(a), _ptr(b)), "kg_model_set_quantum")
        if active != self._quantum_active:
            self._caches = {}
        self._quantum_active = active

    def get_quantum_active(self):
        return self._quantum_active

    def set_JKWB_limits(self, E, l):
        """Quantum::set_JKWB_limits (cpp/Quantum.h:128-130): JKWB phase shifts for E / eps > E or angular momentum >= l."""
        self.dev._chk(self.dev.L.kg_model_set_jkwb_limits(self.dev.h, float(E), int(l)), "kg_model_set_jkwb_limits")
        self._caches = {}

    def omega(self, i, j, l, r, T):
        """Quantum::omega: quantum or classical by the active switch."""
        if self._quantum_active:
            ws, wa = self.symmetry_weights()
            return float(self.dev.quantum_omega(0, [T], [l], [r], ws, wa)[0, 0])
        return float(self.dev.omega_batch([T], [l], [r])[0, 0, 0])

    def cross_section(self, i, j, n, E_red):
        """Quantum::cross_section at reduced energy E / eps (quantum engine)."""
        ws, wa = self.symmetry_weights()
        return float(self.dev.quantum_cross_sections(0, [n], [E_red], ws, wa)[0, 0])

    def phase_shift(self, i, j, l, E_red):
        return float(self.dev.quantum_phase_shifts(0, [l], [E_red])[0])

    def potential(self, i, j, r, T=300.0):
        return self.dev.potential_points(0, T, r)[0]

    def potential_derivative_r(self, i, j, r, T=300.0):
        return self.dev.potential_points(0, T, r)[1]

    def chi(self, i, j, T, g, b):
        """Spherical::chi (cpp/Spherical.cpp:455-465); b in metres like the reference binding."""
        return self.dev.chi_points(0, T, g, np.asarray(b, float) / self.sigma_ij[0, 0])

    def get_R(self, i, j, T, g, b):
        return self.dev.chi_points(0, T, g, np.asarray(b, float) / self.sigma_ij[0, 0], want_R=True)[1] * self.sigma_ij[0, 0]


class LJSpline(_GenericSpherical):
    """pykingas.LJSpline collision dynamics (cpp/LJSpline.h:13-30: argon-like sigma, eps, mass are fixed by the class)."""
    KIND = 1

    def __init__(self, N=4, is_idealgas=False, device=0, use_eos=None):
        self._setup("LJS", 6.64e-26 * AVOGADRO * 1e3, 3.42e-10, 1.71200476e-21 / BOLTZMANN, [], 0, N, device, use_eos)
        # RET for the LJ/s fluid: own RDF at contact (cpp/LJSpline.h:89-125) and the omega correlation of LJSpline::omega
        # (cpp/LJSpline.cpp:123-134), which the reference applies unconditionally
        self.is_idealgas = is_idealgas
        self.omega_correlation_active = True


class LJTS(_GenericSpherical):
    """cpp/LJTS.h: Lennard-Jones truncated at 2.5 sigma and shifted, dilute gas only."""
    KIND = 2

    def __init__(self, mole_weight=None, sigma=3.42e-10, eps_div_k=None, N=4, device=0, use_eos=None):
        # defaults: the argon-like values pykingas passes (same numbers as cpp/LJSpline.h:16-18)
        mole_weight = 6.64e-26 * AVOGADRO * 1e3 if mole_weight is None else mole_weight
        eps_div_k = 1.71200476e-21 / BOLTZMANN if eps_div_k is None else eps_div_k
        self._setup("LJTS", mole_weight, sigma, eps_div_k, [], 0, N, device, use_eos)


class ModTangToennis(_GenericSpherical):
    """pykingas.multiparam.ModTangToennis / FH_ModTangToennies (cpp/multiparam.cpp:5-89): Ar, Ne, He."""
    KIND = 3

    def __init__(self, comps, FH_order=0, N=4, device=0, use_eos=None, quantum_active=True):
        from .fluids import TANG_TOENNIES
        from .model import tang_toennies_params
        p = TANG_TOENNIES[comps]
        self._setup(comps, MIE[comps][0], p["sigma"], p["eps_div_k"], tang_toennies_params(p), FH_order, N, device, use_eos)
        self.set_quantum_active(quantum_active)          # pykingas/multiparam.py:94-104: quantum by default


class HFD_B2(_GenericSpherical):
    """pykingas.multiparam.HFD_B2 / FH_HFD_B2 (cpp/multiparam.cpp:91-182): helium."""
    KIND = 4

    def __init__(self, comps="HE", FH_order=0, N=4, device=0, use_eos=None, quantum_active=True):
        from .fluids import HFD_B2 as TAB
        from .model import hfd_b2_params
        p = TAB[comps]
        self._setup(comps, MIE[comps][0], p["sigma"], p["eps_div_k"], hfd_b2_params(p), FH_order, N, device, use_eos)
        self.set_quantum_active(quantum_active)          # pykingas/multiparam.py:126-129


class Patowski(_GenericSpherical):
    """pykingas.multiparam.Patowski / PatowskiFH (cpp/multiparam.cpp:184-279): hydrogen."""
    KIND = 5

    def __init__(self, comps="H2", FH_order=0, N=4, device=0, use_eos=None, quantum_active=True):
        from .fluids import PATKOWSKI
        from .model import patkowski_params
        p = PATKOWSKI[comps]
        self._setup(comps, MIE[comps][0], p["sigma"], p["eps_div_k"], patkowski_params(p), FH_order, N, device, use_eos)
        self.set_quantum_active(quantum_active)
