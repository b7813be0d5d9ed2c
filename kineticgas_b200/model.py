"""Thin object layer over the C-ABI (include/kgcuda.h): a device model (one per KineticGas object) and an
omega plan (a batched sweep resident in HBM)."""
import ctypes as C
import numpy as np
from . import _lib
from ._lib import PotentialDesc, PotentialDescEx, dp, ip

STATUS_TEXT = {1: "get_R reached max iter!", 2: "tanh-sinh node table exhausted", 4: "non-finite collision integral",
               8: "Unable to converge bmid!", 16: "singular Sonine-expansion system", 32: "temperature not present in the device cache"}


def _p(a):
    return None if a is None else a.ctypes.data_as(dp)


def _pi(a):
    return None if a is None else a.ctypes.data_as(ip)


def power_sum_desc(sigma, eps, C_, lam):
    d = PotentialDesc()
    d.kind = 0
    d.nterms = len(C_)
    d.sigma = sigma
    d.eps = eps
    for k, (c, l) in enumerate(zip(C_, lam)):
        d.C[k] = c
        d.lam[k] = l
    return d


def mie_desc(sigma, eps, la, lr):
    """Mie block: u = C eps [(sigma/r)^lr - (sigma/r)^la], C as in cpp/MieKinGas.cpp:46-56."""
    Cm = (lr / (lr - la)) * (lr / la) ** (la / (lr - la))
    return power_sum_desc(sigma, eps, [Cm, -Cm], [lr, la])


KIND_LJ_SPLINE, KIND_LJTS, KIND_TANG_TOENNIES, KIND_HFD_B2, KIND_PATKOWSKI = 1, 2, 3, 4, 5
HBAR = 1.054571817e-34      # cpp/global_params.h


def generic_desc(kind, sigma, eps, params=(), fh_order=0, fh_D=0.0):
    """Block of a non-power-law potential (include/kgcuda.h: kg_potential_desc_ex)."""
    d = PotentialDescEx()
    d.kind, d.fh_order, d.sigma, d.eps, d.fh_D, d.nparams = kind, fh_order, sigma, eps, fh_D, len(params)
    for k, v in enumerate(params):
        d.params[k] = v
    return d


def tang_toennies_params(p):
    """Parameter list of KG_POT_TANG_TOENNIES from a fluid-file block (TangToennisParam(json), cpp/multiparam.h:27-41)."""
    return [p["A_div_k"], p["b"], p["A_tilde_div_k"], p["a_tilde"], p["a1"], p["a2"], p["am1"], p["am2"], p["L_unit"],
            p["short_range_lim"]] + list(p["C"])


def hfd_b2_params(p):
    """cpp/multiparam.cpp:91-131"""
    return [p["A"], p["alpha"], p["c6"], p["c8"], p["c10"], p["beta_star"], p["D"], p["rm"]]


def patkowski_params(p):
    """cpp/multiparam.cpp:184-226"""
    return [p["Rc"], p["Ac"], p["Bc"], p["Cex1"], p["Cex2"], p["Csp1"], p["Csp2"], p["Csp3"], p["Csp4"], p["delta"],
            p["C6"], p["C8"], p["C10"]]


def describe_generic(desc, T=300.0):
    """Host-only: the effective piecewise P(r) exp(Q(r)) form the device evaluates, as a dict (kg_genpot_describe)."""
    L = _lib.lib()
    n = -L.kg_genpot_describe(C.byref(desc), float(T), None, 0)
    buf = np.zeros(n)
    if L.kg_genpot_describe(C.byref(desc), float(T), _p(buf), n) != n:
        raise RuntimeError("kg_genpot_describe failed")
    MAXC, MAXQ, MAXT, MAXP = 40, 6, 4, 3
    ints = buf.view(np.int32)
    out = {"len_scale": buf[0], "e_scale": buf[1], "npieces": int(ints[4]), "zero_on_negative": int(ints[5]), "pieces": []}
    term_d = 4 + 2 * MAXC + MAXQ
    piece_d = 2 + MAXT * term_d
    for p in range(out["npieces"]):
        o = 3 + p * piece_d
        piece = {"r_hi": buf[o], "terms": []}
        for t in range(int(ints[2 * (o + 1)])):
            ot = o + 2 + t * term_d
            pmin, pn, dmin, dn, qmin, qn = [int(v) for v in ints[2 * ot:2 * ot + 6]]
            piece["terms"].append({"pmin": pmin, "pc": buf[ot + 4:ot + 4 + pn].copy(), "dmin": dmin, "dc": buf[ot + 4 + MAXC:ot + 4 + MAXC + dn].copy(),
                                   "qmin": qmin, "qc": buf[ot + 4 + 2 * MAXC:ot + 4 + 2 * MAXC + qn].copy()})
        out["pieces"].append(piece)
    return out


class DeviceModel:
    def __init__(self, descs, masses, is_singlecomp=False, device=0):
        self.L = _lib.lib()
        self.ncomps = len(masses)
        npairs = self.ncomps * (self.ncomps + 1) // 2
        assert len(descs) == npairs
        m = np.ascontiguousarray(masses, float)
        h = C.c_void_p()
        if isinstance(descs[0], PotentialDescEx):
            arr = (PotentialDescEx * npairs)(*descs)
            rc = self.L.kg_model_create_ex(arr, self.ncomps, _p(m), int(is_singlecomp), device, C.byref(h))
        else:
            arr = (PotentialDesc * npairs)(*descs)
            rc = self.L.kg_model_create(arr, self.ncomps, _p(m), int(is_singlecomp), device, C.byref(h))
        if rc != 0:
            raise RuntimeError("kg_model_create failed: " + self.L.kg_last_error(None).decode())
        self.h = h
        self.npairs = npairs
        self.descs = descs

    def close(self):
        if getattr(self, "h", None):
            self.L.kg_model_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _chk(self, rc, what):
        if rc != 0:
            raise RuntimeError(f"{what} failed ({rc}): " + self.L.kg_last_error(self.h).decode())

    def set_stream(self, cuda_stream_ptr):
        self._chk(self.L.kg_model_set_stream(self.h, C.c_void_p(int(cuda_stream_ptr))), "kg_model_set_stream")

    def set_temperature_dependent(self, flag=True):
        """Declare that the power-law coefficients of this model depend on T (Feynman-Hibbs Mie): calls without coeff_T are refused."""
        self._chk(self.L.kg_model_set_temperature_dependent(self.h, int(bool(flag))), "kg_model_set_temperature_dependent")

    def pair_index(self, i, j):
        i, j = min(i, j), max(i, j)
        return i * self.ncomps - i * (i - 1) // 2 + (j - i)

    def chi_points(self, pair, T, g, b_red, coeff=None, want_R=False):
        g = np.ascontiguousarray(np.atleast_1d(g), float)
        b = np.ascontiguousarray(np.atleast_1d(b_red), float)
        chi, R = np.empty_like(g), np.empty_like(g)
        st = np.zeros(g.size, np.int32)
        cf = None if coeff is None else np.ascontiguousarray(coeff, float)
        self._chk(self.L.kg_chi_points(self.h, pair, float(T), g.size, _p(g), _p(b), _p(cf), _p(chi), _p(R), _pi(st)), "kg_chi_points")
        raise_on_status(st)
        return (chi, R) if want_R else chi

    def quantum_phase_shifts(self, pair, l, E_red, which=0):
        """Partial-wave phase shifts (Quantum::phase_shift, cpp/Quantum.cpp:522-528) at angular momenta l and reduced energies E / eps;
        which = 1 forces the Numerov solution, 2 the JKWB approximation."""
        l = np.ascontiguousarray(np.atleast_1d(l), np.int32)
        E = np.ascontiguousarray(np.broadcast_to(np.asarray(E_red, float), l.shape))
        out, st = np.empty(l.size), np.zeros(l.size, np.int32)
        self._chk(self.L.kg_quantum_phase_shifts(self.h, pair, l.size, _pi(l), _p(E), which, _p(out), _pi(st)), "kg_quantum_phase_shifts")
        raise_on_status(st)
        return out

    def quantum_cross_sections(self, pair, n, E_red, w_sym=0.5, w_anti=0.5):
        """Q^(n)(E) [m^2] (Quantum::cross_section, cpp/Quantum.cpp:1200-1236): [nE][nn]."""
        n = np.ascontiguousarray(np.atleast_1d(n), np.int32)
        E = np.ascontiguousarray(np.atleast_1d(E_red), float)
        out, st = np.empty((E.size, n.size)), np.zeros(E.size, np.int32)
        self._chk(self.L.kg_quantum_cross_sections(self.h, pair, w_sym, w_anti, n.size, _pi(n), E.size, _p(E), _p(out), _pi(st)), "kg_quantum_cross_sections")
        raise_on_status(st)
        return out

    def quantum_omega(self, pair, T, l, r, w_sym=0.5, w_anti=0.5):
        """Omega^(l,r)(T) from partial-wave phase shifts (Quantum::quantum_omega, cpp/Quantum.cpp:1242-1259): [nT][nlr]."""
        T = np.ascontiguousarray(np.atleast_1d(T), float)
        l, r = np.ascontiguousarray(l, np.int32), np.ascontiguousarray(r, np.int32)
        out, st = np.empty((T.size, l.size)), np.zeros(T.size, np.int32)
        self._chk(self.L.kg_quantum_omega(self.h, pair, w_sym, w_anti, T.size, _p(T), l.size, _pi(l), _pi(r), _p(out), _pi(st)), "kg_quantum_omega")
        raise_on_status(st)
        return out

    def potential_points(self, pair, T, r):
        """u(r) [J] and du/dr [J/m] (Spherical::potential / potential_derivative_r) at temperature T."""
        r = np.ascontiguousarray(np.atleast_1d(r), float)
        u, du = np.empty_like(r), np.empty_like(r)
        self._chk(self.L.kg_potential_points(self.h, pair, float(T), r.size, _p(r), _p(u), _p(du)), "kg_potential_points")
        return u, du

    def omega_batch(self, T, l, r, coeff_T=None, want_w=False):
        T = np.ascontiguousarray(np.atleast_1d(T), float)
        l = np.ascontiguousarray(l, np.int32)
        r = np.ascontiguousarray(r, np.int32)
        om = np.empty((T.size, self.npairs, l.size))
        w = np.empty_like(om)
        st = np.zeros((T.size, self.npairs), np.int32)
        cf = None if coeff_T is None else np.ascontiguousarray(coeff_T, float)
        self._chk(self.L.kg_omega_batch(self.h, T.size, _p(T), l.size, _pi(l), _pi(r), _p(cf), _p(om), _p(w), _pi(st)), "kg_omega_batch")
        raise_on_status(st)
        return (om, w) if want_w else om

    def transfer_lengths(self, T, coeff_T=None):
        """(mtl, etl) in metres, each [nT][npairs] (EWCA model, cpp/transfer_lengths.cpp:165-179)."""
        T = np.ascontiguousarray(np.atleast_1d(T), float)
        mtl, etl = np.empty((T.size, self.npairs)), np.empty((T.size, self.npairs))
        st = np.zeros((T.size, self.npairs), np.int32)
        cf = None if coeff_T is None else np.ascontiguousarray(coeff_T, float)
        self._chk(self.L.kg_transfer_lengths_batch(self.h, T.size, _p(T), _p(cf), _p(mtl), _p(etl), _pi(st)), "kg_transfer_lengths_batch")
        raise_on_status(st)
        return mtl, etl

    def cross_sections(self, pair, E_red, l, coeff=None):
        """Q^(l)(E) [m^2], E in units of the pair's eps (Spherical::cross_section, cpp/Spherical.cpp:104-127): [nE][nl]."""
        E = np.ascontiguousarray(np.atleast_1d(E_red), float)
        l = np.ascontiguousarray(np.atleast_1d(l), np.int32)
        out = np.empty((E.size, l.size))
        st = np.zeros(E.size, np.int32)
        cf = None if coeff is None else np.ascontiguousarray(coeff, float)
        self._chk(self.L.kg_cross_sections(self.h, pair, E.size, _p(E), l.size, _pi(l), _p(cf), _p(out), _pi(st)), "kg_cross_sections")
        raise_on_status(st & ~64)
        return out

    def second_virial(self, pair, T, coeff_T=None):
        """Classical second virial coefficient [m^3/mol] (Spherical::second_virial, cpp/Spherical.cpp:140-160)."""
        T = np.ascontiguousarray(np.atleast_1d(T), float)
        out = np.empty(T.size)
        cf = None if coeff_T is None else np.ascontiguousarray(coeff_T, float)
        self._chk(self.L.kg_second_virial(self.h, pair, T.size, _p(T), _p(cf), _p(out)), "kg_second_virial")
        return out

    def last_timings(self):
        out = np.zeros(8)
        self.L.kg_last_timings(self.h, _p(out))
        return dict(zip(["ms_state_kernels", "ms_omega_stage", "ms_tl_stage", "ms_bracket_stage", "launches_tcache", "launches_batch", "ms_batch_call"], out.tolist()))

    def omega_plan(self, T, l, r, coeff_T=None):
        return OmegaPlan(self, T, l, r, coeff_T)


class OmegaTable:
    """Device-resident Omega(T) interpolation table (kg_omega_table): built once by exact quadrature at nodes uniform in ln T, segmented at
    the lattice-regime temperatures, evaluated on the device, persistent through save() / load(). Accuracy: see include/kgcuda.h."""

    def __init__(self, model, handle):
        self.model, self.h = model, handle
        n = model.L.kg_omega_table_lr(handle, None, None)
        self.l, self.r = np.zeros(n, np.int32), np.zeros(n, np.int32)
        model.L.kg_omega_table_lr(handle, _pi(self.l), _pi(self.r))

    @classmethod
    def build(cls, model, Tmin, Tmax, l, r, nodes_per_decade=64):
        l = np.ascontiguousarray(l, np.int32)
        r = np.ascontiguousarray(r, np.int32)
        h = C.c_void_p()
        model._chk(model.L.kg_omega_table_create(model.h, float(Tmin), float(Tmax), int(nodes_per_decade), l.size, _pi(l), _pi(r), C.byref(h)),
                   "kg_omega_table_create")
        return cls(model, h)

    @classmethod
    def load(cls, model, path):
        h = C.c_void_p()
        model._chk(model.L.kg_omega_table_load(model.h, str(path).encode(), C.byref(h)), "kg_omega_table_load")
        return cls(model, h)

    def save(self, path):
        self.model._chk(self.model.L.kg_omega_table_save(self.h, str(path).encode()), "kg_omega_table_save")

    def __call__(self, T):
        T = np.ascontiguousarray(np.atleast_1d(T), float)
        out = np.empty((T.size, self.model.npairs, self.l.size))
        self.model._chk(self.model.L.kg_omega_table_eval(self.h, T.size, _p(T), _p(out)), "kg_omega_table_eval")
        return out

    def info(self):
        out = np.zeros(8)
        self.model.L.kg_omega_table_info(self.h, _p(out))
        return dict(zip(["Tmin", "Tmax", "nodes", "segments", "npairs", "nlr", "bytes"], out.tolist()))

    def close(self):
        if getattr(self, "h", None):
            self.model.L.kg_omega_table_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def raise_on_status(st):
    bad = int(np.bitwise_or.reduce(np.asarray(st).ravel())) if np.size(st) else 0
    if bad:
        msgs = [t for b, t in STATUS_TEXT.items() if bad & b]
        raise RuntimeError("; ".join(msgs))     # same text as the reference's throw sites (cpp/Spherical.cpp:432)


class OmegaPlan:
    def __init__(self, model, T, l, r, coeff_T=None):
        self.model = model
        self.T = np.ascontiguousarray(np.atleast_1d(T), float)
        self.l = np.ascontiguousarray(l, np.int32)
        self.r = np.ascontiguousarray(r, np.int32)
        self.cf = None if coeff_T is None else np.ascontiguousarray(coeff_T, float)
        h = C.c_void_p()
        model._chk(model.L.kg_omega_plan_create(model.h, self.T.size, _p(self.T), self.l.size, _pi(self.l), _pi(self.r), _p(self.cf),
                                                C.byref(h)), "kg_omega_plan_create")
        self.h = h

    def run(self, sync=True):
        self.model._chk(self.model.L.kg_omega_plan_run(self.h), "kg_omega_plan_run")
        if sync:
            self.model._chk(self.model.L.kg_omega_plan_sync(self.h), "kg_omega_plan_sync")

    def fetch(self, want_w=False):
        shape = (self.T.size, self.model.npairs, self.l.size)
        om, w = np.empty(shape), np.empty(shape)
        st = np.zeros(shape[:2], np.int32)
        self.model._chk(self.model.L.kg_omega_plan_fetch(self.h, _p(om), _p(w), _pi(st)), "kg_omega_plan_fetch")
        raise_on_status(st)
        return (om, w) if want_w else om

    def export_dev(self, omega_ptr, w_ptr=None):
        """D2D copy of the result tables into caller-owned device buffers (e.g. torch tensors' data_ptr())."""
        self.model._chk(self.model.L.kg_omega_plan_export_dev(self.h, C.c_void_p(int(omega_ptr)), C.c_void_p(int(w_ptr)) if w_ptr else None),
                        "kg_omega_plan_export_dev")

    def counters(self):
        out = np.zeros(16)
        self.model._chk(self.model.L.kg_omega_plan_counters(self.h, _p(out)), "kg_omega_plan_counters")
        keys = ["chi_evals", "chi_nodes", "lim_evals", "lim_nodes", "ms_theta", "ms_lim", "ms_mesh", "launches", "passes",
                "h2d_bytes", "d2h_bytes", "pool_bytes", "max_cells_per_level", "cell_capacity", "cell_capacity_limit", "cell_regrows"]
        return dict(zip(keys, out.tolist()))

    def quadrature_stats(self):
        out = np.zeros(8)
        self.model._chk(self.model.L.kg_omega_plan_quadrature_stats(self.h, _p(out)), "kg_omega_plan_quadrature_stats")
        return dict(zip(["reference_nodes", "executed_nodes", "redo_integrals", "theta_integrals", "stride8_integrals", "mesh_cells"], out.tolist()))

    def close(self):
        if getattr(self, "h", None):
            self.model.L.kg_omega_plan_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def set_exact_quadrature(exact):
    """Validation switch (kg_set_exact_quadrature): node-by-node tanh-sinh sums instead of the strided ones."""
    if _lib.lib().kg_set_exact_quadrature(int(bool(exact))) != 0:
        raise RuntimeError("kg_set_exact_quadrature failed (no CUDA device?)")


def measure_fp64_peak(device=0):
    t, f = C.c_double(), C.c_double()
    rc = _lib.lib().kg_measure_fp64_peak(device, C.byref(t), C.byref(f))
    if rc != 0:
        raise RuntimeError("kg_measure_fp64_peak failed (no CUDA device?)")
    return t.value, f.value
