"""SAFT-VR Mie equation of state on the GPU: the EoS-derived inputs of the RET property code for batches of state points.

Host-side mirror of the part of thermopack's `saftvrmie` object that pykingas uses (pykingas/MieKinGas.py:55-65 constructs it;
cpp/eos_interface.h:121-181 calls chemical_potential_tv(dmudn), idealenthalpysingle(dhdt), specific_volume(dvdn), pressure_tv) --
same method names and return conventions for scalar calls, plus batched calls that take arrays. The arithmetic runs in
kineticgas_b200/csrc/kg_eos.cu behind the C-ABI (include/kgcuda.h kg_eos_*); there is no CPU fallback.

Parameters: `DeviceSaftVRMie(comps)` loads what `thermopack.saftvrmie(comps)` would load -- segment numbers, Mie parameters, binary
k_ij / l_ij / gamma_ij, molar masses, ideal-gas Cp correlations -- from kineticgas_b200/saft_db.json, which the fixture generator
wrote from the thermopack 2.2.3 wheel vendored in the reference tree; `set_pure_fluid_param` overrides them like thermopack's.
"""
import ctypes as C
import json
import os

import numpy as np

from . import _lib
from ._lib import dp, ip

_DB = None


def _db():
    global _DB
    if _DB is None:
        _DB = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "saft_db.json")))
    return _DB


def _ptr(a):
    return None if a is None else a.ctypes.data_as(dp)


class DevicePtr:
    """A device-resident double array handed from one C-ABI call to the next (address + logical shape)."""

    def __init__(self, address, shape):
        self.address, self.shape = int(address), tuple(shape)


class DeviceSaftVRMie:
    VAPPH, LIQPH = 2, 1                     # thermopack's phase flags (only the vapour-like root is implemented)
    Rgas = 6.02214076e23 * 1.380649e-23

    def __init__(self, comps, device=0):
        self.comps = comps.split(",")
        self.nc = len(self.comps)
        db = _db()
        for c in self.comps:
            if c not in db["pure"]:
                raise KeyError(f"no SAFT-VR Mie parameters for {c} in saft_db.json")
        pure = [db["pure"][c] for c in self.comps]
        self.m = np.array([p["m"] for p in pure], float)
        self._sigma = np.array([p["sigma"] for p in pure], float)
        self._eps = np.array([p["eps_div_k"] for p in pure], float)
        self._la = np.array([p["la"] for p in pure], float)
        self._lr = np.array([p["lr"] for p in pure], float)
        self.mw = np.array([p["mw"] for p in pure], float)
        self.cp_type = np.array([p["cp_type"] for p in pure], np.int32)
        self.cp_par = np.array([p["cp_par"] for p in pure], float)
        nc = self.nc
        self.kij, self.lij, self.gammaij = np.zeros((nc, nc)), np.zeros((nc, nc)), np.zeros((nc, nc))
        for i in range(nc):
            for j in range(nc):
                b = db["binary"].get(self.comps[i] + "," + self.comps[j]) or db["binary"].get(self.comps[j] + "," + self.comps[i])
                if b and i != j:
                    self.kij[i, j], self.lij[i, j], self.gammaij[i, j] = b["kij"], b["lij"], b["gammaij"]
        self._device = device
        self.h = None
        self._build()

    # ------------------------------------------------------------------------------------------------ parameters
    def pair_matrices(self):
        """thermopack's combining rules (Lafitte et al. 2013; src/saftvrmie_containers.f90): sigma_ij arithmetic with l_ij, epsilon_ij
        geometric with the sigma^3 weighting and k_ij, lambda_ij - 3 geometric with gamma_ij."""
        s, e, la, lr = self._sigma, self._eps, self._la, self._lr
        sig = 0.5 * (s[:, None] + s[None, :]) * (1 - self.lij)
        eps = (1 - self.kij) * np.sqrt(s[:, None] ** 3 * s[None, :] ** 3) / sig ** 3 * np.sqrt(e[:, None] * e[None, :])
        laij = 3 + np.sqrt((la[:, None] - 3) * (la[None, :] - 3))
        lrij = 3 + (1 - self.gammaij) * np.sqrt((lr[:, None] - 3) * (lr[None, :] - 3))
        return sig, eps, laij, lrij

    def get_pure_fluid_param(self, ic):
        i = ic - 1
        return self.m[i], self._sigma[i], self._eps[i], self._la[i], self._lr[i]

    def set_pure_fluid_param(self, ic, m, sigma, eps_div_kb, lambda_a, lambda_r):
        i = ic - 1
        self.m[i], self._sigma[i], self._eps[i], self._la[i], self._lr[i] = m, sigma, eps_div_kb, lambda_a, lambda_r
        self._build()

    def _build(self):
        L = _lib.lib()
        self.L = L
        if self.h:
            L.kg_eos_destroy(self.h)
            self.h = None
        sig, eps, la, lr = (np.ascontiguousarray(a, float) for a in self.pair_matrices())
        h = C.c_void_p()
        rc = L.kg_eos_saftvrmie_create(self._device, self.nc, _ptr(np.ascontiguousarray(self.m)), _ptr(sig), _ptr(eps), _ptr(la), _ptr(lr),
                                       self.cp_type.ctypes.data_as(ip), _ptr(np.ascontiguousarray(self.cp_par)), _ptr(np.ascontiguousarray(self.mw)),
                                       C.byref(h))
        if rc:
            raise RuntimeError("kg_eos_saftvrmie_create: " + L.kg_eos_last_error(None).decode())
        self.h = h

    def close(self):
        if getattr(self, "h", None):
            self.L.kg_eos_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _chk(self, rc, what):
        if rc:
            raise RuntimeError(what + ": " + self.L.kg_eos_last_error(self.h).decode())

    # ------------------------------------------------------------------------------------------------ batched calls
    def tv_batch(self, T, Vm, x, dmudn=True, cp_ideal=False, pressure=False, dvdn=False):
        """dict of arrays for n state points (T [n], Vm [n] m^3/mol, x [n][nc]): 'dmudn' [n][nc][nc] J/mol^2, 'cp_ideal' [n][nc] J/(mol K),
        'pressure' [n] Pa, 'dvdn' [n][nc] m^3/mol -- the layout kineticgas_b200's `eos_inputs=` takes."""
        T = np.ascontiguousarray(np.atleast_1d(T), float)
        n, nc = T.size, self.nc
        Vm = np.ascontiguousarray(np.broadcast_to(np.asarray(Vm, float), (n,)))
        x = np.ascontiguousarray(np.broadcast_to(np.asarray(x, float), (n, nc)))
        out = {}
        if dmudn:
            out["dmudn"] = _lib.pinned_empty((n, nc, nc)) if n >= 32768 else np.empty((n, nc, nc))
        if cp_ideal:
            out["cp_ideal"] = np.empty((n, nc))
        if pressure:
            out["pressure"] = np.empty(n)
        if dvdn:
            out["dvdn"] = np.empty((n, nc))
        self._chk(self.L.kg_eos_tv_batch(self.h, n, _ptr(T), _ptr(Vm), _ptr(x), _ptr(out.get("dmudn")), _ptr(out.get("cp_ideal")),
                                         _ptr(out.get("pressure")), _ptr(out.get("dvdn"))), "kg_eos_tv_batch")
        return out

    def tv_batch_dev(self, T, Vm, x, dmudn=True, cp_ideal=False, dvdn=False):
        """As tv_batch, results left on the GPU: dict of DevicePtr (valid until the next call on this object), accepted by
        MieKinGas.transport_batch(eos_inputs=...) -- the 5x5 matrices of 10^5 points never cross the PCIe bus."""
        T = np.ascontiguousarray(np.atleast_1d(T), float)
        n, nc = T.size, self.nc
        Vm = np.ascontiguousarray(np.broadcast_to(np.asarray(Vm, float), (n,)))
        x = np.ascontiguousarray(np.broadcast_to(np.asarray(x, float), (n, nc)))
        dev = (C.c_void_p * 4)()
        want = (1 if dmudn else 0) | (2 if cp_ideal else 0) | (8 if dvdn else 0)
        self._chk(self.L.kg_eos_tv_batch_dev(self.h, n, _ptr(T), _ptr(Vm), _ptr(x), want, dev), "kg_eos_tv_batch_dev")
        shapes = {"dmudn": (n, nc, nc), "cp_ideal": (n, nc), "dvdn": (n, nc)}
        return {k: DevicePtr(dev[i], shapes[k]) for i, k in ((0, "dmudn"), (1, "cp_ideal"), (3, "dvdn")) if dev[i]}

    def specific_volume_batch(self, T, p, x):
        """Vapour-like molar volume [m^3/mol] of n state points (T, p, x)."""
        T = np.ascontiguousarray(np.atleast_1d(T), float)
        n, nc = T.size, self.nc
        p = np.ascontiguousarray(np.broadcast_to(np.asarray(p, float), (n,)))
        x = np.ascontiguousarray(np.broadcast_to(np.asarray(x, float), (n, nc)))
        Vm, st = np.empty(n), np.zeros(n, np.int32)
        self._chk(self.L.kg_eos_specific_volume_batch(self.h, n, _ptr(T), _ptr(p), _ptr(x), _ptr(Vm), st.ctypes.data_as(ip)), "kg_eos_specific_volume_batch")
        if st.any():
            raise RuntimeError("specific_volume: the density solver did not converge for %d state points" % int(st.sum()))
        return Vm

    def last_kernel_ms(self):
        return float(self.L.kg_eos_last_kernel_ms(self.h))

    # ------------------------------------------------------------------------------------------------ thermopack-style scalar calls
    def chemical_potential_tv(self, temp, volume, n, dmudt=None, dmudv=None, dmudn=None, property_flag="IR"):
        """Only what cpp/eos_interface.h:143-151 uses: returns (None, dmudn) with dmudn [nc][nc]; n must be one mole in total."""
        if dmudt is not None or dmudv is not None or property_flag != "IR":
            raise NotImplementedError("DeviceSaftVRMie.chemical_potential_tv: only dmudn of the total chemical potential")
        return None, self.tv_batch([temp], [volume], [list(n)])["dmudn"][0]

    def idealenthalpysingle(self, temp, j, dhdt=None):
        """(None, Cp_ideal of component j) -- cpp/eos_interface.h:153-160 reads element [1]."""
        return None, float(self.tv_batch([temp], [1.0], [[1.0 / self.nc] * self.nc], dmudn=False, cp_ideal=True)["cp_ideal"][0, j - 1])

    def pressure_tv(self, temp, volume, n, dpdt=None, dpdv=None, dpdn=None, property_flag="IR"):
        return (float(self.tv_batch([temp], [volume], [list(n)], dmudn=False, pressure=True)["pressure"][0]),)

    def specific_volume(self, temp, press, x, phase, dvdt=None, dvdp=None, dvdn=None):
        if phase != self.VAPPH:
            raise NotImplementedError("DeviceSaftVRMie.specific_volume: vapour-like root only (pykingas calls it with VAPPH)")
        v = float(self.specific_volume_batch([temp], [press], [list(x)])[0])
        if dvdn is None:
            return (v,)
        return v, self.tv_batch([temp], [v], [list(x)], dmudn=False, dvdn=True)["dvdn"][0]
