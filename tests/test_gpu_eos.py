"""SURVEY 8f row 1 on the GPU: kineticgas_b200.DeviceSaftVRMie (C-ABI kg_eos_*, kernels in csrc/kg_eos.cu) against thermopack 2.2.3 --
the EoS object pykingas constructs and the reference's C++ calls through cpp/eos_interface.h:121-181 -- via tests/golden/eos_reference.npz
(oracle/make_golden.py eos). Tolerance 1e-9 relative: the device model evaluates one generic Helmholtz function in hyper-dual numbers, so
derivatives are exact to rounding (measured agreement 1e-12); the density solver stops at a relative step of 1e-13."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
G = dict(np.load(os.path.join(os.path.dirname(__file__), "golden", "eos_reference.npz")))
TAGS = ["ar", "co2c1n2", "c1nc10", "five"]
TOL = 1e-9


@pytest.mark.parametrize("tag", TAGS)
def test_database_parameters_and_combining_rules(tag):
    from kineticgas_b200 import DeviceSaftVRMie
    eos = DeviceSaftVRMie(str(G[tag + "_comps"]))
    sig, eps, la, lr = eos.pair_matrices()
    assert np.array_equal(eos.m, G[tag + "_m"])
    assert np.max(np.abs(sig / G[tag + "_sigma"] - 1)) < 1e-14 and np.max(np.abs(eps / G[tag + "_eps"] - 1)) < 1e-13


@pytest.mark.parametrize("tag", TAGS)
def test_tv_properties_match_thermopack(tag):
    from kineticgas_b200 import DeviceSaftVRMie
    eos = DeviceSaftVRMie(str(G[tag + "_comps"]))
    r = eos.tv_batch(G[tag + "_T"], G[tag + "_Vm"], G[tag + "_x"], dmudn=True, cp_ideal=True, pressure=True)
    ref = G[tag + "_dmudn"]
    e_dm = np.max(np.abs(r["dmudn"] - ref), axis=(1, 2)) / np.max(np.abs(ref), axis=(1, 2))
    e_cp = np.abs(r["cp_ideal"] / G[tag + "_cp"] - 1)
    e_p = np.abs(r["pressure"] / G[tag + "_p"] - 1)
    print(tag, "dmudn", e_dm.max(), "cp", e_cp.max(), "p", e_p.max())
    assert e_dm.max() < TOL and e_cp.max() < TOL and e_p.max() < TOL
    # dmudn alone takes the smaller Hessian (no volume row): same numbers
    r2 = eos.tv_batch(G[tag + "_T"], G[tag + "_Vm"], G[tag + "_x"])
    assert np.max(np.abs(r2["dmudn"] - r["dmudn"]) / np.abs(r["dmudn"]).max()) < 1e-13


@pytest.mark.parametrize("tag", TAGS)
def test_specific_volume_and_dvdn_match_thermopack(tag):
    from kineticgas_b200 import DeviceSaftVRMie
    eos = DeviceSaftVRMie(str(G[tag + "_comps"]))
    Vm = eos.specific_volume_batch(G[tag + "_Tp"], G[tag + "_pp"], G[tag + "_x"])
    e_v = np.abs(Vm / G[tag + "_vm_tp"] - 1)
    dv = eos.tv_batch(G[tag + "_Tp"], Vm, G[tag + "_x"], dmudn=False, dvdn=True)["dvdn"]
    e_d = np.max(np.abs(dv - G[tag + "_dvdn_tp"]), axis=1) / np.max(np.abs(G[tag + "_dvdn_tp"]), axis=1)
    print(tag, "Vm", e_v.max(), "dvdn", e_d.max())
    assert e_v.max() < TOL and e_d.max() < 1e-8
    # the thermopack-style scalar calls the reference's eos_interface makes
    v, dvdn = eos.specific_volume(G[tag + "_Tp"][0], G[tag + "_pp"][0], list(G[tag + "_x"][0]), eos.VAPPH, dvdn=True)
    assert abs(v / G[tag + "_vm_tp"][0] - 1) < TOL
    _, dm = eos.chemical_potential_tv(G[tag + "_T"][0], G[tag + "_Vm"][0], list(G[tag + "_x"][0]), dmudn=True)
    assert np.max(np.abs(dm - G[tag + "_dmudn"][0])) / np.max(np.abs(G[tag + "_dmudn"][0])) < TOL
    assert abs(eos.idealenthalpysingle(G[tag + "_T"][0], 1, dhdt=True)[1] / G[tag + "_cp"][0, 0] - 1) < TOL


def test_transport_with_device_eos_equals_transport_with_thermopack_inputs():
    """RET-Mie diffusion + conductivity of CO2/C1/N2 with use_eos='device' (EoS inputs computed on the GPU, fed to kg_transport_batch)
    against the same call with thermopack's dmudn / Cp_ideal passed as arrays."""
    from kineticgas_b200 import MieKinGas
    tag = "co2c1n2"
    T, Vm, x = G[tag + "_T"], G[tag + "_Vm"], G[tag + "_x"]
    keep = Vm > 2e-4                                           # gas-like states (the RET model is a gas-phase theory)
    T, Vm, x = T[keep], Vm[keep], x[keep]
    a = MieKinGas("CO2,C1,N2", N=2, use_eos="device")
    b = MieKinGas("CO2,C1,N2", N=2)
    ra = a.transport_batch(T, Vm, x, N=2)
    rb = b.transport_batch(T, Vm, x, N=2, eos_inputs={"dmudn": G[tag + "_dmudn"][keep], "cp_ideal": G[tag + "_cp"][keep]})
    for k in ("viscosity", "conductivity", "D", "DT"):
        ok = np.isfinite(rb[k])
        err = np.max(np.abs(ra[k][ok] - rb[k][ok]) / np.max(np.abs(rb[k][ok])))
        print(k, err)
        assert err < 1e-8
    # *_tp calls: Vm(T, p) from the device EoS
    eta_tp = a.viscosity_tp(G[tag + "_Tp"][:4], G[tag + "_pp"][:4], G[tag + "_x"][:4], N=2)
    eta_tv = b.viscosity(G[tag + "_Tp"][:4], G[tag + "_vm_tp"][:4], G[tag + "_x"][:4], N=2)
    assert np.max(np.abs(eta_tp / eta_tv - 1)) < 1e-8


def test_device_resident_hand_over_for_large_batches():
    """n >= 4096: the EoS results stay on the GPU (kg_eos_tv_batch_dev) and kg_transport_batch reads them there (its EoS arrays may be
    device pointers); same numbers as the host-array route."""
    from kineticgas_b200 import MieKinGas, DeviceSaftVRMie
    tag = "co2c1n2"
    keep = G[tag + "_Vm"] > 2e-4
    rep = 6000 // int(keep.sum()) + 1
    T, Vm, x = (np.tile(G[tag + k][keep], (rep,) + (1,) * (G[tag + k].ndim - 1)) for k in ("_T", "_Vm", "_x"))
    order = np.argsort(T, kind="stable")                        # runs of equal temperature, as sweeps have them
    T, Vm, x = T[order], Vm[order], np.ascontiguousarray(x[order])
    a = MieKinGas("CO2,C1,N2", N=2, use_eos="device")
    ra = a.transport_batch(T, Vm, x, N=2)                       # device hand-over
    eos = DeviceSaftVRMie("CO2,C1,N2")
    host = eos.tv_batch(T, Vm, x, dmudn=True, cp_ideal=True)
    rb = MieKinGas("CO2,C1,N2", N=2).transport_batch(T, Vm, x, N=2, eos_inputs=host)
    for k in ("viscosity", "conductivity", "D", "DT"):
        ok = np.isfinite(rb[k])
        # not bit-equal: two model objects integrate Omega separately and the leaf sums are accumulated with atomics (order-dependent last bits)
        assert np.max(np.abs(ra[k][ok] - rb[k][ok])) / np.max(np.abs(rb[k][ok])) < 1e-12, k
    dev = eos.tv_batch_dev(T, Vm, x, dmudn=True, cp_ideal=True)
    assert set(dev) == {"dmudn", "cp_ideal"} and dev["dmudn"].shape == (T.size, 3, 3)
