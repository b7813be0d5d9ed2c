#!/usr/bin/env python
"""bench.py -- headline benchmark of the collision-integral hot path (BASELINE.json configs[1]):
Lennard-Jones (fluid LJF, Mie 12-6) collision integrals Omega^(l,r), 1 <= l <= r <= 4 (10 per temperature), on a
4096-point T* = 0.3..100 grid (T = 30 T* K: all five lattice regimes of cpp/Spherical.cpp:60-84).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]

A step = one pass of the hot path over the whole grid. N>1 (torchrun): weak scaling -- rank k owns every N-th point of
an N-times denser grid (per-GPU work fixed); the only collective is one all_gather of the Omega tables per step.
Prints ONE JSON line on rank 0.
"""
import argparse
import json
import multiprocessing as mp
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

NT_PER_GPU = 4096
LR = [(l, r) for l in range(1, 5) for r in range(l, 5)]
# FP64-pipe lane instructions (DFMA + DMUL + DADD, predicated-on threads) per unit of work, calibrated on ncu's
# smsp__sass_thread_inst_executed_op_{dfma,dmul,dadd}_pred_on summed over ALL launches of one full-size headline step of this build
# (tools/gpu_calib.sh; profiles/r2_summary.md) and divided by the step's device counters:
NCU_THETA_TRAFFIC_BYTES = 757.5e6 + 261.1e6 + 153.7e6 + 412.3e6     # ncu DRAM read + write of the largest theta + set-up launch pair (see roofline.traffic_note)
FP64_PER_NODE = 8.99         # theta_kernel<2>: 5.402e10 lane instructions / 6.012e9 executed node evaluations (all-in: coarse nodes 8 each,
                             # end nodes, difference formulas, continuation to the stop, node-by-node tier)
FP64_PER_SETUP = 228.0       # theta_setup_kernel<2>: 7.43e9 lane instructions / 32.63 M root findings (integer-exponent get_R in registers)
FP64_PER_CELL = 406.0        # mesh_level_kernel: 8.42e9 lane instructions / 20.73 M cell records of all levels (a group record holds up to 4 leaves;
                             # ten (l,r) members each)
F_NODE = 98.0    # nominal FP64 flop per tanh-sinh node for a 2-term (Mie/LJ) potential, SURVEY.md 8(d): 52 + 23 n_terms
METRIC = "collision integrals/s"


def hbm_peak():
    """(GB/s, source): MEASURED_PEAKS.json (driver-written copy bandwidth of this pool's B200s), else the profiling recipe's fallback."""
    try:
        return float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs"
    except Exception:
        return 6550.0, "fallback of B200_PROFILING.md (MEASURED_PEAKS.json absent)"


# DRAM bytes per state point moved by the state kernels, from ncu --set full captures of this build (dram__bytes_read.sum + dram__bytes_write.sum
# of rdf_kernel + state_kernel over the launch's points; profiles/r2_summary.md)
STATE_TRAFFIC_PER_POINT = {"state_points": None, "five_component": 3360.0}


def state_roofline(leg, n_points, h2d, d2h, ms_kernels):
    """The state-point kernels against the HBM roofline: algorithmic bytes = the caller's inputs and outputs of one step (what h2d/d2h count;
    per-temperature tables are shared by ~100-1500 points and stay in L2), over the summed duration of the rdf + state kernels."""
    peak, src = hbm_peak()
    ach = (h2d + d2h) / (ms_kernels * 1e-3) / 1e9 if ms_kernels > 0 else 0.0
    tpp = STATE_TRAFFIC_PER_POINT.get(leg)
    return {"bound": "hbm", "kernel": "rdf_kernel + state_kernel<N nc> (one thread per state point: RDF at contact, Sonine systems assembled in shared "
            "memory, LU in Eigen's pivot order, properties)", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
            "traffic": None if tpp is None else tpp * n_points, "algorithmic_bytes_per_step": int(h2d + d2h), "kernel_ms_per_step": ms_kernels,
            "peak_source": src,
            "note": "far below the HBM roofline by construction: a state point is ~2e4 FP64 operations (matrix assembly + two dense LU solves of "
                    "order N nc) on ~0.5 kB of input; the kernels are latency / FP64-issue bound, the fraction is reported for completeness"}


def grid(n_gpus, rank):
    tstar = np.geomspace(0.3, 100.0, NT_PER_GPU * n_gpus)[rank::n_gpus]
    return 30.0 * tstar


def ljf():
    from kineticgas_b200.fluids import mie_parameters, BOLTZMANN, AVOGADRO
    ids, Mw, s, e, la, lr = mie_parameters("LJF")
    return Mw[0], s[0], e[0], la[0], lr[0], BOLTZMANN, AVOGADRO


# ------------------------------------------------------------------------------------------------ CPU reference arm
def _cpu_worker(task):
    """One Omega through the reference's own Spherical::omega (oracle/_ref) or the C port of it."""
    T, l, r, kind = task
    Mw, s, e, la, lr, KB, NA = ljf()
    t0 = time.perf_counter()
    if kind == "reference":
        from oracle import ref
        m = ref.RefMie([Mw], [s], [e], [la], [lr], is_idealgas=True, py_constants=False)     # same eps [J] and m [kg] as the device model
        m.set_omega_correlation_active(False)
        val = m.omega(0, 0, l, r, T)
    else:
        from oracle import port
        p = port.mie_pair(s, e * KB, la, lr)
        val = port.w_integral(p, T, l, r) * port.lib().kgo_omega_prefactor(s, T, Mw * 1e-3 / NA, Mw * 1e-3 / NA, 1)
    return T, l, r, val, time.perf_counter() - t0


def cpu_sample(n_gpus, n_sample):
    """Stratified sample of the workload: the grid's index range cut into n_sample equal strata, the middle point of
    each, (l,r) cycling through the 10 combinations -> same regime mix as the full grid."""
    T = np.sort(np.concatenate([grid(n_gpus, k) for k in range(n_gpus)]))
    idx = ((np.arange(n_sample) + 0.5) * T.size / n_sample).astype(int)
    return [(float(T[i]), LR[k % len(LR)][0], LR[k % len(LR)][1]) for k, i in enumerate(idx)]


def run_cpu_reference(n_gpus, n_sample, cores, warm=False):
    """Throughput mode of SURVEY 8(d): one worker process per host core, one Omega per task (longest first). The reported value
    is the LOAD-BALANCED throughput n x cores / (sum of per-task core-seconds): what the box sustains on a long grid, where the
    tail of a short sample (one 20 s Omega at T < 20 K) does not set the wall clock. The raw n / wall of the sample is kept.
    warm=True: a cheap untimed pass (hot temperatures only) that starts the pool and loads the library."""
    from oracle import ref
    kind = "reference" if ref.available() else "port"
    if warm:
        tasks = [(3000.0 - 10.0 * k, 1, 1, kind) for k in range(cores)]
    else:
        tasks = [(T, l, r, kind) for (T, l, r) in cpu_sample(n_gpus, n_sample)]
    tasks.sort(key=lambda t: t[0])          # coldest (most expensive) first
    t0 = time.perf_counter()
    with mp.get_context("spawn").Pool(cores) as pool:
        res = pool.map(_cpu_worker, tasks, chunksize=1)
    wall = time.perf_counter() - t0
    core_s = sum(r[4] for r in res)
    return {"value": len(tasks) * cores / core_s, "unit": METRIC, "cores": cores, "nproc": os.cpu_count(), "kind": kind,
            "value_wall_clock_of_sample": len(tasks) / wall, "core_seconds": core_s, "wall_seconds": wall,
            "longest_task_seconds": max(r[4] for r in res),
            "sample": f"{len(tasks)} Omega: one T per equal-count stratum of the sorted T grid, (l,r) cycling over the 10 combinations, one "
                      f"process per core, longest first; value = n x cores / sum(core-s) = load-balanced throughput "
                      f"({core_s:.1f} core-s, {wall:.1f} s wall of which the slowest single Omega is {max(r[4] for r in res):.1f} s)"}, res


def _as_shipped_worker(Ts):
    """Mode (i) of SURVEY 8(d): ONE process, the reference as shipped -- viscosity + thermal conductivity of LJF (ideal gas, N = 2,
    correlation off) per temperature; the reference fans the uncached Omega of one property call out over <= 8 std::threads
    (cpp/KineticGas_mthr.cpp:13, 58-143). 7 distinct Omega per temperature."""
    from oracle import ref
    Mw, s, e, la, lr, KB, NA = ljf()
    m = ref.RefMie([Mw], [s], [e], [la], [lr], is_idealgas=True)
    m.set_omega_correlation_active(False)
    R = 8.31446261815324
    t0 = time.perf_counter()
    for T in Ts:
        Vm = R * T / 1e5
        m.set_eos(dmudn=np.array([[R * T]]), cp_id=np.array([2.5 * R]), p=1e5, vm=Vm, dvdn=np.array([Vm]))
        m.viscosity(T, Vm, [0.5, 0.5], 2)
        m.thermal_conductivity(T, Vm, [0.5, 0.5], 2)
    return time.perf_counter() - t0


def run_cpu_as_shipped(n_gpus, n_T=4):
    from oracle import ref
    if not ref.available():
        return None
    T = np.sort(np.concatenate([grid(n_gpus, k) for k in range(n_gpus)]))
    Ts = [float(T[int((k + 0.5) * T.size / n_T)]) for k in range(n_T)]
    with mp.get_context("spawn").Pool(1) as pool:
        wall = pool.apply(_as_shipped_worker, (Ts,))
    return {"value": 7 * n_T / wall, "unit": METRIC, "threads": 8, "kind": "reference",
            "sample": f"as shipped: one process, viscosity + thermal_conductivity (N = 2, 7 distinct Omega) at {n_T} stratified temperatures "
                      f"{[round(t, 2) for t in Ts]} K, the reference's own <= 8-thread fan-out inside each property call; {wall:.1f} s wall"}


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


# ------------------------------------------------------------------------------------------------ RET state points (configs[2])
SP_COMPS, SP_N = "C1,NC10", 3
SP_GRID = (64, 64, 16)        # T x rho* x x1
R_GAS = 8.31446261815324


def sp_grid(n_gpus, rank, shape=SP_GRID):
    """BASELINE.json configs[2] / SURVEY 8(d).3: T = linspace(300,700,64) K, rho* = rho sigma_11^3 = linspace(0.01,0.6,64),
    x_1 = linspace(0.02,0.98,16) -> 65 536 state points per GPU (weak scaling: rank k owns every n_gpus-th temperature of an
    n_gpus-times denser T grid)."""
    from kineticgas_b200.fluids import mie_parameters, AVOGADRO
    ids, Mw, s, e, la, lr = mie_parameters(SP_COMPS)
    nT, nr, nx = shape
    T = np.linspace(300.0, 700.0, nT * n_gpus)[rank::n_gpus]
    rho = np.linspace(0.01, 0.6, nr) / s[0] ** 3
    x1 = np.linspace(0.02, 0.98, nx)
    Tg, rg, xg = np.meshgrid(T, rho, x1, indexing="ij")
    Tn, Vm = Tg.ravel(), AVOGADRO / rg.ravel()
    x = np.stack([xg.ravel(), 1.0 - xg.ravel()], axis=1)
    return np.ascontiguousarray(Tn), np.ascontiguousarray(Vm), np.ascontiguousarray(x), Mw


def sp_eos(T, Vm, x):
    """Synthetic EoS-derived inputs (host-side inputs at the boundary): ideal-gas mixture plus a symmetric
    second-virial-like term. Same formula as oracle/make_golden.py synthetic_eos, vectorised."""
    nc = x.shape[1]
    B = -1.0e-4 * (1.0 + 0.1 * (np.arange(nc)[:, None] + np.arange(nc)[None, :]))
    eye = np.eye(nc)[None, :, :] / x[:, :, None]
    return np.ascontiguousarray(R_GAS * T[:, None, None] * (eye + 2.0 * B[None] / Vm[:, None, None]))


def _sp_cpu_worker(task):
    """Reference arm of the state-point leg: interdiffusion + thermal_diffusion_coeff at the sample's state points of ONE
    temperature through the unmodified reference (oracle/_ref): the first point pays the Omega / transfer-length fan-out."""
    T, pts = task
    from oracle import ref
    from kineticgas_b200.fluids import mie_parameters
    ids, Mw, s, e, la, lr = mie_parameters(SP_COMPS)
    m = ref.RefMie(Mw, s, e, la, lr, is_idealgas=False)
    times = []
    for Vm, x1 in pts:
        x = np.array([x1, 1.0 - x1])
        dm = sp_eos(np.array([T]), np.array([Vm]), x[None, :])[0]
        m.set_eos(dmudn=dm, cp_id=np.zeros(2), p=1e5, vm=Vm, dvdn=np.zeros(2))
        t0 = time.perf_counter()
        m.interdiffusion(T, Vm, x, SP_N)
        m.thermal_diffusion_coeff(T, Vm, x, SP_N)
        times.append(time.perf_counter() - t0)
    return times


def sp_cpu_baseline(cores, n_T=4, pts_per_T=4):
    from oracle import ref
    if not ref.available():
        return None
    T, Vm, x, _ = sp_grid(1, 0)
    Tu = np.unique(T)
    Tsel = Tu[((np.arange(n_T) + 0.5) * Tu.size / n_T).astype(int)]
    rng = np.random.default_rng(7)
    tasks = []
    for Tk in Tsel:
        idx = rng.choice(np.flatnonzero(T == Tk), pts_per_T, replace=False)
        tasks.append((float(Tk), [(float(Vm[i]), float(x[i, 0])) for i in idx]))
    t0 = time.perf_counter()
    with mp.get_context("spawn").Pool(min(cores, len(tasks))) as pool:
        res = pool.map(_sp_cpu_worker, tasks, chunksize=1)
    wall = time.perf_counter() - t0
    cold = float(np.mean([r[0] for r in res]))
    warm = float(np.mean([t for r in res for t in r[1:]]))
    nT, n = Tu.size, T.size
    est = (nT * cold + (n - nT) * warm) / cores         # one worker per core, temperatures split over workers
    return {"value": n / est, "unit": "state points/s", "cores": cores, "kind": "reference",
            "sample": f"{n_T} temperatures x {pts_per_T} state points through the reference's interdiffusion + thermal_diffusion_coeff "
                      f"(N={SP_N}); measured {cold:.2f} s for the first point of a temperature (Omega + transfer-length fan-out, 8 threads) "
                      f"and {warm * 1e3:.1f} ms per further point; extrapolated to {nT} T / {n} points on {cores} cores; {wall:.1f} s wall"}


def run_state_points(local_rank, rank, world, n_gpus, steps):
    """configs[2] on this rank's GPU: per step, build the device tables for the 64 temperatures (Omega by quadrature where the
    reference integrates, correlation elsewhere, transfer lengths, brackets) and assemble + solve all state points. Host buffers
    in, host buffers out (this IS the end-to-end path). Returns per-step milliseconds and stage split."""
    from kineticgas_b200 import MieKinGas
    from kineticgas_b200.kingas import PROP_DIFFUSION, PROP_THERMAL_DIFFUSION
    kin = MieKinGas(SP_COMPS, N=SP_N, is_idealgas=False, device=local_rank)
    from kineticgas_b200 import pinned_copy
    T, Vm, x, _ = sp_grid(n_gpus, rank)
    T, Vm, x = pinned_copy(T), pinned_copy(Vm), pinned_copy(x)          # the caller's arrays live in page-locked memory (kg_host_alloc)
    eos = {"dmudn": pinned_copy(sp_eos(T, Vm, x))}
    props = PROP_DIFFUSION | PROP_THERMAL_DIFFUSION
    out = {"ms": [], "ms_tables": [], "ms_state_kernels": [], "launches": 0}
    res = None
    for step in range(steps + 1):                 # first pass is the warm-up
        kin._caches = {}
        res = None                                # results of the previous step are consumed: their page-locked blocks return to the pool
        t0 = time.perf_counter()
        cache = kin.tcache(T, SP_N, props)
        t1 = time.perf_counter()
        res = kin.transport_batch(T, Vm, x, N=SP_N, properties=props, eos_inputs=eos, cache=cache)
        t2 = time.perf_counter()
        tm = kin.dev.last_timings()
        if step > 0:
            out["ms"].append(1e3 * (t2 - t0)); out["ms_tables"].append(1e3 * (t1 - t0)); out["ms_state_kernels"].append(tm["ms_state_kernels"])
            out["launches"] += int(tm["launches_tcache"] + tm["launches_batch"])
    out["n"] = int(T.size)
    out["stage_ms"] = {k: tm[k] for k in ("ms_omega_stage", "ms_tl_stage", "ms_bracket_stage", "ms_state_kernels", "ms_batch_call")}
    out["h2d"] = int(T.nbytes + Vm.nbytes + x.nbytes + eos["dmudn"].nbytes)
    out["d2h"] = int(res["D"].nbytes + res["DT"].nbytes + 4 * T.size)
    out["finite"] = bool(np.isfinite(res["D"]).all() and np.isfinite(res["DT"]).all())
    return out


# ------------------------------------------------------------------------------------------------ Feynman-Hibbs sweep (configs[3])
QM_NT = 10000
QM_LR = [(1, 1), (1, 2), (1, 3), (2, 2), (2, 3), (2, 4), (3, 3)]     # N = 2 viscosity + conductivity (cpp/KineticGas_mthr.cpp:62-125)
QM_F_NODE = 144.0                                                    # 52 + 23 * 4 terms (FH1-Mie)


def run_qmie_sweep(local_rank, rank, n_gpus, steps):
    """configs[3]: QuantumMie('HE', FH1) and QuantumMie('H2', FH1), Omega(l,r) of the N = 2 viscosity + conductivity set on a
    10^4-point T grid 10-500 K per GPU (T-dependent potential: effective coefficients per temperature, every Omega integrated).
    Host T array in, host Omega table out (omega_table = kg_omega_batch)."""
    from kineticgas_b200 import QuantumMie
    T = np.linspace(10.0, 500.0, QM_NT * n_gpus)[rank::n_gpus]
    ls, rs = [l for l, r in QM_LR], [r for l, r in QM_LR]
    out = {"ms": [], "finite": True}
    kins = [QuantumMie(c, FH_orders=1, N=2, is_idealgas=True, device=local_rank) for c in ("HE", "H2")]
    for step in range(steps + 1):                 # first pass is the warm-up
        t0 = time.perf_counter()
        for kin in kins:
            om = kin.omega_table(T, ls, rs)
            out["finite"] = out["finite"] and bool(np.isfinite(om).all())
        if step > 0:
            out["ms"].append(1e3 * (time.perf_counter() - t0))
    out["n_omega"] = int(2 * T.size * len(QM_LR))
    out["h2d"] = int(2 * (T.nbytes + T.size * 6 * 8)); out["d2h"] = int(2 * T.size * len(QM_LR) * 8)
    return out


# ------------------------------------------------------------------------------------------------ 5-component surface (configs[4])
FC_COMPS, FC_N = "C1,C2,C3,N2,CO2", 2
FC_GRID = (25, 40, 100)       # T x rho* x compositions = 10^5 state points


def fc_grid(n_gpus, rank):
    """25 T (250-700 K) x 40 rho* (0.01-0.6) x 100 Dirichlet(1,1,1,1,1) compositions (seed 1234, clipped to
    x_i >= 1e-3, renormalised). Weak scaling: rank k owns every n_gpus-th temperature of an n_gpus-times denser T grid."""
    from kineticgas_b200.fluids import mie_parameters, AVOGADRO
    ids, Mw, s, e, la, lr = mie_parameters(FC_COMPS)
    nT, nr, nx = FC_GRID
    T = np.linspace(250.0, 700.0, nT * n_gpus)[rank::n_gpus]
    rho = np.linspace(0.01, 0.6, nr) / s[0] ** 3
    xs = np.random.default_rng(1234).dirichlet(np.ones(5), nx)
    xs = np.clip(xs, 1e-3, None); xs /= xs.sum(axis=1, keepdims=True)
    Tg, rg, ig = np.meshgrid(T, rho, np.arange(nx), indexing="ij")
    return np.ascontiguousarray(Tg.ravel()), np.ascontiguousarray(AVOGADRO / rg.ravel()), np.ascontiguousarray(xs[ig.ravel()])


def run_five_comp(local_rank, rank, n_gpus, steps):
    """configs[4]: full transport surface (eta, lambda, D 4x4 CoN, D_T) of the 5-component RET-Mie mixture, 10^5 state points per
    GPU, device tables (15 pairs; the 5 CO2 pairs fully integrated) rebuilt every step. Host buffers in and out."""
    from kineticgas_b200 import MieKinGas
    kin = MieKinGas(FC_COMPS, N=FC_N, is_idealgas=False, device=local_rank)
    from kineticgas_b200 import pinned_copy
    T, Vm, x = fc_grid(n_gpus, rank)
    T, Vm, x = pinned_copy(T), pinned_copy(Vm), pinned_copy(x)          # the caller's arrays live in page-locked memory (kg_host_alloc)
    eos = {"dmudn": pinned_copy(sp_eos(T, Vm, x)), "cp_ideal": pinned_copy(np.full((T.size, 5), 4.5 * R_GAS))}
    out = {"ms": [], "ms_tables": []}
    res = None
    for step in range(steps + 1):
        kin._caches = {}
        res = None                                # results of the previous step are consumed: their page-locked blocks return to the pool
        t0 = time.perf_counter()
        cache = kin.tcache(T, FC_N)
        t1 = time.perf_counter()
        res = kin.transport_batch(T, Vm, x, N=FC_N, eos_inputs=eos, cache=cache)
        t2 = time.perf_counter()
        if step > 0:
            out["ms"].append(1e3 * (t2 - t0)); out["ms_tables"].append(1e3 * (t1 - t0))
    tm = kin.dev.last_timings()
    out["n"] = int(T.size)
    out["stage_ms"] = {k: tm[k] for k in ("ms_omega_stage", "ms_tl_stage", "ms_bracket_stage", "ms_state_kernels", "ms_batch_call")}
    out["h2d"] = int(T.nbytes + Vm.nbytes + x.nbytes + eos["dmudn"].nbytes + eos["cp_ideal"].nbytes)
    out["d2h"] = int(sum(res[k].nbytes for k in ("viscosity", "conductivity", "conductivity_contribs", "D", "DT")) + 4 * T.size)
    out["finite"] = bool(all(np.isfinite(res[k]).all() for k in ("viscosity", "conductivity", "D", "DT")))
    return out


# ------------------------------------------------------------------------------------------------ EoS-derived inputs on the device
# thermopack 2.2.3 (the EoS object pykingas constructs) cannot travel to the GPU box; its cost per state point of this 5-component mixture was
# measured in the build container on one core through the python calls cpp/eos_interface.h makes: chemical_potential_tv(dmudn) 203 us,
# five idealenthalpysingle calls 72 us, specific_volume(T, p) 851 us.
THERMOPACK_US_PER_POINT = {"dmudn": 203.0, "cp_ideal": 72.0, "specific_volume": 851.0}


# FP64 lane instructions (DFMA + DMUL + DADD) of one eos_hess_kernel thread = one hyper-dual Helmholtz evaluation of the 5-component model:
# ncu's SASS op counters of a 10^5-point launch, 2.4111e10 / (10^5 points x 15 Hessian entries) (profiles/r2_eos_fp64_ops.csv)
FP64_PER_EOS_EVAL = 16074.0


def run_eos_leg(local_rank, steps, peak_tflops):
    """dmu/dn (5x5) + Cp_ideal of configs[4]'s 10^5 state points from DeviceSaftVRMie (host arrays in, host arrays out), and the whole
    configs[4] surface with the EoS inputs computed on the GPU (use_eos='device') instead of handed in as arrays."""
    from kineticgas_b200 import DeviceSaftVRMie, MieKinGas
    T, Vm, x = fc_grid(1, 0)
    eos = DeviceSaftVRMie(FC_COMPS, device=local_rank)
    ms, kms = [], []
    for step in range(steps + 1):
        t0 = time.perf_counter()
        r = eos.tv_batch(T, Vm, x, dmudn=True, cp_ideal=True)
        t1 = time.perf_counter()
        if step > 0:
            ms.append(1e3 * (t1 - t0)); kms.append(eos.last_kernel_ms())
    finite = bool(np.isfinite(r["dmudn"]).all() and np.isfinite(r["cp_ideal"]).all())
    t0 = time.perf_counter()
    Vtp = eos.specific_volume_batch(T, np.full(T.size, 1e5), x)
    ms_vol = 1e3 * (time.perf_counter() - t0)
    kin = MieKinGas(FC_COMPS, N=FC_N, is_idealgas=False, use_eos="device", device=local_rank)
    full = []
    for step in range(steps + 1):
        kin._caches = {}
        t0 = time.perf_counter()
        res = kin.transport_batch(T, Vm, x, N=FC_N)
        if step > 0:
            full.append(1e3 * (time.perf_counter() - t0))
    n = int(T.size)
    tp_s = n * (THERMOPACK_US_PER_POINT["dmudn"] + THERMOPACK_US_PER_POINT["cp_ideal"]) * 1e-6
    return {"workload": f"SURVEY 8f row 1: SAFT-VR Mie dmu/dn (5x5) + Cp_ideal of the {n} state points of configs[4] ({FC_COMPS}), host arrays in and out",
            "value": n / (np.mean(ms) * 1e-3), "unit": "state points/s", "ms_per_step": float(np.mean(ms)), "kernel_ms": float(np.mean(kms)),
            "h2d_bytes_per_step": int(T.nbytes + Vm.nbytes + x.nbytes), "d2h_bytes_per_step": int(r["dmudn"].nbytes + r["cp_ideal"].nbytes),
            "all_finite": finite and bool(np.isfinite(Vtp).all()),
            "roofline": {"bound": "fp64", "kernel": "eos_tpre_kernel + eos_hess_kernel + eos_out_kernel (timed together; the Hessian kernel is 95 %)",
                         "achieved": n * 15 * FP64_PER_EOS_EVAL * 2.0 / (np.mean(kms) * 1e-3) / 1e12, "peak": peak_tflops, "unit": "TFLOP/s",
                         "frac": n * 15 * FP64_PER_EOS_EVAL / (np.mean(kms) * 1e-3) / (peak_tflops * 1e12 / 2.0), "traffic": 11.3e6,
                         "definition": "FP64 lane instructions of the 15 hyper-dual evaluations per state point (ncu-calibrated count) / (peak/2 x summed "
                                       "kernel time); ncu of the Hessian kernel alone: FP64 pipe 71.9 % busy, 77.5 % of its warp instructions are FP64, "
                                       "DRAM traffic 11 MB (inputs 5.6 MB + Hessian 12 MB stay in L2)"},
            "specific_volume": {"value": n / (ms_vol * 1e-3), "unit": "state points/s", "ms": ms_vol, "workload": "Vm(T, 1 bar, x) of the same points"},
            "five_component_with_device_eos": {"value": n / (np.mean(full) * 1e-3), "unit": "state points/s", "ms_per_step": float(np.mean(full)),
                                               "all_finite": bool(all(np.isfinite(res[k]).all() for k in ("viscosity", "conductivity", "D", "DT")))},
            "cpu_reference": {"kind": "reference", "where": "build container, one core (thermopack 2.2.3 is not on the GPU box)",
                              "us_per_point": THERMOPACK_US_PER_POINT, "seconds_for_this_batch": tp_s, "value": n / tp_s, "unit": "state points/s"}}


# ------------------------------------------------------------------------------------------------ quantum collision integrals
def run_quantum_leg(local_rank):
    """Omega^(l,r)(T) of HFD-B2 helium from partial-wave phase shifts (the pykingas default for the ab-initio potentials) on a 64-point
    T grid 10-1000 K x the five (l,r) of the N = 2 viscosity / conductivity set; host arrays in and out."""
    from kineticgas_b200 import HFD_B2
    kin = HFD_B2("HE", device=local_rank)
    T = np.geomspace(10.0, 1000.0, 64)
    l, r = [1, 1, 2, 2, 3], [1, 2, 2, 3, 3]
    ws, wa = kin.symmetry_weights()
    kin.dev.quantum_omega(0, T, l, r, ws, wa)                  # untimed warm-up pass of the same size (grows the device pool: 0.5 GB of phase-shift cache)
    t0 = time.perf_counter()
    om = kin.dev.quantum_omega(0, T, l, r, ws, wa)
    dt = time.perf_counter() - t0
    out = {"workload": "SURVEY 8f row 4: quantum Omega(l,r)(T) of HFD-B2 helium, 64 temperatures 10-1000 K x 5 (l,r), phase shifts by Numerov / JKWB",
           "value": om.size / dt, "unit": "Omega/s", "seconds": dt, "all_finite": bool(np.isfinite(om).all())}
    g = np.load(os.path.join(ROOT, "tests", "golden", "quantum_reference.npz"))
    ref_om = kin.dev.quantum_omega(0, g["hfd_he_Tom"], g["hfd_he_lr"][:, 0], g["hfd_he_lr"][:, 1], ws, wa)
    out["parity"] = {"against": "Quantum::omega of the unmodified reference build, fixture tests/golden/quantum_reference.npz (HFD-B2 He, 15 / 80 / 300 K, 5 (l,r))",
                     "tolerance": 1e-6, "max_rel_err": float(np.max(np.abs(ref_om / g["hfd_he_omega"] - 1)))}
    # The reference's ab-initio potential classes read their parameters from fluids/*.json under the reference tree, which is not on the GPU box:
    # its quantum Omega was timed in the build container on one core (fresh object per value; oracle/make_golden.py quantum prints the same calls)
    out["cpu_reference"] = {"kind": "reference", "where": "build container, one core (the reference's fluid files are not on the GPU box)",
                            "seconds_each": {"Omega(2,2) at 300 K": 1.22, "Omega(2,2) at 20 K": 13.95}, "value": 2 / (1.22 + 13.95), "unit": "Omega/s"}
    return out


# ------------------------------------------------------------------------------------------------ strong scaling (fixed total work)
def _rank_values(v, world):
    import torch
    import torch.distributed as dist
    t = torch.tensor([float(v)], dtype=torch.float64, device="cuda")
    if world == 1:
        return [float(v)]
    out = [torch.zeros_like(t) for _ in range(world)]
    dist.all_gather(out, t)
    return [float(o.item()) for o in out]


def run_strong_omega(model, rank, world, steps, barrier):
    """configs[1] as BASELINE.json states it: ONE 4096-point T* grid, dealt over the ranks by kineticgas_b200.sharding (cost-balanced),
    every rank ends up with the full Omega table. Fixed total work: strong scaling."""
    import torch
    from kineticgas_b200 import sharding
    T1 = grid(1, 0)
    deal = sharding.TemperatureDeal(T1, world, device="cuda")
    Tl = deal.local(rank)
    ls = np.array([l for l, r in LR], np.int32)
    rs = np.array([r for l, r in LR], np.int32)
    plan = model.omega_plan(Tl, ls, rs)
    loc = torch.empty((Tl.size, len(LR)), dtype=torch.float64, device="cuda")

    def share(_):
        plan.run(sync=False)
        plan.export_dev(loc.data_ptr())
        return loc

    for _ in range(3):
        sharding.sharded_temperature_tables(share, T1, rank, world, device="cuda", deal=deal)
    barrier()
    stream = torch.cuda.current_stream()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    th = 0.0
    for _ in range(steps):
        full = sharding.sharded_temperature_tables(share, T1, rank, world, device="cuda", deal=deal)
        th += plan.counters()["ms_theta"]
    e1.record(stream)
    barrier()
    ms = max(_rank_values(e0.elapsed_time(e1), world)) / steps
    th_ranks = [v / steps for v in _rank_values(th, world)]
    plan.close()
    return {"workload": "configs[1], ONE 4096-point grid x 10 (l,r) dealt over the ranks (strong scaling)", "value": T1.size * len(LR) / (ms * 1e-3),
            "unit": "Omega/s", "ms_per_step": ms, "theta_ms_per_rank": th_ranks,
            "imbalance_max_over_mean_theta_ms": max(th_ranks) / (sum(th_ranks) / len(th_ranks)),
            "modelled_cost_max_over_mean": float(deal.modelled_cost.max() / deal.modelled_cost.mean()), "all_finite": bool(torch.isfinite(full).all().item())}


def run_strong_five(local_rank, rank, world, steps, barrier):
    """configs[4] as BASELINE.json states it: 10^5 state points IN TOTAL, contiguous blocks over the ranks (kineticgas_b200.sharding.
    sharded_state_points), each rank builds the device tables of the temperatures in its block, the results (eta, lambda, D 4x4, D_T) are
    all-gathered. Host buffers in, gathered device table out."""
    import torch
    from kineticgas_b200 import MieKinGas, pinned_copy, sharding
    kin = MieKinGas(FC_COMPS, N=FC_N, is_idealgas=False, device=local_rank)
    T, Vm, x = fc_grid(1, 0)
    T, Vm, x = pinned_copy(T), pinned_copy(Vm), pinned_copy(x)
    dm, cp = pinned_copy(sp_eos(T, Vm, x)), pinned_copy(np.full((T.size, 5), 4.5 * R_GAS))
    n = int(T.size)

    phase = {"tables": 0.0, "batch": 0.0, "pack_h2d": 0.0, "gather": 0.0}

    def block(lo, hi):
        kin._caches = {}
        t0 = time.perf_counter()
        cache = kin.tcache(T[lo:hi], FC_N)
        t1 = time.perf_counter()
        r = kin.transport_batch(T[lo:hi], Vm[lo:hi], x[lo:hi], N=FC_N, eos_inputs={"dmudn": dm[lo:hi], "cp_ideal": cp[lo:hi]}, cache=cache)
        t2 = time.perf_counter()
        packed = torch.empty((hi - lo, 23), dtype=torch.float64, device="cuda")
        packed[:, 0] = torch.from_numpy(np.asarray(r["viscosity"])).cuda()
        packed[:, 1] = torch.from_numpy(np.asarray(r["conductivity"])).cuda()
        packed[:, 2:18] = torch.from_numpy(np.asarray(r["D"]).reshape(hi - lo, 16)).cuda()
        packed[:, 18:23] = torch.from_numpy(np.asarray(r["DT"])).cuda()
        torch.cuda.synchronize()
        t3 = time.perf_counter()
        phase["tables"] += t1 - t0; phase["batch"] += t2 - t1; phase["pack_h2d"] += t3 - t2
        return packed

    out = sharding.sharded_state_points(block, n, rank, world, device="cuda")     # warm-up
    barrier()
    for k in phase:
        phase[k] = 0.0
    t0 = time.perf_counter()
    for _ in range(steps):
        tg = time.perf_counter()
        out = sharding.sharded_state_points(block, n, rank, world, device="cuda")
        torch.cuda.synchronize()
        phase["gather"] += time.perf_counter() - tg
    barrier()
    ms = max(_rank_values(1e3 * (time.perf_counter() - t0), world)) / steps
    ph = {k: 1e3 * v / steps for k, v in phase.items()}
    ph["gather"] -= ph["tables"] + ph["batch"] + ph["pack_h2d"]
    ph["compute_ms_per_rank"] = _rank_values(ph["tables"] + ph["batch"] + ph["pack_h2d"], world)
    return {"workload": f"configs[4], {n} state points IN TOTAL in contiguous blocks over the ranks (strong scaling), tables of each block's "
                        "temperatures rebuilt every step, results all-gathered on the device", "value": n / (ms * 1e-3), "unit": "state points/s",
            "ms_per_step": ms, "rank0_phase_ms": ph, "all_finite": bool(torch.isfinite(out).all().item()),
            "note": "rank0_phase_ms: device tables of the block's temperatures / kg_transport_batch / results back to the device for the collective / "
                    "all_gather + wait for the slowest rank. The table stage (Omega passes and transfer lengths of ~25/N temperatures) has a fixed "
                    "latency floor, which is what bends the strong-scaling curve"}


# ------------------------------------------------------------------------------------------------ parity + configs[0]
def parity_block(model):
    """Omega of the headline workload against the reference-built fixture (tests/golden/regimes_reference.npz: 48 temperatures of this
    grid on all five lattice regimes + the 8 regime edges, all ten (l,r); generated by oracle/make_golden.py from the unmodified
    reference). Runs outside the timed region; max relative error per lattice regime."""
    path = os.path.join(ROOT, "tests", "golden", "regimes_reference.npz")
    if not os.path.exists(path):
        return None
    g = np.load(path)
    T = np.concatenate([g["grid_T"], g["edges_T"]])
    ref = np.concatenate([g["grid_omega"], g["edges_omega"]])
    om = model.omega_batch(T, g["lr"][:, 0], g["lr"][:, 1])[:, 0, :]
    err = np.abs(om / ref - 1.0)
    reg = np.where(T < 20, 4, np.where(T < 35, 3, np.where(T < 50, 2, np.where(T < 70, 1, 0))))
    names = ["T>=70K", "50-70K", "35-50K", "20-35K", "<20K"]
    e = np.sort(err.ravel())
    return {"against": "Omega from the unmodified reference build (oracle/_ref), fixture tests/golden/regimes_reference.npz",
            "tolerance": 1e-6, "n_checked": int(err.size), "max_rel_err": float(e[-1]), "p999_rel_err": float(e[int(0.999 * (e.size - 1))]),
            "max_rel_err_per_regime": {names[k]: float(err[reg == k].max()) for k in range(5) if (reg == k).any()},
            "n_per_regime": {names[k]: int((reg == k).sum() * err.shape[1]) for k in range(5) if (reg == k).any()}}


def config0_inputs():
    """Vm(300 K, 1 bar) and Cp_ideal of argon from the EoS pykingas would construct (thermopack 2.2.3 saftvrmie('AR'), evaluated once by
    `python -m oracle.make_golden eos0` from the wheel vendored in the reference tree; tests/golden/config0_eos.npz)."""
    g = np.load(os.path.join(ROOT, "tests", "golden", "config0_eos.npz"))
    return float(g["T"]), float(g["p"]), float(g["Vm"]), float(g["cp_ideal"])


def config0_latency(local_rank):
    """BASELINE configs[0]: MieKinGas('AR'), N = 2, T = 300 K, p = 1 bar, viscosity_tp + thermal_conductivity_tp of ONE state point, cold
    (new model object, empty caches) and warm. The two EoS-derived numbers the calls need, Vm(T, p) and Cp_ideal, are the ones the
    reference's own EoS returns (config0_inputs); the work is <= 4 numeric Omega + correlation values + MTL + ETL + one state point.
    Seconds, host clock around the public calls."""
    from kineticgas_b200 import MieKinGas
    T, p, Vm, cp = config0_inputs()
    eos = {"cp_ideal": np.array([[cp]])}
    t0 = time.perf_counter()
    kin = MieKinGas("AR", N=2, is_idealgas=False, device=local_rank)
    t1 = time.perf_counter()
    eta = kin.viscosity(T, Vm, [1.0], N=2)
    lam = kin.thermal_conductivity(T, Vm, [1.0], N=2, eos_inputs=eos)
    t2 = time.perf_counter()
    kin._caches = {}
    t3 = time.perf_counter()
    kin.viscosity(T, Vm, [1.0], N=2)
    kin.thermal_conductivity(T, Vm, [1.0], N=2, eos_inputs=eos)
    t4 = time.perf_counter()
    return {"workload": "configs[0]: MieKinGas('AR') viscosity_tp + thermal_conductivity_tp, T = 300 K, p = 1 bar, N = 2, one state point; "
                        "Vm(T,p) and Cp_ideal from thermopack 2.2.3 saftvrmie('AR') (fixture)",
            "seconds_cold": t2 - t0, "seconds_model_create": t1 - t0, "seconds_tables_rebuilt": t4 - t3, "unit": "s",
            "viscosity_Pa_s": float(eta), "conductivity_W_mK": float(lam)}


def _cfg0_ref_worker(_):
    from oracle import ref
    from kineticgas_b200.fluids import mie_parameters
    ids, Mw, s, e, la, lr = mie_parameters("AR")
    T, p, Vm, cp = config0_inputs()
    R = 8.31446261815324
    t0 = time.perf_counter()
    m = ref.RefMie(Mw, s, e, la, lr, is_idealgas=False)
    m.set_eos(dmudn=np.array([[R * T]]), cp_id=np.array([cp]), p=p, vm=Vm, dvdn=np.array([Vm]))
    eta = m.viscosity(T, Vm, [0.5, 0.5], 2)
    lam = m.thermal_conductivity(T, Vm, [0.5, 0.5], 2)
    return time.perf_counter() - t0, eta, lam


def config0_reference():
    from oracle import ref
    if not ref.available():
        return None
    with mp.get_context("spawn").Pool(1) as pool:
        sec, eta, lam = pool.apply(_cfg0_ref_worker, (0,))
    return {"seconds_cold": sec, "viscosity_Pa_s": eta, "conductivity_W_mK": lam, "kind": "reference", "threads": 8}


# ------------------------------------------------------------------------------------------------ clocks
class ClockSampler:
    """nvidia-smi polled every 50 ms from before the warm-up (its start-up takes longer than a short timed region); stop(t0, t1)
    keeps the samples taken between the wall-clock marks of the timed region, widened by 150 ms on both sides (the GPU is busy
    with warm-up steps before and the end-to-end loop after)."""
    Q = "timestamp,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "50"],
                                      stdout=self.f, stderr=subprocess.DEVNULL)
        except Exception:
            self.p = None

    def stop(self, t0=None, t1=None):
        import datetime
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": []}
        if self.p is None:
            return out
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except Exception:
            self.p.kill()
        self.f.flush()
        rows = [l.strip().split(", ") for l in open(self.f.name) if l.strip()]
        os.unlink(self.f.name)
        sm, reasons = [], set()
        for r in rows:
            try:
                ts = datetime.datetime.strptime(r[0].strip(), "%Y/%m/%d %H:%M:%S.%f").timestamp()
                if t0 is not None and not (t0 - 0.15 <= ts <= t1 + 0.15):
                    continue
                r = r[1:]
                sm.append(float(r[0]))
                out["sm_max_mhz"] = float(r[1])
                for name, v in zip(["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"], r[3:7]):
                    if v.strip().lower() == "active":
                        reasons.add(name)
            except Exception:
                pass
        if sm:
            busy = [x for x in sm if x > 0.5 * max(sm)]
            out["sm_mhz"] = float(np.median(busy))
            out["samples"] = len(sm)
        out["reasons"] = sorted(reasons)
        return out


# ------------------------------------------------------------------------------------------------ main
_REAL_STDOUT = None


def guard_stdout():
    """Libraries (NCCL's version banner, for one) write to file descriptor 1; the contract is ONE JSON line on stdout. Everything written
    to fd 1 from here on goes to stderr, emit() writes the line to the original stdout."""
    global _REAL_STDOUT
    if _REAL_STDOUT is None:
        sys.stdout.flush()
        _REAL_STDOUT = os.dup(1)
        os.dup2(2, 1)


def emit(line):
    sys.stdout.flush()
    fd = _REAL_STDOUT if _REAL_STDOUT is not None else 1
    os.write(fd, (line + "\n").encode())


def main():
    guard_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--cpu-sample", type=int, default=0, help="Omega count of the CPU sample (default: scaled to the host cores)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-state-points", action="store_true", help="skip the RET state-point leg (profiling runs of the Omega kernels)")
    ap.add_argument("--no-extra-configs", action="store_true", help="skip the configs[3] (Feynman-Hibbs sweep) and configs[4] (5-component surface) legs")
    ap.add_argument("--nt", type=int, default=0, help="profiling only: shrink the per-GPU T grid (a line produced with --nt is not a bench value)")
    args = ap.parse_args()
    if args.nt:
        global NT_PER_GPU
        NT_PER_GPU = args.nt
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local_rank = int(os.environ.get("LOCAL_RANK", 0))
    n_gpus = max(args.gpus, world)
    config = {"workload": "configs[1]: LJF (Mie 12-6) Omega(l,r), 1<=l<=r<=4, 4096-point T* grid 0.3-100 (T = 30 T* K) per GPU",
              "n_T_per_gpu": NT_PER_GPU, "n_lr": len(LR), "omega_per_step": NT_PER_GPU * len(LR) * n_gpus,
              "l2": "working set (cos(chi) lattices + masks, ~0.5 GB per GPU) is rebuilt every step and exceeds the 126 MB L2",
              "sharding": "kineticgas_b200.sharding: T grid dealt cost-balanced over ranks (equal counts per lattice regime), no data-path "
                          "collective; one all_gather of the Omega tables per step puts the full table on every rank"}
    cores = host_cores()

    if args.impl == "reference":
        if rank != 0:
            return
        n_sample = args.cpu_sample or min(128, max(32, 2 * cores))
        vals = []
        for step in range(args.warmup):
            run_cpu_reference(n_gpus, n_sample, cores, warm=True)      # untimed: starts the pool, pages the library in (hot temperatures only)
        for step in range(args.steps):
            cb, _ = run_cpu_reference(n_gpus, n_sample, min(cores, n_sample))
            vals.append(cb)
        v = float(np.mean([c["value"] for c in vals]))
        cb = dict(vals[-1]); cb["value"] = v
        cb["as_shipped_single_process"] = run_cpu_as_shipped(n_gpus)
        emit(json.dumps({"impl": "reference", "metric": METRIC, "value": v, "unit": "Omega/s", "n_gpus": n_gpus, "steps": args.steps,
                          "warmup": args.warmup, "ms_per_step": 1e3 * n_sample / v, "higher_is_better": True, "scaling": "weak",
                          "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": config, "cpu_baseline": cb,
                          "e2e": {"value": v, "unit": "Omega/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))
        return

    import torch
    import torch.distributed as dist
    from kineticgas_b200 import DeviceModel, mie_desc, measure_fp64_peak
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    Mw, s, e, la, lr, KB, NA = ljf()
    model = DeviceModel([mie_desc(s, e * KB, la, lr)], [Mw * 1e-3 / NA], True, device=local_rank)
    stream = torch.cuda.current_stream()
    model.set_stream(stream.cuda_stream)
    # The multi-GPU path is the product's own sharding layer (kineticgas_b200/sharding.py): the temperatures of the N-times denser grid
    # are dealt cost-balanced over the ranks (every rank gets the same number of points of every lattice regime), each rank integrates
    # its share, one all_gather puts the table together in grid order on every rank.
    from kineticgas_b200 import sharding
    T_all = np.sort(np.concatenate([grid(world, k) for k in range(world)]))
    deal = sharding.TemperatureDeal(T_all, world, device="cuda")
    T = deal.local(rank)
    ls = np.array([l for l, r in LR], np.int32)
    rs = np.array([r for l, r in LR], np.int32)
    plan = model.omega_plan(T, ls, rs)
    n_out = T.size * len(LR)
    om_local = torch.empty((T.size, len(LR)), dtype=torch.float64, device="cuda")
    om_all = None

    def compute_share(_T_local):
        plan.run(sync=False)
        plan.export_dev(om_local.data_ptr())
        return om_local

    def step():
        nonlocal om_all
        om_all = sharding.sharded_temperature_tables(compute_share, T_all, rank, world, device="cuda", deal=deal)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    sampler = ClockSampler(local_rank) if rank == 0 else None
    for _ in range(max(args.warmup, 3)):
        step()
    barrier()
    peak_tflops, peak_mhz = measure_fp64_peak(local_rank)
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    theta_ms, nodes, mesh_ms, launches, exec_nodes, redo, n_theta, tier2, mesh_cells = 0.0, 0.0, 0.0, 0, 0.0, 0.0, 0.0, 0.0, 0.0
    barrier()
    wall0 = time.time()
    ev0.record(stream)
    for _ in range(args.steps):
        step()
        c = plan.counters()
        theta_ms += c["ms_theta"]; mesh_ms += c["ms_mesh"]; launches += int(c["launches"]) + 1
        nodes += c["chi_nodes"] + c["lim_nodes"]
        q = plan.quadrature_stats()
        exec_nodes += q["executed_nodes"]; redo += q["redo_integrals"]; n_theta += q["theta_integrals"]; tier2 += q["stride8_integrals"]
        mesh_cells += q["mesh_cells"]
    ev1.record(stream)
    barrier()
    wall1 = time.time()
    ms = ev0.elapsed_time(ev1)
    tmax = torch.tensor([ms], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
    ms = float(tmax.item())
    n_total = T_all.size * len(LR)
    value = n_total * args.steps / (ms * 1e-3)

    # end to end through the public host-buffer call: plan creation (H2D of the work description), kernels, D2H of the tables
    for _ in range(max(args.warmup, 3)):          # untimed warm-up of this path too (first plan of this size grows the device pool)
        p2 = model.omega_plan(T, ls, rs)
        p2.run(sync=False)
        p2.fetch()
        p2.close()
    barrier()
    e2e0 = time.perf_counter()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    h2d = d2h = 0
    for _ in range(args.steps):
        p2 = model.omega_plan(T, ls, rs)
        p2.run(sync=False)
        om_host = p2.fetch()
        cc = p2.counters()
        h2d, d2h = int(cc["h2d_bytes"]) + T.nbytes, int(cc["d2h_bytes"])
        p2.close()
        if world > 1:
            deal.gather(torch.from_numpy(om_host.reshape(T.size, len(LR))).cuda(), rank)
    e1.record(stream)
    barrier()
    e2e_ms = max(e0.elapsed_time(e1), 1e3 * (time.perf_counter() - e2e0))
    t2 = torch.tensor([e2e_ms], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t2, op=dist.ReduceOp.MAX)
    e2e_value = n_total * args.steps / (float(t2.item()) * 1e-3)

    clocks = sampler.stop(wall0, wall1) if sampler else None

    # RET state points (configs[2]): every rank runs its own grid; max over ranks
    sp = None
    if not args.no_state_points:
        plan.close()
        barrier()
        sp = run_state_points(local_rank, rank, world, n_gpus, max(1, min(args.steps, 3)))
        barrier()
        t3 = torch.tensor([float(np.mean(sp["ms"]))], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(t3, op=dist.ReduceOp.MAX)
        sp["ms_max"] = float(t3.item())

    qm = fc = None
    if not args.no_extra_configs and not args.no_state_points:
        barrier()
        qm = run_qmie_sweep(local_rank, rank, n_gpus, max(1, min(args.steps, 2)))
        barrier()
        fc = run_five_comp(local_rank, rank, n_gpus, max(1, min(args.steps, 2)))
        barrier()
        t4 = torch.tensor([float(np.mean(qm["ms"])), float(np.mean(fc["ms"]))], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(t4, op=dist.ReduceOp.MAX)
        qm["ms_max"], fc["ms_max"] = float(t4[0].item()), float(t4[1].item())

    par = cfg0 = eos_leg = quantum_leg = None
    if rank == 0 and not args.nt and not args.no_extra_configs and not args.no_state_points:
        eos_leg = run_eos_leg(local_rank, max(1, min(args.steps, 2)), peak_tflops)
        quantum_leg = run_quantum_leg(local_rank)
    if rank == 0 and not args.nt:
        par = parity_block(model)
        if not args.no_state_points:
            cfg0 = config0_latency(local_rank)
            if not args.no_cpu_baseline:
                cfg0["cpu_reference"] = config0_reference()

    strong = None
    if not args.no_extra_configs and not args.no_state_points and not args.nt:
        barrier()
        strong = {"omega": run_strong_omega(model, rank, world, max(1, min(args.steps, 3)), barrier),
                  "five_component": run_strong_five(local_rank, rank, world, max(1, min(args.steps, 2)), barrier)}
        barrier()
    th_ranks = _rank_values(theta_ms / args.steps, world)

    if rank == 0:
        from kineticgas_b200 import _lib
        import __graft_entry__ as ge
        # FP64-pipe instructions issued per executed node evaluation, per theta integral and per root finding, counted in the SASS of
        # this build (profiles/r2_theta_sass.md) and cross-checked against ncu's smsp__sass_thread_inst_executed_op_{dfma,dmul,dadd}
        fp64_theta = exec_nodes * FP64_PER_NODE
        fp64_setup = n_theta * FP64_PER_SETUP
        issue_peak = peak_tflops * 1e12 / 2.0            # FP64 lane instructions per second (a DFMA counts 2 flop)
        theta_s = theta_ms * 1e-3
        algo_tflops = nodes * F_NODE / theta_s / 1e12 if theta_ms > 0 else None
        out = {"metric": METRIC, "value": value, "unit": "Omega/s", "n_gpus": n_gpus, "steps": args.steps, "warmup": max(args.warmup, 3),
               "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
               "data": "synthetic", "config": config, "clocks": clocks, "gpu_launches": launches,
               "build": {"library": _lib.lib().kg_build_info().decode(), "tree_src_sha256": ge.source_sha()},
               "e2e": {"value": e2e_value, "unit": "Omega/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h},
               "parity": par,
               "roofline": {"bound": "fp64", "kernel": "theta_setup_kernel<2> + theta_kernel<2> (root finding + tanh-sinh sums of one pass, timed together)",
                            "achieved": (fp64_theta + fp64_setup) * 2.0 / theta_s / 1e12 if theta_ms > 0 else None, "peak": peak_tflops, "unit": "TFLOP/s",
                            "frac": (fp64_theta + fp64_setup) / theta_s / issue_peak if theta_ms > 0 else None,
                            "frac_step": (fp64_theta + fp64_setup + mesh_cells * FP64_PER_CELL) / (ms * 1e-3) / issue_peak,
                            "traffic": NCU_THETA_TRAFFIC_BYTES,
                            "traffic_note": "dram__bytes_read.sum + dram__bytes_write.sum of the largest theta_kernel<2> + theta_setup_kernel<2> launch pair of a "
                                            "full-size step (pass 5: 19.2 M integrals), one `ncu --set full` capture (profiles/r2_ncu_theta_kernel_digest.txt, "
                                            "r2_ncu_theta_setup_digest.txt): 1.02 + 0.57 GB against 1.38 GB algorithmic (per integral: set-up reads the 8 B "
                                            "work-list entry and writes 24 B of parameters, the sum reads both and writes an 8 B result) -- 0.4 TB/s over "
                                            "the pair's 4.2 ms, the kernels are not HBM-bound",
                            "definition": "frac = FP64-pipe lane instructions executed by the two theta kernels (device counters of node evaluations, incl. "
                                          "rejected attempts and node-by-node sums, and of root findings, x the per-unit instruction counts calibrated with "
                                          "ncu) / (peak/2 x their summed duration); `achieved` prices every such instruction as one DFMA (2 flop). "
                                          "frac_step = the same plus the mesh kernel's FP64 instructions over the whole step time",
                            "peak_source": "kg_measure_fp64_peak: DFMA chain measured in this run (MEASURED_PEAKS.json has no FP64 figure; "
                                           "datasheet 37 TFLOP/s)",
                            "algorithmic_equivalent_tflops": algo_tflops,
                            "algorithmic": f"{F_NODE:.0f} nominal flop per tanh-sinh node of the reference's sums (SURVEY 8d) x touched nodes / kernel time: "
                                           "NOT a hardware fraction (the kernel removes the per-node log/exp/div and strides the sum)",
                            "nodes_per_step": nodes / args.steps, "chi_evals_per_step": c["chi_evals"], "kernel_ms_per_step": theta_ms / args.steps,
                            "kernel_share_of_step": theta_ms / ms, "mesh_ms_per_step": mesh_ms / args.steps, "passes": c["passes"],
                            "executed": {"node_evaluations_per_step": exec_nodes / args.steps, "fraction_of_reference_nodes": exec_nodes / nodes if nodes else None,
                                         "theta_integrals_per_step": n_theta / args.steps, "second_tier_stride8_per_step": tier2 / args.steps,
                                         "summed_node_by_node_per_step": redo / args.steps,
                                         "fp64_instr_per_node_evaluation": FP64_PER_NODE,
                                         "fp64_instr_per_root_finding": FP64_PER_SETUP, "fp64_instr_per_mesh_cell": FP64_PER_CELL,
                                         "mesh_cells_per_step": mesh_cells / args.steps}}}
        if cfg0 is not None:
            out["config0_latency"] = cfg0
        out["load_balance"] = {"theta_ms_per_rank": th_ranks, "max_over_mean": max(th_ranks) / (sum(th_ranks) / len(th_ranks)),
                               "modelled_cost_max_over_mean": float(deal.modelled_cost.max() / deal.modelled_cost.mean())}
        if strong is not None:
            out["strong_scaling"] = strong
        if sp is not None:
            out["gpu_launches"] = launches + sp["launches"]
            n_sp = sp["n"] * world
            out["state_points"] = {"workload": f"configs[2]: RET-Mie {SP_COMPS}, N={SP_N}, interdiffusion (CoN) + thermal_diffusion_coeff on a "
                                               f"{SP_GRID[0]}x{SP_GRID[1]}x{SP_GRID[2]} T x rho* x x grid per GPU, device tables rebuilt every step",
                                   "value": n_sp / (sp["ms_max"] * 1e-3), "unit": "state points/s", "ms_per_step": sp["ms_max"],
                                   "ms_tables": float(np.mean(sp["ms_tables"])), "stage_ms": sp["stage_ms"], "n_per_gpu": sp["n"],
                                   "e2e": {"value": n_sp / (sp["ms_max"] * 1e-3), "unit": "state points/s", "h2d_bytes_per_step": sp["h2d"],
                                           "d2h_bytes_per_step": sp["d2h"]},
                                   "all_finite": sp["finite"],
                                   "roofline": state_roofline("state_points", sp["n"], sp["h2d"], sp["d2h"], float(np.mean(sp["ms_state_kernels"]))),
                                   "note": "host buffers in and out through kg_tcache_create + kg_transport_batch (the e2e path); EoS inputs (dmu/dn) "
                                           "are synthetic arrays prepared outside the timed region"}
            if not args.no_cpu_baseline:
                out["state_points"]["cpu_baseline"] = sp_cpu_baseline(cores)
        if qm is not None:
            out["qmie_sweep"] = {"workload": f"configs[3]: QuantumMie HE + H2, Feynman-Hibbs order 1 (4 power-law terms, C_eff(T) per temperature), "
                                             f"{len(QM_LR)} Omega(l,r) of the N=2 viscosity + conductivity set on a {QM_NT}-point T grid 10-500 K per GPU",
                                 "value": qm["n_omega"] * world / (qm["ms_max"] * 1e-3), "unit": "Omega/s", "ms_per_step": qm["ms_max"],
                                 "e2e": {"value": qm["n_omega"] * world / (qm["ms_max"] * 1e-3), "unit": "Omega/s",
                                         "h2d_bytes_per_step": qm["h2d"], "d2h_bytes_per_step": qm["d2h"]},
                                 "f_node": QM_F_NODE, "all_finite": qm["finite"],
                                 "note": "host T grid in, host Omega table out through kg_omega_batch (plan creation, kernels, D2H inside the timed region)"}
            n_fc = fc["n"] * world
            out["five_component"] = {"workload": f"configs[4]: RET-Mie {FC_COMPS}, N={FC_N}, eta + lambda + D (4x4, CoN) + D_T on "
                                                 f"{FC_GRID[0]} T x {FC_GRID[1]} rho* x {FC_GRID[2]} compositions = {fc['n']} state points per GPU, device tables rebuilt every step",
                                     "value": n_fc / (fc["ms_max"] * 1e-3), "unit": "state points/s", "ms_per_step": fc["ms_max"],
                                     "ms_tables": float(np.mean(fc["ms_tables"])), "stage_ms": fc["stage_ms"],
                                     "e2e": {"value": n_fc / (fc["ms_max"] * 1e-3), "unit": "state points/s", "h2d_bytes_per_step": fc["h2d"],
                                             "d2h_bytes_per_step": fc["d2h"]},
                                     "all_finite": fc["finite"],
                                     "roofline": state_roofline("five_component", fc["n"], fc["h2d"], fc["d2h"], fc["stage_ms"]["ms_state_kernels"]),
                                     "note": "host buffers in and out through kg_tcache_create + kg_transport_batch; EoS inputs (dmu/dn, Cp_id) are synthetic "
                                             "arrays prepared outside the timed region"}
        if eos_leg is not None:
            out["eos_inputs"] = eos_leg
        if quantum_leg is not None:
            out["quantum_omega"] = quantum_leg
        if not args.no_cpu_baseline:
            n_sample = args.cpu_sample or min(64, max(16, cores))
            cb, _ = run_cpu_reference(n_gpus, n_sample, min(cores, n_sample))
            out["cpu_baseline"] = cb
        emit(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
