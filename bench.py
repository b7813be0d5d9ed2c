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
F_NODE = 98.0    # nominal FP64 flop per tanh-sinh node for a 2-term (Mie/LJ) potential, SURVEY.md 8(d): 52 + 23 n_terms
METRIC = "collision integrals/s"


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
        m = ref.RefMie([Mw], [s], [e], [la], [lr], is_idealgas=True)
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


def run_cpu_reference(n_gpus, n_sample, cores):
    from oracle import ref
    kind = "reference" if ref.available() else "port"
    tasks = [(T, l, r, kind) for (T, l, r) in cpu_sample(n_gpus, n_sample)]
    tasks.sort(key=lambda t: t[0])          # coldest (most expensive) first
    t0 = time.perf_counter()
    with mp.get_context("spawn").Pool(cores) as pool:
        res = pool.map(_cpu_worker, tasks, chunksize=1)
    wall = time.perf_counter() - t0
    return {"value": len(tasks) / wall, "unit": METRIC, "cores": cores, "kind": kind,
            "sample": f"{len(tasks)} Omega: one T per equal-count stratum of the sorted T grid, (l,r) cycling over the 10 "
                      f"combinations, {wall:.1f} s wall, {sum(r[4] for r in res):.1f} core-s"}, res


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


# ------------------------------------------------------------------------------------------------ clocks
class ClockSampler:
    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100"],
                                      stdout=self.f, stderr=subprocess.DEVNULL)
        except Exception:
            self.p = None

    def stop(self):
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
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--cpu-sample", type=int, default=0, help="Omega count of the CPU sample (default: scaled to the host cores)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
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
              "sharding": "T grid interleaved over ranks, no data-path collective; one all_gather of the Omega tables per step"}
    cores = host_cores()

    if args.impl == "reference":
        if rank != 0:
            return
        n_sample = args.cpu_sample or min(128, max(32, cores))
        vals = []
        for step in range(args.warmup + args.steps):
            cb, _ = run_cpu_reference(n_gpus, n_sample, min(cores, n_sample))
            if step >= args.warmup:
                vals.append(cb)
        v = float(np.mean([c["value"] for c in vals]))
        cb = dict(vals[-1]); cb["value"] = v
        print(json.dumps({"impl": "reference", "metric": METRIC, "value": v, "unit": "Omega/s", "n_gpus": n_gpus, "steps": args.steps,
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
    T = grid(n_gpus, rank)
    ls = np.array([l for l, r in LR], np.int32)
    rs = np.array([r for l, r in LR], np.int32)
    plan = model.omega_plan(T, ls, rs)
    n_out = T.size * len(LR)
    om_local = torch.empty(n_out, dtype=torch.float64, device="cuda")
    om_all = torch.empty(n_out * world, dtype=torch.float64, device="cuda") if world > 1 else None

    def step():
        plan.run(sync=False)
        plan.export_dev(om_local.data_ptr())
        if world > 1:
            dist.all_gather_into_tensor(om_all, om_local)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(args.warmup, 3)):
        step()
    barrier()
    peak_tflops, peak_mhz = measure_fp64_peak(local_rank)
    sampler = ClockSampler(local_rank) if rank == 0 else None
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    theta_ms, nodes, mesh_ms, launches = 0.0, 0.0, 0.0, 0
    barrier()
    ev0.record(stream)
    for _ in range(args.steps):
        step()
        c = plan.counters()
        theta_ms += c["ms_theta"]; mesh_ms += c["ms_mesh"]; launches += int(c["launches"]) + 1
        nodes += c["chi_nodes"] + c["lim_nodes"]
    ev1.record(stream)
    barrier()
    ms = ev0.elapsed_time(ev1)
    clocks = sampler.stop() if sampler else None
    tmax = torch.tensor([ms], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
    ms = float(tmax.item())
    value = n_out * world * args.steps / (ms * 1e-3)

    # end to end through the public host-buffer call: plan creation (H2D of the work description), kernels, D2H of the tables
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
            dist.all_gather_into_tensor(om_all, torch.from_numpy(om_host.ravel()).cuda())
    e1.record(stream)
    barrier()
    e2e_ms = max(e0.elapsed_time(e1), 1e3 * (time.perf_counter() - e2e0))
    t2 = torch.tensor([e2e_ms], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t2, op=dist.ReduceOp.MAX)
    e2e_value = n_out * world * args.steps / (float(t2.item()) * 1e-3)

    if rank == 0:
        c = plan.counters()
        achieved = nodes * F_NODE / (theta_ms * 1e-3) / 1e12 if theta_ms > 0 else None
        out = {"metric": METRIC, "value": value, "unit": "Omega/s", "n_gpus": n_gpus, "steps": args.steps, "warmup": max(args.warmup, 3),
               "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
               "data": "synthetic", "config": config, "clocks": clocks, "gpu_launches": launches,
               "e2e": {"value": e2e_value, "unit": "Omega/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h},
               "roofline": {"bound": "fp64", "kernel": "theta_kernel<2>", "achieved": achieved, "peak": peak_tflops, "unit": "TFLOP/s",
                            "frac": (achieved / peak_tflops) if achieved else None, "traffic": None,
                            "peak_source": "kg_measure_fp64_peak: DFMA chain measured in this run (MEASURED_PEAKS.json has no FP64 figure; "
                                           "datasheet 37 TFLOP/s)",
                            "algorithmic": f"{F_NODE:.0f} nominal flop per tanh-sinh node (SURVEY 8d) x touched nodes",
                            "nodes_per_step": nodes / args.steps, "chi_evals_per_step": c["chi_evals"], "kernel_ms_per_step": theta_ms / args.steps,
                            "kernel_share_of_step": theta_ms / ms, "mesh_ms_per_step": mesh_ms / args.steps, "passes": c["passes"]}}
        if not args.no_cpu_baseline:
            n_sample = args.cpu_sample or min(64, max(16, cores))
            cb, _ = run_cpu_reference(n_gpus, n_sample, min(cores, n_sample))
            out["cpu_baseline"] = cb
        print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
