"""Multi-GPU plumbing of the hot path: the path shards, it has no exchange step inside the math.

Stage 1 (collision integrals / transfer lengths): work items are (pair, T); all pairs of one T live on one rank, so the
unit that is dealt out is a temperature. Cost depends on the lattice regime of cpp/Spherical.cpp:60-84 (a T < 20 K item is
~30x a T >= 70 K item), so temperatures are dealt longest-first to the least-loaded rank; the result tables are put back
in caller order by ONE all_gather. Stage 2 (state points): contiguous blocks, one all_gather of the result arrays.
One process per GPU; `torch.distributed` (NCCL on the GPUs, gloo in the CPU tests) is only the transport.
"""
import numpy as np
import torch
import torch.distributed as dist

# relative cost of one (pair, T) item by lattice regime (touched nodes x theta-integral length, measured on configs[1])
REGIME_COST = np.array([1.0, 2.5, 2.5, 6.0, 30.0])


def regime_index(T):
    """0: T >= 70 K, 1: 50-70, 2: 35-50, 3: 20-35, 4: < 20 (cpp/Spherical.cpp:60-84)."""
    T = np.asarray(T, float)
    return np.where(T < 20, 4, np.where(T < 35, 3, np.where(T < 50, 2, np.where(T < 70, 1, 0))))


def partition_temperatures(T, world):
    """Cost-balanced deal of temperature indices over ranks. The cost of an item takes one of five values (its lattice regime), so
    dealing every regime's items round-robin -- most expensive regime first, each regime starting at the rank where the previous one
    stopped -- gives every rank the same number of items of every regime to within one. Returns a list of index arrays, one per
    rank, each sorted ascending; every index appears exactly once."""
    T = np.asarray(T, float)
    reg = regime_index(T)
    owner = np.empty(T.size, dtype=np.int64)
    start = 0
    for k in np.argsort(-REGIME_COST, kind="stable"):
        idx = np.flatnonzero(reg == k)
        owner[idx] = (start + np.arange(idx.size)) % world
        start = (start + idx.size) % world
    return [np.flatnonzero(owner == r) for r in range(world)]


def partition_cost(T, parts):
    """Modelled cost of every rank's share (sum of REGIME_COST): what the deal balances; bench.py compares it with measured kernel time."""
    reg = regime_index(np.asarray(T, float))
    return np.array([REGIME_COST[reg[p]].sum() for p in parts])


def partition_blocks(n, world):
    """Contiguous blocks of state points: rank r owns [starts[r], starts[r+1])."""
    base, rem = divmod(n, world)
    sizes = np.array([base + (1 if r < rem else 0) for r in range(world)], dtype=np.int64)
    return np.concatenate([[0], np.cumsum(sizes)])


def _all_gather_padded(local, counts, group=None):
    """all_gather of per-rank tensors whose leading sizes differ (counts[r] rows on rank r)."""
    world = len(counts)
    nmax = int(max(counts))
    pad = torch.zeros((nmax,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    pad[: local.shape[0]] = local
    buf = torch.empty((world * nmax,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    dist.all_gather_into_tensor(buf, pad.contiguous(), group=group)
    return buf.view((world, nmax) + tuple(local.shape[1:]))


class TemperatureDeal:
    """The deal of one temperature grid over `world` ranks, with the index tensors of the gather kept on `device` (a sweep that
    repeats the same grid pays for the partition once)."""

    def __init__(self, T, world, device="cpu"):
        self.T = np.asarray(T, float)
        self.world = world
        self.parts = partition_temperatures(self.T, world)
        self.counts = [len(p) for p in self.parts]
        self.index = [torch.as_tensor(p, device=device) for p in self.parts]
        self.modelled_cost = partition_cost(self.T, self.parts)

    def local(self, rank):
        return self.T[self.parts[rank]]

    def gather(self, local, rank, group=None):
        """local [counts[rank], ...] of this rank -> the full table [nT, ...] in the caller's temperature order, on every rank."""
        out = torch.empty((self.T.size,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
        if self.world == 1:
            out[self.index[0]] = local
            return out
        g = _all_gather_padded(local, self.counts, group)
        for r in range(self.world):
            if self.counts[r]:
                out[self.index[r]] = g[r, : self.counts[r]]
        return out


def sharded_temperature_tables(compute_fn, T, rank, world, device="cpu", group=None, deal=None):
    """Every rank returns the full table [nT, ...] in the caller's temperature order. compute_fn(T_local) -> tensor
    [n_local, ...] (on `device`) is this rank's share of stage 1 (e.g. OmegaPlan run + export). deal: a TemperatureDeal of the
    same grid built earlier (sweeps that repeat one grid)."""
    if deal is None:
        deal = TemperatureDeal(T, world, device)
    local = torch.as_tensor(compute_fn(deal.local(rank)), device=device)
    return deal.gather(local, rank, group)


def sharded_state_points(compute_fn, n, rank, world, device="cpu", group=None):
    """Stage 2: compute_fn(lo, hi) -> tensor [hi - lo, ...] for this rank's block; returns the gathered [n, ...]."""
    starts = partition_blocks(n, world)
    lo, hi = int(starts[rank]), int(starts[rank + 1])
    local = torch.as_tensor(compute_fn(lo, hi), device=device)
    if world == 1:
        return local
    counts = [int(starts[r + 1] - starts[r]) for r in range(world)]
    g = _all_gather_padded(local, counts, group)
    return torch.cat([g[r, : counts[r]] for r in range(world)], dim=0)
