"""Host-side multi-GPU logic (kineticgas_b200/sharding.py) on CPU: world_size 2 and 3 over gloo. The device work is replaced
by a deterministic function of the inputs; what is tested is the deal, the unequal-size all_gather and the re-ordering."""
import os
import socket
import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from kineticgas_b200 import sharding


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _fake_omega(T):
    T = torch.as_tensor(np.asarray(T, float))
    return torch.stack([T, T ** 2, torch.sqrt(T)], dim=1)


def _worker(rank, world, port, T, n_sp, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        tab = sharding.sharded_temperature_tables(_fake_omega, T, rank, world)
        sp = sharding.sharded_state_points(lambda lo, hi: torch.arange(lo, hi, dtype=torch.float64)[:, None] * torch.ones(1, 2, dtype=torch.float64),
                                           n_sp, rank, world)
        q.put((rank, tab.numpy(), sp.numpy()))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_sharded_tables_match_unsharded(world):
    T = 30.0 * np.geomspace(0.3, 100.0, 37)          # all five lattice regimes, size not divisible by world
    n_sp = 101
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, T, n_sp, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    want = _fake_omega(T).numpy()
    for rank, tab, sp in res:
        assert np.array_equal(tab, want)
        assert np.array_equal(sp[:, 0], np.arange(n_sp)) and sp.shape == (n_sp, 2)


def test_partition_is_balanced_and_complete():
    T = 30.0 * np.geomspace(0.3, 100.0, 4096)
    for world in (1, 2, 4, 8):
        parts = sharding.partition_temperatures(T, world)
        allidx = np.sort(np.concatenate(parts))
        assert np.array_equal(allidx, np.arange(T.size))
        cost = sharding.REGIME_COST[sharding.regime_index(T)]
        loads = np.array([cost[p].sum() for p in parts])
        assert loads.max() / loads.mean() < 1.02


def test_block_partition():
    for n, world in ((10, 3), (65536, 8), (5, 8)):
        s = sharding.partition_blocks(n, world)
        assert s[0] == 0 and s[-1] == n and np.all(np.diff(s) >= 0) and np.diff(s).max() - np.diff(s).min() <= 1
