"""CPU, world_size 2 over gloo: the host-side shard logic of the multi-GPU path.  The
collision step itself has no collective (cells are independent), so what is checked is
(a) the partition is exact and balanced, (b) stepping the shards independently with the
CPU oracle and gathering the per-cell counters equals stepping the whole population at
once -- i.e. sharding by cell changes nothing as long as the per-cell random streams are
the same, (c) the per-cell all-gather."""

import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from pysdm_b200 import sharding


@pytest.mark.parametrize("n_cell, world", ((10, 2), (7, 3), (5, 8), (131072, 8), (1, 1)))
def test_cell_range_partition(n_cell, world):
    ranges = [sharding.cell_range(n_cell, r, world) for r in range(world)]
    assert ranges[0][0] == 0 and ranges[-1][1] == n_cell
    for (_, e), (b, _) in zip(ranges[:-1], ranges[1:]):
        assert e == b
    sizes = [e - b for b, e in ranges]
    assert max(sizes) - min(sizes) <= 1
    owners = sharding.owner_of_cell(np.arange(n_cell), n_cell, world)
    for r, (b, e) in enumerate(ranges):
        assert (owners[b:e] == r).all()


def test_shard_population_is_a_partition():
    rng = np.random.default_rng(0)
    n_sd, n_cell, world = 1000, 13, 4
    cid = rng.integers(0, n_cell, n_sd)
    mult = rng.integers(1, 9, n_sd)
    attrs = rng.random((2, n_sd))
    seen = []
    for r in range(world):
        m, a, lc, nlc, gi = sharding.shard_population(mult, attrs, cid, n_cell, r, world)
        b, e = sharding.cell_range(n_cell, r, world)
        assert nlc == e - b and lc.min() >= 0 and lc.max() < nlc
        np.testing.assert_array_equal(m, mult[gi])
        np.testing.assert_array_equal(a, attrs[:, gi])
        np.testing.assert_array_equal(lc + b, cid[gi])
        seen.append(gi)
    np.testing.assert_array_equal(np.sort(np.concatenate(seen)), np.arange(n_sd))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, n_cell, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        b, e = sharding.cell_range(n_cell, rank, world)
        local = torch.arange(b, e, dtype=torch.int64) * 10 + rank
        full = sharding.gather_per_cell(local, n_cell)
        expect = torch.cat([torch.arange(*sharding.cell_range(n_cell, r, world)) * 10 + r
                            for r in range(world)])
        assert torch.equal(full, expect)
        # timing protocol of bench.py: max over ranks of the elapsed time
        t = torch.tensor([float(rank + 1)], dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        assert t.item() == world
        # products of the whole domain: rate totals, dt_min / healthy, substep count
        sums, minima, maxima = sharding.reduce_products(
            sums=torch.tensor([float(e - b), 2.5 * (rank + 1)], dtype=torch.float64),
            minima=torch.tensor([10.0 / (rank + 1), float(rank != 1)], dtype=torch.float64),
            maxima=torch.tensor([3 + rank], dtype=torch.int64))
        assert sums.tolist() == [float(n_cell), 2.5 * world * (world + 1) / 2]
        assert minima.tolist() == [10.0 / world, 0.0] and maxima.tolist() == [2 + world]
        assert sharding.reduce_products() == (None, None, None)
        with open(os.path.join(out_dir, f"ok{rank}"), "w", encoding="utf-8") as f:
            f.write("ok")
    finally:
        dist.destroy_process_group()


def test_gather_per_cell_gloo_world2(tmp_path):
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), 7, str(tmp_path)), nprocs=world, join=True)
    assert all((tmp_path / f"ok{r}").exists() for r in range(world))


def test_sharded_steps_equal_global_step():
    """cells are independent: the oracle stepped per shard (with the uniforms of its own
    positions) gives the per-cell counters and multiplicities of the unsharded run"""
    from oracle.sdm_step import Golovin, OracleCollision  # pylint: disable=import-outside-toplevel

    rng = np.random.default_rng(3)
    n_cell, per_cell, world = 12, 16, 2
    n_sd = n_cell * per_cell
    cid = np.repeat(np.arange(n_cell), per_cell)  # cell-contiguous, equal sizes
    mult = rng.integers(1, 5, n_sd).astype(np.int64)
    mass = rng.exponential(1000 * 4 / 3 * np.pi * 30.531e-6 ** 3, n_sd)

    class Replay:  # uniforms indexed by global position / pair slot
        def __init__(self, u, r, lo_pos):
            self.u, self.r, self.lo, self.k = u, r, lo_pos, 0

        def __call__(self, array):
            src = self.u if self.k % 2 == 0 else self.r
            lo = self.lo if self.k % 2 == 0 else self.lo // 2
            array[:] = src[lo:lo + len(array)]
            self.k += 1

    u, r = rng.random(n_sd), rng.random(n_sd // 2)
    kw = dict(dt=1.0, dv=2e-8, kernel=Golovin(1.5e3), seed=0)
    whole = OracleCollision(multiplicity=mult, water_mass=mass, cell_id=cid, n_cell=n_cell,
                            rnd=Replay(u, r, 0), **kw)
    whole()
    for rank in range(world):
        m, a, lc, nlc, gi = sharding.shard_population(mult, mass, cid, n_cell, rank, world)
        part = OracleCollision(multiplicity=m, water_mass=a[0], cell_id=lc, n_cell=nlc,
                               rnd=Replay(u, r, int(gi[0])), **kw)
        part()
        b, e = sharding.cell_range(n_cell, rank, world)
        np.testing.assert_array_equal(part.coalescence_rate, whole.coalescence_rate[b:e])
        np.testing.assert_array_equal(part.multiplicity, whole.multiplicity[gi])
        np.testing.assert_array_equal(part.attributes[0], whole.attributes[0][gi])
    assert whole.coalescence_rate.sum() > 0


def _exchange_worker(rank, world, port, out_dir):
    """pysdm_b200.migration.exchange (the all-to-all-v of migrating super-droplet records) on
    CPU tensors over gloo: rank r sends r + d + 1 rows to rank d (none to itself), every word
    tagged with (source, destination, row)"""
    from pysdm_b200.migration import exchange  # pylint: disable=import-outside-toplevel

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        n_words = 5

        def rows(src, dst):
            n = 0 if src == dst else src + dst + 1
            base = (src * 100 + dst * 10) * 1000
            return (base + torch.arange(n, dtype=torch.int64)[:, None] * n_words
                    + torch.arange(n_words, dtype=torch.int64)[None, :])

        send = torch.cat([rows(rank, d) for d in range(world)])
        counts = [0 if d == rank else rank + d + 1 for d in range(world)]
        recv, recv_counts = exchange(send, counts)
        assert recv_counts == [0 if s == rank else s + rank + 1 for s in range(world)]
        assert torch.equal(recv, torch.cat([rows(s, rank) for s in range(world)]))
        # a rank with nothing to send or receive still takes part
        recv, recv_counts = exchange(torch.empty((0, n_words), dtype=torch.int64), [0] * world)
        assert recv.shape == (0, n_words) and recv_counts == [0] * world
        with open(os.path.join(out_dir, f"ok{rank}"), "w", encoding="utf-8") as f:
            f.write("ok")
    finally:
        dist.destroy_process_group()


def test_migration_exchange_gloo_world2(tmp_path):
    world = 2
    mp.spawn(_exchange_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    assert all((tmp_path / f"ok{r}").exists() for r in range(world))
