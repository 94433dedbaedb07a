"""GPU: migration of super-droplets between slabs (pysdm_b200.migration, csrc/migration.cu).
One GPU: the ranks of a 3-slab world live in one process and the exchange is a loopback that
routes the packed records exactly as all_to_all_single does (the NCCL call itself is covered by
tests/test_sharding.py over gloo and, on a box with >= 2 GPUs, by tests/mp_migration_nccl.py
under torchrun)."""

import os
import subprocess
import sys

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def test_migration_loopback_three_slabs():
    from .migration_common import Shard, global_population, moved, sort_and_count  # pylint: disable=import-outside-toplevel

    world = 3
    pop = global_population()
    shards = [Shard(pop, r, world) for r in range(world)]
    for s in shards:
        s.displace()
    packed = [s.migrator.pack() for s in shards]
    for dst, s in enumerate(shards):  # what all_to_all_single delivers to rank `dst`
        parts = []
        for src in range(world):
            records, counts = packed[src]
            begin = sum(counts[:dst])
            parts.append(records[begin:begin + counts[dst]])
        s.migrator.unpack(torch.cat(parts))
    origin, _ = moved(pop)
    total = sum(s.check(pop) for s in shards)
    assert total == pop["multiplicity"].size
    assert sum(s.migrator.n_sent for s in shards) == sum(s.migrator.n_received for s in shards) > 0
    sizes = np.concatenate([sort_and_count(s) for s in shards])
    want = np.bincount(origin[0] * 6 + origin[1], minlength=66)
    np.testing.assert_array_equal(sizes, want)


@pytest.mark.parametrize("world", (1, 2, 3, 6))
def test_migration_neighbour_path_loopback(world):
    """the device-driven neighbour exchange (no host round trip before the data is in place:
    64-bit scan of the three classes, fixed-capacity buffers with the count in their header,
    device free stack) on a loopback world; with 6 thin slabs some droplets move farther than a
    neighbour: they must be left in place, reported, and moved by the general path afterwards"""
    from .migration_common import Shard, global_population, moved, sort_and_count  # pylint: disable=import-outside-toplevel

    pop = global_population()
    shards = [Shard(pop, r, world) for r in range(world)]
    for s in shards:
        s.displace()
        s.migrator.split()
    for dst, s in enumerate(shards):  # what the two send/recv pairs deliver to rank `dst`
        if world == 1:
            s.migrator.recv.zero_()
        else:
            s.migrator.recv[0].copy_(shards[(dst - 1) % world].migrator.send[0])
            s.migrator.recv[1].copy_(shards[(dst + 1) % world].migrator.send[1])
    far = [s.migrator.merge() for s in shards]
    assert any(far) == (world == 6)
    if any(far):  # every rank enters the general path together
        packed = [s.migrator.pack() for s in shards]
        for dst, s in enumerate(shards):
            parts = []
            for src in range(world):
                records, counts = packed[src]
                begin = sum(counts[:dst])
                parts.append(records[begin:begin + counts[dst]])
            s.migrator.unpack(torch.cat(parts))
    origin, _ = moved(pop)
    total = sum(s.check(pop) for s in shards)
    assert total == pop["multiplicity"].size
    sent = sum(s.migrator.n_sent for s in shards)
    assert sent == sum(s.migrator.n_received for s in shards) and (sent > 0 or world == 1)
    sizes = np.concatenate([sort_and_count(s) for s in shards])
    want = np.bincount(origin[0] * 6 + origin[1], minlength=66)
    np.testing.assert_array_equal(sizes, want)


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs (NCCL all-to-all)")
def test_migration_nccl_two_gpus():
    script = os.path.join(os.path.dirname(__file__), "mp_migration_nccl.py")
    out = subprocess.run(
        [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
         "--master-addr", "127.0.0.1", "--master-port", "29541", script],
        capture_output=True, text=True, timeout=600, check=False)
    assert out.returncode == 0, out.stdout[-2000:] + out.stderr[-4000:]
    assert "migration ok rank 0" in out.stdout and "migration ok rank 1" in out.stdout


def test_migration_neighbours_only_world_1():
    """`neighbours_only=True` (no agreement between the ranks on far movers): migrate() is the
    neighbour exchange alone; a droplet beyond a neighbour slab raises instead of being moved"""
    from .migration_common import Shard, global_population  # pylint: disable=import-outside-toplevel

    pop = global_population()
    shard = Shard(pop, 0, 1)
    shard.migrator.neighbours_only = True
    shard.displace()
    shard.migrator.migrate()
    assert shard.check(pop) == pop["multiplicity"].size and shard.migrator.n_general == 0
    # a far mover, as merge() reports it (status bit 1 of the device counters)
    shard.migrator.merge = lambda all_reduce_far=False: True
    with pytest.raises(RuntimeError, match="neighbours_only"):
        shard.migrator.migrate()
