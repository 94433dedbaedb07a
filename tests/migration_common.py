"""shared by tests/test_gpu_migration.py (loopback on one GPU) and tests/mp_migration_nccl.py
(NCCL, one process per GPU): a seeded global population on a 2-D grid, its slab shards on the
device, and the check that after displacement + migration every rank holds exactly the
super-droplets of its slab, bit for bit."""

import numpy as np
import torch

from pysdm_b200 import methods as m
from pysdm_b200.migration import SlabMigrator
from pysdm_b200.sharding import cell_range

GRID = (11, 6)
N_ATTR = 3


def global_population(seed=5, n_sd=4000):
    rng = np.random.default_rng(seed)
    origin = np.stack([rng.integers(0, n, n_sd) for n in GRID]).astype(np.int64)
    return {
        "multiplicity": rng.integers(1, 10 ** 6, n_sd).astype(np.int64),
        "attributes": rng.exponential(1e-12, (N_ATTR, n_sd)),
        "cell_origin": origin,
        "position_in_cell": rng.uniform(0, 1, (2, n_sd)),
        "displacement": rng.uniform(-1.7, 1.7, (2, n_sd)),  # up to two cells per step
    }


def moved(pop):
    """host restatement of Displacement.update_position/update_cell_origin/boundary_condition
    (PySDM/dynamics/displacement.py:138-153; pinned by tests/test_transport_golden.py)"""
    pos = pop["position_in_cell"] + pop["displacement"]
    floor = np.floor(pos).astype(np.int64)
    origin = (pop["cell_origin"] + floor) % np.asarray(GRID)[:, None]
    return origin, pos - floor


class Shard:  # pylint: disable=too-many-instance-attributes
    """the super-droplets of rank's slab on the device, with spare capacity"""

    def __init__(self, pop, rank, world, exchange_fn=None, device="cuda"):
        b, e = cell_range(GRID[0], rank, world)
        mine = np.flatnonzero((pop["cell_origin"][0] >= b) & (pop["cell_origin"][0] < e))
        n, cap = len(mine), 2 * len(mine) + 64

        def dev(a, fill):
            out = np.full(a.shape[:-1] + (cap,), fill, dtype=a.dtype)
            out[..., :n] = a[..., mine]
            return torch.from_numpy(out).to(device)

        self.rank, self.world, self.rows = rank, world, (b, e)
        self.multiplicity = dev(pop["multiplicity"], 0)
        self.attributes = dev(pop["attributes"], np.nan)
        self.cell_origin = dev(pop["cell_origin"], -1)
        self.position_in_cell = dev(pop["position_in_cell"], np.nan)
        self.displacement = dev(pop["displacement"], 0.0)
        self.cell_id = torch.full((cap,), -1, dtype=torch.int64, device=device)
        self.idx = torch.arange(cap, dtype=torch.int64, device=device)
        kw = {} if exchange_fn is None else {"exchange_fn": exchange_fn}
        self.migrator = SlabMigrator(
            grid=GRID, rank=rank, world=world, multiplicity=self.multiplicity,
            attributes=self.attributes, cell_origin=self.cell_origin,
            position_in_cell=self.position_in_cell, cell_id=self.cell_id, idx=self.idx, length=n,
            **kw)

    def displace(self):
        """Displacement's Storage operator chain through the backend's own Storage class"""
        from pysdm_b200.storage import Storage, StorageSignature  # pylint: disable=import-outside-toplevel

        def wrap(t, dtype):
            return Storage(StorageSignature(t, tuple(t.shape), dtype))

        pos, org = wrap(self.position_in_cell, Storage.FLOAT), wrap(self.cell_origin, Storage.INT)
        tmp = Storage.from_ndarray(np.zeros(tuple(self.cell_origin.shape), dtype=np.int64))
        self.position_in_cell.nan_to_num_(nan=0.5)  # free slots: keep the arithmetic finite
        pos += wrap(self.displacement, Storage.FLOAT)
        tmp.floor(pos)
        org += tmp
        pos -= tmp
        org %= Storage.from_ndarray(np.asarray(GRID, dtype=np.int64))

    def check(self, pop):
        """idx[:length] = exactly the SDs whose NEW cell is in this slab, records bit-identical"""
        origin, pos = moved(pop)
        b, e = self.rows
        expect = np.flatnonzero((origin[0] >= b) & (origin[0] < e))
        n = self.migrator.length
        assert n == len(expect), (self.rank, n, len(expect))
        idx = self.idx[:n].cpu().numpy()
        assert len(np.unique(idx)) == n
        got = np.concatenate([
            self.multiplicity.cpu().numpy()[None, idx].view(np.int64),
            self.attributes.cpu().numpy()[:, idx].view(np.int64),
            self.cell_origin.cpu().numpy()[:, idx],
            self.position_in_cell.cpu().numpy()[:, idx].view(np.int64)])
        want = np.concatenate([
            pop["multiplicity"][None, expect], pop["attributes"][:, expect].view(np.int64),
            origin[:, expect], pos[:, expect].view(np.int64)])
        # the order of arrivals is by source rank: compare as sets of records
        def canon(a):
            return a[:, np.lexsort(a[::-1])]

        np.testing.assert_array_equal(canon(got), canon(want))
        cid = self.cell_id.cpu().numpy()[idx]
        org = self.cell_origin.cpu().numpy()[:, idx]
        np.testing.assert_array_equal(cid, (org[0] - b) * GRID[1] + org[1])
        assert cid.min() >= 0 and cid.max() < self.migrator.n_local_cells
        free = self.migrator.free.cpu().numpy()
        assert len(np.intersect1d(free, idx)) == 0 and len(free) + n == self.migrator.n_cap
        assert (self.multiplicity.cpu().numpy()[free] == 0).all()
        return n


def sort_and_count(shard):
    """the migrated population goes straight into the collision path: sort by (local) cell"""
    mg = shard.migrator
    n_cell = mg.n_local_cells
    cell_start = torch.zeros(n_cell + 1, dtype=torch.int64, device=shard.idx.device)
    cell_idx = torch.arange(n_cell, dtype=torch.int64, device=shard.idx.device)
    new_idx = torch.empty_like(shard.idx)
    m.counting_sort_by_cell_id(new_idx, shard.idx, shard.cell_id, cell_idx, mg.length, cell_start)
    sizes = np.diff(cell_start.cpu().numpy())
    assert sizes.sum() == mg.length
    return sizes
