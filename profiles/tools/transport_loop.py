"""Multi-GPU loop around the collision step: BASELINE configs[3] (Sc-rain, Arabas et al. 2015)
in kind, sharded over the ranks.  One process per GPU, the 2-D grid cut into slabs of rows
(64 rows x 512 columns x 128 super-droplets = 2^22 per GPU; 8 GPUs = the 512 x 512 grid), and per
time step

    calculate_displacement (both dimensions)      csrc/transport.cu
    position += displacement; floor; origin; wrap  Storage operators (csrc/storage_ops.cu)
    migrate                                        csrc/migration.cu + NCCL all-to-all-v
    derived attributes of the arrivals             csrc/fused.cu (one pass)
    sort by cell + adaptive collision sub-steps    csrc/index.cu, csrc/fused*.cu

with `Coalescence(Geometric(1), adaptive=True, optimized_random=True)`
(Morrison_and_Grabowski_2007/common.py:30-35), A = 3 extensive attributes (water mass, "dry
volume", "kappa times dry volume": environments/kinematic_2d.py:63-64,84), dt = 5 s, cells of
(1500 m / 512)^2 x 1 m, a synthetic rotating Courant field (the flow solver and condensation
are out of scope).  Checks the global invariants (no super-droplet lost or duplicated by the
exchange, total water mass conserved, every rank holds only droplets of its slab).

    python -m torch.distributed.run --nproc-per-node N profiles/tools/transport_loop.py \
        [--rows-per-gpu 64] [--cols 512] [--sd-per-cell 128] [--steps 20] [--kernel geometric]

`bench_entry` is what `bench.py --gpus N` (N > 1) calls for its `transport` key.
"""

import argparse
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from pysdm_b200 import methods as m  # noqa: E402
from pysdm_b200.collision import DeviceCollision, Geometric, Golovin  # noqa: E402
from pysdm_b200.migration import SlabMigrator  # noqa: E402
from pysdm_b200.sharding import cell_range, reduce_products  # noqa: E402
from pysdm_b200.storage import Storage, StorageSignature  # noqa: E402


def wrap(tensor, dtype):
    return Storage(StorageSignature(tensor, tuple(tensor.shape), dtype))


class TransportLoop:  # pylint: disable=too-many-instance-attributes
    def __init__(self, rank, world, *, rows_per_gpu=64, cols=512, sd_per_cell=128,
                 kernel="geometric", fused_displacement=True):
        self.rank, self.world = rank, world
        self.fused_displacement = fused_displacement
        dev = self.dev = torch.device("cuda")
        grid = self.grid = (rows_per_gpu * world, cols)
        self.rows = rows = cell_range(grid[0], rank, world)
        n_cell = (rows[1] - rows[0]) * grid[1]
        n0 = self.n0 = n_cell * sd_per_cell
        cap = int(n0 * 1.25) + 1024
        # every slab starts from the same droplet population (the collision random streams differ):
        # the number of adaptive sub-steps a slab needs is set by the tail of its size
        # distribution, and a weak-scaling run should give every GPU the same work at every N
        rng = np.random.default_rng(100)

        def padded(values, fill, dtype):
            out = np.full(values.shape[:-1] + (cap,), fill, dtype=dtype)
            out[..., :n0] = values
            return out

        cell = np.arange(n0) // sd_per_cell
        origin = np.stack([rows[0] + cell // grid[1], cell % grid[1]]).astype(np.int64)
        if kernel == "geometric":  # configs[3] in kind
            dv, dt = (1500 / 512) ** 2, 5.0
            volume = 4 / 3 * np.pi * (10e-6 * np.exp(0.47 * rng.standard_normal(n0))) ** 3
            mult = np.full(n0, int(100e6 * dv / sd_per_cell), dtype=np.int64)
            extra = [padded(volume * 1e-3, 0.0, np.float64), padded(volume * 0.6e-3, 0.0, np.float64)]
            kw = dict(kernel=Geometric(1.0), adaptive=True, optimized_random=True,
                      dt_coal_range=(0.1, dt), extra_attributes=extra)
            self.what = ("Geometric(1) + adaptive + optimized_random, A=3, dt=5 s "
                         "(BASELINE configs[3] in kind)")
        else:  # the round-1 loop: Golovin, fixed time step, A = 1
            dv, dt = 1.0, 1.0
            volume = rng.exponential(4 / 3 * np.pi * 30.531e-6 ** 3, n0)
            mult = np.full(n0, 2 ** 20, dtype=np.int64)
            kw = dict(kernel=Golovin(1.5e3))
            self.what = "Golovin, fixed time step, A=1"
        self.sim = sim = DeviceCollision(
            multiplicity=padded(mult, 0, np.int64), water_mass=padded(1000.0 * volume, 0.0, np.float64),
            cell_id=padded(cell.astype(np.int64), 0, np.int64), n_cell=n_cell, dt=dt, dv=dv,
            seed=7 + rank, fused=True, **kw)
        assert sim.length == n0  # the empty slots (multiplicity 0) are outside idx[:length]
        self.cell_origin = torch.from_numpy(padded(origin, 0, np.int64)).to(dev)
        self.position = torch.from_numpy(padded(rng.uniform(0, 1, (2, n0)), 0.5, np.float64)).to(dev)
        self.displacement = torch.zeros((2, cap), dtype=torch.float64, device=dev)
        self.floor_tmp = Storage.from_ndarray(np.zeros((2, cap), dtype=np.int64))
        self.grid_storage = Storage.from_ndarray(np.asarray(grid, dtype=np.int64))
        # Courant field on the Arakawa-C grid of the GLOBAL domain (every rank holds a copy; it
        # is 2 x 8 B per cell): a smooth rotating flow, up to 0.4 cells per step
        ii, jj = np.meshgrid(np.arange(grid[0] + 1), np.arange(grid[1]), indexing="ij")
        courant_x = 0.4 * np.sin(2 * np.pi * (jj + 0.5) / grid[1]) * np.ones_like(ii, dtype=float)
        ii, jj = np.meshgrid(np.arange(grid[0]), np.arange(grid[1] + 1), indexing="ij")
        courant_z = -0.4 * np.sin(2 * np.pi * (ii + 0.5) / grid[0]) * np.ones_like(jj, dtype=float)
        self.courant = [torch.from_numpy(np.ascontiguousarray(c)).to(dev)
                        for c in (courant_x, courant_z)]
        self.mig = SlabMigrator(grid=grid, rank=rank, world=world, multiplicity=sim.multiplicity,
                                attributes=sim.attributes, cell_origin=self.cell_origin,
                                position_in_cell=self.position, cell_id=sim.cell_id, idx=sim.idx,
                                length=sim.length,
                                # |Courant| <= 0.4 against slabs of 64 rows: ring neighbours only
                                neighbours_only=True)
        self.ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
        self.phase_ms = np.zeros(3)
        self.migration_pieces = np.zeros(3)  # split+pack, exchange, merge (+ read-back)

    def totals(self):
        """domain-wide: droplet number, water mass, super-droplet count (NCCL all-reduce)"""
        sim = self.sim
        idx = sim.idx[:sim.length]
        n = sim.multiplicity[idx]
        t = torch.stack([n.sum().double(), (n.double() * sim.attributes[0][idx]).sum(),
                         torch.tensor(float(sim.length), device=self.dev, dtype=torch.float64)])
        reduce_products(sums=t)
        return t.cpu().numpy()

    def step(self, timed=False):
        sim, mig = self.sim, self.mig
        if timed:
            self.ev[0].record()
        if self.fused_displacement:
            # Displacement.__call__ (dynamics/displacement.py:103-153) in one launch
            m.displacement_step(self.displacement, self.courant, self.cell_origin, self.position,
                                self.grid, 1)
        else:  # the same as the reference issues it: seven passes
            for dim in range(2):  # Particulator.calculate_displacement, particulator.py:428-440
                m.calculate_displacement(dim, self.displacement, self.courant[dim],
                                         self.cell_origin, self.position, 1)
            pos, org = wrap(self.position, Storage.FLOAT), wrap(self.cell_origin, Storage.INT)
            pos += wrap(self.displacement, Storage.FLOAT)  # Displacement.update_position
            self.floor_tmp.floor(pos)                      # Displacement.update_cell_origin
            org += self.floor_tmp
            pos -= self.floor_tmp
            org %= self.grid_storage                       # Displacement.boundary_condition
        if timed:
            self.ev[1].record()
        mig.idx, mig.length = sim.idx, sim.length
        if timed:  # the pieces of migrate(), each synchronised (outside the timed region)
            import time  # pylint: disable=import-outside-toplevel

            t = [time.perf_counter()]
            mig.split()
            torch.cuda.synchronize(); t.append(time.perf_counter())
            mig.neighbour_exchange_fn(mig.send, mig.recv, mig.rank, mig.world)
            torch.cuda.synchronize(); t.append(time.perf_counter())
            if mig.merge(all_reduce_far=mig.world > 1 and not mig.neighbours_only):
                mig.migrate_general()
            torch.cuda.synchronize(); t.append(time.perf_counter())
            self.migration_pieces += np.diff(t) * 1e3
            sim.length = sim.valid_n_sd = mig.length
        else:
            sim.length = sim.valid_n_sd = mig.migrate()  # also rewrites the slab-local cell ids
        sim.sorted = False
        if sim.dependents is not None:
            sim.refresh_dependents()                   # radius / fall velocity of the arrivals
        if timed:
            self.ev[2].record()
        sim()                                          # sort by cell + collision (sub-)steps
        if timed:
            self.ev[3].record()
            torch.cuda.current_stream().synchronize()
            self.phase_ms += [self.ev[i].elapsed_time(self.ev[i + 1]) for i in range(3)]

    def check_slab(self):
        own = self.cell_origin[0][self.sim.idx[:self.sim.length]]
        unique = int(torch.unique(self.sim.idx[:self.sim.length]).numel())
        return bool(((own >= self.rows[0]) & (own < self.rows[1])).all()), unique == self.sim.length


def bench_entry(rank, world, steps, barrier, gather, **loop_kw):
    """times `steps` steps of the loop (CUDA events, max over ranks) and checks its invariants;
    returns the dict of bench.py's `transport` key (identical on every rank)"""
    loop = TransportLoop(rank, world, **loop_kw)
    sim, mig = loop.sim, loop.mig
    t0 = loop.totals()
    for _ in range(3):
        loop.step()
    substeps0 = sim.n_fused_steps
    sent0 = mig.n_sent
    beg, end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    beg.record()
    for _ in range(steps):
        loop.step()
    end.record()
    barrier()
    per_rank = gather(beg.elapsed_time(end))
    ms = max(per_rank)
    t1 = loop.totals()
    in_slab, no_duplicates = loop.check_slab()
    for _ in range(3):  # phase breakdown (synchronises every step: outside the timed region)
        loop.step(timed=True)
    phases = [max(gather(x)) / 3 for x in loop.phase_ms]
    flags = [min(gather(float(in_slab))) > 0, min(gather(float(no_duplicates))) > 0]
    sent = sum(gather(float(mig.n_sent - sent0)))
    substeps = gather(float(sim.n_fused_steps - substeps0))
    n_total = int(t1[2])
    mass_err = float(abs(t1[1] - t0[1]) / t0[1])
    return {
        "loop": "displacement (one launch) + migration (device-driven neighbour exchange, NCCL "
                "send/recv) + derived attributes + sort by cell + collision sub-steps",
        "collision": loop.what, "grid": list(loop.grid), "super_droplets": n_total,
        "super_droplets_per_gpu": loop.n0, "steps": steps, "ms_per_step": ms / steps,
        "sd_steps_per_s": n_total * steps / (ms * 1e-3), "per_rank_ms": per_rank,
        "collision_substeps_per_step": float(np.mean(substeps)) / steps,
        "phase_ms": {"displacement": phases[0], "migration": phases[1], "collision": phases[2]},
        "migration_ms": dict(zip(("split_pack", "exchange", "merge_readback"),
                                 [max(gather(x)) / 3 for x in loop.migration_pieces])),
        "migration_capacity_records": mig.cap, "migration_general_path_runs": mig.n_general,
        "migrating_fraction_per_step": sent / steps / max(n_total, 1),
        "invariants": {"every_rank_holds_only_its_slab": flags[0],
                       "no_super_droplet_duplicated": flags[1],
                       "super_droplets_before_after": [int(t0[2]), n_total],
                       "droplet_number_ratio": float(t1[0] / t0[0]),
                       "mass_conservation_rel_err": mass_err,
                       "ok": bool(flags[0] and flags[1] and mass_err <= 1e-11
                                  and t1[0] <= t0[0] and t1[2] <= t0[2])},
        "limiting_collective": "two fixed-size send/recv pairs with the ring neighbours (NCCL; "
                               "counts travel in the buffers' headers, no host round trip): "
                               "migration_ms.exchange"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows-per-gpu", type=int, default=64)
    ap.add_argument("--cols", type=int, default=512)
    ap.add_argument("--sd-per-cell", type=int, default=128)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--kernel", default="geometric", choices=("geometric", "golovin"))
    a = ap.parse_args()
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", 0)))
    sys.stdout.flush()
    out_fd = os.dup(1)
    os.dup2(2, 1)  # NCCL's banner goes to stderr
    dist.init_process_group("nccl", device_id=torch.device("cuda", torch.cuda.current_device()))

    def barrier():
        dist.barrier()
        torch.cuda.synchronize()

    def gather(value):
        t = torch.zeros(world, dtype=torch.float64, device="cuda")
        t[rank] = value
        dist.all_reduce(t)
        return [float(x) for x in t.tolist()]

    result = bench_entry(rank, world, a.steps, barrier, gather, rows_per_gpu=a.rows_per_gpu,
                         cols=a.cols, sd_per_cell=a.sd_per_cell, kernel=a.kernel)
    result["n_gpus"] = world
    if rank == 0:
        sys.stdout.flush()
        os.dup2(out_fd, 1)
        print(json.dumps(result), flush=True)
        os.dup2(2, 1)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
