"""Multi-GPU loop around the collision step (supplementary; NOT the bench.py contract):
one process per GPU, the grid cut into slabs of rows, and per time step

    calculate_displacement (both dimensions)      csrc/transport.cu
    position += displacement; floor; origin; wrap  Storage operators (csrc/storage_ops.cu)
    migrate                                        csrc/migration.cu + NCCL all-to-all-v
    sort by cell + fused collision step            csrc/index.cu, csrc/fused.cu

i.e. the Sc-rain-like flow of BASELINE configs[3] with a synthetic, divergence-free-ish Courant
field and the Golovin kernel (the part of that flow this repo covers).  Checks the global
invariants every step (super-droplets only leave through coalescence, total water mass is
conserved, every rank holds only droplets of its slab) and prints one JSON line on rank 0.

    python -m torch.distributed.run --nproc-per-node N profiles/tools/transport_loop.py \
        [--rows-per-gpu 256] [--cols 512] [--sd-per-cell 128] [--steps 20]
"""

import argparse
import json
import os
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from pysdm_b200 import methods as m  # noqa: E402
from pysdm_b200.collision import DeviceCollision, Golovin  # noqa: E402
from pysdm_b200.migration import SlabMigrator  # noqa: E402
from pysdm_b200.sharding import cell_range, reduce_products  # noqa: E402
from pysdm_b200.storage import Storage, StorageSignature  # noqa: E402


def wrap(tensor, dtype):
    return Storage(StorageSignature(tensor, tuple(tensor.shape), dtype))


def main():  # pylint: disable=too-many-locals,too-many-statements
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows-per-gpu", type=int, default=256)
    ap.add_argument("--cols", type=int, default=512)
    ap.add_argument("--sd-per-cell", type=int, default=128)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--check", action="store_true", help="verify the invariants after every step")
    a = ap.parse_args()
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", 0)))
    sys.stdout.flush()
    out_fd = os.dup(1)
    os.dup2(2, 1)  # NCCL's banner goes to stderr
    dist.init_process_group("nccl", device_id=torch.device("cuda", torch.cuda.current_device()))
    dev = torch.device("cuda")
    grid = (a.rows_per_gpu * world, a.cols)
    rows = cell_range(grid[0], rank, world)
    n_rows = rows[1] - rows[0]
    n_cell = n_rows * grid[1]
    n0 = n_cell * a.sd_per_cell
    cap = int(n0 * 1.25) + 1024
    rng = np.random.default_rng(100 + rank)

    def padded(values, fill, dtype):
        out = np.full(values.shape[:-1] + (cap,), fill, dtype=dtype)
        out[..., :n0] = values
        return out

    cell = np.arange(n0) // a.sd_per_cell
    origin = np.stack([rows[0] + cell // grid[1], cell % grid[1]]).astype(np.int64)
    x0 = 4 / 3 * np.pi * 30.531e-6 ** 3
    sim = DeviceCollision(
        multiplicity=padded(np.full(n0, 2 ** 20, dtype=np.int64), 0, np.int64),
        water_mass=padded(1000.0 * rng.exponential(x0, n0), 0.0, np.float64),
        cell_id=padded(cell.astype(np.int64), 0, np.int64), n_cell=n_cell, dt=1.0, dv=1.0,
        kernel=Golovin(1.5e3), seed=7 + rank, fused=True)
    assert sim.length == n0  # the empty slots (multiplicity 0) are outside idx[:length]
    cell_origin = torch.from_numpy(padded(origin, 0, np.int64)).to(dev)
    position = torch.from_numpy(padded(rng.uniform(0, 1, (2, n0)), 0.5, np.float64)).to(dev)
    displacement = torch.zeros((2, cap), dtype=torch.float64, device=dev)
    floor_tmp = Storage.from_ndarray(np.zeros((2, cap), dtype=np.int64))
    grid_storage = Storage.from_ndarray(np.asarray(grid, dtype=np.int64))
    # Courant field on the Arakawa-C grid of the GLOBAL domain (every rank holds a copy; it is
    # 2 x 8 B per cell): a smooth rotating flow, up to 0.4 cells per step
    ii, jj = np.meshgrid(np.arange(grid[0] + 1), np.arange(grid[1]), indexing="ij")
    courant_x = 0.4 * np.sin(2 * np.pi * (jj + 0.5) / grid[1]) * np.ones_like(ii, dtype=float)
    ii, jj = np.meshgrid(np.arange(grid[0]), np.arange(grid[1] + 1), indexing="ij")
    courant_z = -0.4 * np.sin(2 * np.pi * (ii + 0.5) / grid[0]) * np.ones_like(jj, dtype=float)
    courant = [torch.from_numpy(np.ascontiguousarray(c)).to(dev) for c in (courant_x, courant_z)]
    mig = SlabMigrator(grid=grid, rank=rank, world=world, multiplicity=sim.multiplicity,
                       attributes=sim.attributes, cell_origin=cell_origin,
                       position_in_cell=position, cell_id=sim.cell_id, idx=sim.idx,
                       length=sim.length)

    def totals():
        idx = sim.idx[:sim.length]
        n = sim.multiplicity[idx]
        t = torch.stack([n.sum().double(), (n.double() * sim.attributes[0][idx]).sum(),
                         torch.tensor(float(sim.length), device=dev, dtype=torch.float64)])
        reduce_products(sums=t)  # NCCL all-reduce of the domain-wide products
        return t.cpu().numpy()

    def step():
        for dim in range(2):  # Particulator.calculate_displacement, particulator.py:428-440
            m.calculate_displacement(dim, displacement, courant[dim], cell_origin, position, 1)
        pos, org = wrap(position, Storage.FLOAT), wrap(cell_origin, Storage.INT)
        pos += wrap(displacement, Storage.FLOAT)      # Displacement.update_position
        floor_tmp.floor(pos)                          # Displacement.update_cell_origin
        org += floor_tmp
        pos -= floor_tmp
        org %= grid_storage                           # Displacement.boundary_condition
        mig.idx, mig.length = sim.idx, sim.length
        sim.length = sim.valid_n_sd = mig.migrate()   # also rewrites the slab-local cell ids
        sim.sorted = False
        sim()                                         # sort by cell + fused collision step

    t0 = totals()
    for k in range(a.warmup):
        step()
        if a.check:
            idx = sim.idx[:sim.length]
            t = totals()
            print(f"[check] rank {rank} step {k}: length {sim.length} unique {int(torch.unique(idx).numel())} "
                  f"zero-mult {int((sim.multiplicity[idx] == 0).sum())} sent {mig.n_sent} "
                  f"received {mig.n_received} free {mig.free.numel()} global count {t[2]:.0f} "
                  f"healthy {int(sim.healthy.item())} status {int(sim.status.item())}",
                  file=sys.stderr, flush=True)
    torch.cuda.synchronize()
    dist.barrier()
    sent0, begin = mig.n_sent, time.perf_counter()
    for _ in range(a.steps):
        step()
    torch.cuda.synchronize()
    dist.barrier()
    elapsed = time.perf_counter() - begin
    t1 = totals()
    # invariants: coalescence only lowers the droplet number, water mass is conserved,
    # super-droplets are neither lost nor duplicated by the exchange (Golovin with 2^20-fold
    # multiplicities never empties a super-droplet), everybody sits in its own slab
    assert t1[0] < t0[0] and abs(t1[1] - t0[1]) <= 1e-11 * t0[1], (t0, t1)
    assert t1[2] == t0[2], (t0, t1)
    own = cell_origin[0][sim.idx[:sim.length]]
    assert bool(((own >= rows[0]) & (own < rows[1])).all())
    assert sim.n_fused_steps == a.steps + a.warmup
    stats = torch.tensor([float(mig.n_sent - sent0), elapsed], device=dev, dtype=torch.float64)
    dist.all_reduce(stats, op=dist.ReduceOp.SUM)
    if rank == 0:
        n_total = int(t1[2])
        sys.stdout.flush()
        os.dup2(out_fd, 1)
        print(json.dumps({
            "loop": "displacement + migration (NCCL) + sort by cell + fused collision step",
            "n_gpus": world, "grid": list(grid), "super_droplets": n_total, "steps": a.steps,
            "ms_per_step": elapsed / a.steps * 1e3,
            "sd_steps_per_s": n_total * a.steps / elapsed,
            "migrating_fraction_per_step": float(stats[0]) / a.steps / n_total,
            "droplet_number_ratio": float(t1[0] / t0[0]),
            "mass_conservation_rel_err": float(abs(t1[1] - t0[1]) / t0[1])}), flush=True)
        os.dup2(2, 1)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
