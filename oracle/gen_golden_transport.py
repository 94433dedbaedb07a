"""Generates tests/golden/transport.npz by RUNNING THE REFERENCE (PySDM Numba backend,
Formulae(fastmath=False)): method-level input/output vectors for the rows either side of the
collision step: `moments`, `spectrum_moments` (products side) and
`calculate_displacement` (1/2/3-D, both advection schemes), `flag_precipitated`,
`flag_out_of_column` plus the Storage operations `Displacement` strings together
(`+=`, `floor`, `-=`, `%=`; PySDM/dynamics/displacement.py:103-153).

TEST INFRASTRUCTURE (see gen_golden.py).  Usage:

    NUMBA_NUM_THREADS=1 python oracle/gen_golden_transport.py
"""

import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)

import gen_golden as gg  # noqa: E402  (sets up the import path of the reference)
import numpy as np  # noqa: E402
from PySDM import Formulae  # noqa: E402
from PySDM.backends import CPU  # noqa: E402


def gen_transport():
    rng = np.random.default_rng(20261105)
    g = {}

    def put(case, **kw):
        for k, v in kw.items():
            g[f"{case}/{k}"] = np.asarray(v)

    backend = CPU(Formulae(fastmath=False))
    S = backend.Storage
    Index = gg.make_Index(backend)
    IndexedStorage = gg.make_IndexedStorage(backend)

    # ---- moments / spectrum_moments (impl_numba/methods/moments_methods.py:12-182) ------------
    for tag, n_sd, n_cell, length in (("a", 500, 7, 500), ("b", 1200, 1, 1000), ("c", 64, 40, 64)):
        cid, idx0, _ = gg.random_cells(rng, n_sd, n_cell)
        mult = rng.integers(1, 10 ** 6, n_sd).astype(np.int64)
        volume = rng.exponential(4e-15, n_sd)
        mass = 1000.0 * volume
        mass[rng.integers(0, n_sd, n_sd // 9)] *= -1  # ice: signed water mass < 0
        idx = Index.from_ndarray(idx0.copy())
        idx.length = S.INT(length)
        put(f"moments_{tag}", idx=idx0, length=length, cell_id=cid, multiplicity=mult,
            volume=volume, mass=mass, n_cell=n_cell)
        variants = {
            "plain": dict(ranks=(0.0, 1.0, 2.0), rng_=(-np.inf, np.inf), wrank=0, skip=False),
            "range": dict(ranks=(1.0, 1 / 3, 2.5), rng_=(0.0, 3e-12), wrank=0, skip=False),
            "weighted": dict(ranks=(1.0,), rng_=(-np.inf, np.inf), wrank=1, skip=True),
            "norank": dict(ranks=(), rng_=(-1e-12, np.inf), wrank=2, skip=False),
        }
        for name, v in variants.items():
            ranks = np.asarray(v["ranks"], dtype=float)
            m0 = S.from_ndarray(np.full(n_cell, np.nan))
            mk = S.from_ndarray(np.full((len(ranks), n_cell), np.nan))
            backend.moments(
                moment_0=m0, moments=mk, multiplicity=IndexedStorage.from_ndarray(idx, mult),
                attr_data=IndexedStorage.from_ndarray(idx, volume) if len(ranks) else
                S.empty((0,), dtype=float),
                cell_id=IndexedStorage.from_ndarray(idx, cid), idx=idx, length=length,
                ranks=S.from_ndarray(ranks), min_x=v["rng_"][0], max_x=v["rng_"][1],
                x_attr=IndexedStorage.from_ndarray(idx, mass),
                weighting_attribute=IndexedStorage.from_ndarray(idx, np.abs(mass)),
                weighting_rank=v["wrank"], skip_division_by_m0=v["skip"])
            put(f"moments_{tag}_{name}", ranks=ranks, min_x=v["rng_"][0], max_x=v["rng_"][1],
                weighting_rank=v["wrank"], skip=int(v["skip"]), moment_0=m0.data,
                moments=mk.data)
        for name, wrank, rank in (("r1", 0, 1.0), ("w1r0", 1, 0.0), ("r2_5", 0, 2.5)):
            bins = np.concatenate([[0.0], np.sort(rng.exponential(4e-12, 6))])
            m0 = S.from_ndarray(np.full((len(bins) - 1, n_cell), np.nan))
            mk = S.from_ndarray(np.full((len(bins) - 1, n_cell), np.nan))
            backend.spectrum_moments(
                moment_0=m0, moments=mk, multiplicity=IndexedStorage.from_ndarray(idx, mult),
                attr_data=IndexedStorage.from_ndarray(idx, volume),
                cell_id=IndexedStorage.from_ndarray(idx, cid), idx=idx, length=length, rank=rank,
                x_bins=S.from_ndarray(bins), x_attr=IndexedStorage.from_ndarray(idx, mass),
                weighting_attribute=IndexedStorage.from_ndarray(idx, np.abs(mass)),
                weighting_rank=wrank)
            put(f"spectrum_{tag}_{name}", bins=bins, rank=rank, weighting_rank=wrank,
                moment_0=m0.data, moments=mk.data)

    # ---- calculate_displacement (displacement_methods.py:14-153) ------------------------------
    for scheme in ("ImplicitInSpace", "ExplicitInSpace"):
        be = CPU(Formulae(fastmath=False, particle_advection=scheme))
        Sb = be.Storage
        for grid in ((13,), (9, 7), (5, 4, 6)):
            n_dims, n_sd = len(grid), 400
            origin = np.stack([rng.integers(0, n, n_sd) for n in grid]).astype(np.int64)
            pos = rng.uniform(0, 1, (n_dims, n_sd))
            for n_substeps in (1, 2):
                disp = Sb.from_ndarray(np.full((n_dims, n_sd), np.nan))
                courants = []
                for dim in range(n_dims):
                    shape = tuple(n + (d == dim) for d, n in enumerate(grid))
                    courant = rng.uniform(-0.4, 0.4, shape)
                    courants.append(courant)
                    be.calculate_displacement(
                        dim=dim, displacement=disp, courant=Sb.from_ndarray(courant),
                        cell_origin=Sb.from_ndarray(origin),
                        position_in_cell=Sb.from_ndarray(pos), n_substeps=n_substeps)
                put(f"displacement_{scheme}_{n_dims}d_{n_substeps}", cell_origin=origin,
                    position_in_cell=pos, n_substeps=n_substeps, displacement=disp.data,
                    **{f"courant{d}": c for d, c in enumerate(courants)})

    # ---- Displacement.update_position / update_cell_origin / boundary_condition ---------------
    # (PySDM/dynamics/displacement.py:138-153) through the reference's Storage operators
    grid = np.array([9, 7], dtype=np.int64)
    n_sd = 300
    origin = np.stack([rng.integers(0, n, n_sd) for n in grid]).astype(np.int64)
    pos = rng.uniform(0, 1, (2, n_sd))
    disp = rng.uniform(-1.6, 1.6, (2, n_sd))
    s_origin, s_pos, s_disp = S.from_ndarray(origin.copy()), S.from_ndarray(pos.copy()), \
        S.from_ndarray(disp)
    s_tmp = S.from_ndarray(np.zeros((2, n_sd), dtype=np.int64))
    s_pos += s_disp
    s_tmp.floor(s_pos)
    s_origin += s_tmp
    s_pos -= s_tmp
    origin_before_bc = s_origin.data.copy()
    s_origin %= S.from_ndarray(grid)
    put("update_position", grid=grid, cell_origin=origin, position_in_cell=pos, displacement=disp,
        floor=s_tmp.data, cell_origin_unbounded=origin_before_bc, cell_origin_out=s_origin.data,
        position_in_cell_out=s_pos.data)

    # ---- flag_precipitated / flag_out_of_column (displacement_methods.py:155-260) -------------
    for tag, n_sd, length in (("a", 300, 300), ("b", 1000, 900)):
        nz = 12
        origin = np.stack([rng.integers(0, 9, n_sd), rng.integers(-1, nz + 1, n_sd)]).astype(np.int64)
        pos = rng.uniform(0, 1, (2, n_sd))
        disp = rng.uniform(-0.5, 0.5, (2, n_sd))
        mult = rng.integers(1, 10 ** 5, n_sd).astype(np.int64)
        mass = 1000.0 * rng.exponential(4e-12, n_sd) * rng.choice([-1.0, 1.0], n_sd)
        idx0 = rng.permutation(n_sd).astype(np.int64)
        idx = Index.from_ndarray(idx0.copy())
        idx.length = S.INT(length)
        healthy = S.from_ndarray(np.full((1,), 1))
        level = 3
        rain = backend.flag_precipitated(
            cell_origin=S.from_ndarray(origin), position_in_cell=S.from_ndarray(pos),
            water_mass=IndexedStorage.from_ndarray(idx, mass),
            multiplicity=IndexedStorage.from_ndarray(idx, mult), idx=idx, length=length,
            healthy=healthy, precipitation_counting_level_index=level,
            displacement=S.from_ndarray(disp))
        put(f"precipitated_{tag}", cell_origin=origin, position_in_cell=pos, displacement=disp,
            multiplicity=mult, water_mass=mass, idx=idx0, length=length, level=level,
            idx_out=idx.data.copy(), healthy=healthy.data.copy(), rainfall_mass=rain)
        idx = Index.from_ndarray(idx0.copy())
        idx.length = S.INT(length)
        healthy = S.from_ndarray(np.full((1,), 1))
        backend.flag_out_of_column(
            cell_origin=S.from_ndarray(origin), position_in_cell=S.from_ndarray(pos), idx=idx,
            length=length, healthy=healthy, domain_top_level_index=nz)
        put(f"out_of_column_{tag}", top=nz, idx_out=idx.data.copy(), healthy=healthy.data.copy())

    os.makedirs(gg.OUT, exist_ok=True)
    np.savez_compressed(os.path.join(gg.OUT, "transport.npz"), **g)
    print("transport.npz:", len(g), "arrays")


if __name__ == "__main__":
    gen_transport()
