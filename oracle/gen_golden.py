"""Generates tests/golden/*.npz by RUNNING THE REFERENCE (PySDM Numba backend,
Formulae(fastmath=False)) in the build container.

TEST INFRASTRUCTURE.  Needs /root/reference (read-only) and the import stand-ins
under oracle/refshims; neither exists on the GPU box, which only sees the
committed .npz files.  Usage:

    NUMBA_NUM_THREADS=1 python oracle/gen_golden.py [methods|scenarios|all]

Every vector is produced by calling the reference's public backend methods or
its Builder/Particulator/Collision objects; nothing is computed by our code.
Uniform random numbers are NOT stored: the reference's Random
(backends/impl_numba/random.py:9-19) is numpy default_rng(seed).uniform, so a
replay regenerates them from the recorded seed in the recorded call order.
"""

import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "refshims"))
sys.path.insert(1, "/root/reference")
os.environ.setdefault("NUMBA_NUM_THREADS", "1")

import numpy as np  # noqa: E402

np.NaN = np.nan  # the reference predates NumPy 2

from PySDM import Builder, Formulae  # noqa: E402
from PySDM.backends import CPU  # noqa: E402
from PySDM.backends.impl_common.index import make_Index  # noqa: E402
from PySDM.backends.impl_common.indexed_storage import make_IndexedStorage  # noqa: E402
from PySDM.backends.impl_common.pair_indicator import make_PairIndicator  # noqa: E402
from PySDM.backends.impl_common.pairwise_storage import make_PairwiseStorage  # noqa: E402
from PySDM.dynamics import Coalescence  # noqa: E402
from PySDM.dynamics.collisions.collision_kernels import Golovin  # noqa: E402
from PySDM.environments import Box  # noqa: E402
from PySDM.impl.mesh import Mesh  # noqa: E402

OUT = os.path.join(os.path.dirname(HERE), "tests", "golden")


def backend_and_types(**formulae_kw):
    backend = CPU(Formulae(fastmath=False, **formulae_kw))
    return (backend, make_Index(backend), make_IndexedStorage(backend),
            make_PairIndicator(backend), make_PairwiseStorage(backend))


def random_cells(rng, n_sd, n_cell, length=None):
    """cell ids with an empty cell and a single-SD cell, a permuted idx"""
    length = n_sd if length is None else length
    cid = rng.integers(0, n_cell, n_sd)
    if n_cell > 3:
        cid[cid == 1] = 0  # cell 1 empty
        cid[cid == 2] = 3
        cid[0] = 2  # cell 2 holds exactly one SD
    idx = rng.permutation(n_sd).astype(np.int64)
    return cid.astype(np.int64), idx, length


def gen_methods():
    """method-level input/output vectors (one npz, keys '<case>/<name>')"""
    backend, Index, IndexedStorage, PairIndicator, PairwiseStorage = backend_and_types()
    S = backend.Storage
    rng = np.random.default_rng(20261104)
    g = {}

    def put(case, **kw):
        for k, v in kw.items():
            g[f"{case}/{k}"] = np.asarray(v)

    for tag, n_sd, n_cell, length in (("a", 57, 6, 57), ("b", 200, 1, 200), ("c", 64, 9, 50),
                                      ("d", 1000, 37, 1000)):
        cid, idx0, length = random_cells(rng, n_sd, n_cell, length)
        cell_idx0 = rng.permutation(n_cell).astype(np.int64) if tag != "a" else np.arange(n_cell)
        # --- counting sort (make_cell_caretaker) for both schemes
        for scheme in ("counting_sort", "counting_sort_parallel"):
            idx = Index.from_ndarray(idx0.copy())
            idx.length = S.INT(length)
            cell_id = IndexedStorage.from_ndarray(idx, cid)
            cell_idx = Index.from_ndarray(cell_idx0.copy())
            cell_start = S.from_ndarray(np.zeros(n_cell + 1, dtype=np.int64))
            caretaker = backend.make_cell_caretaker(idx.shape, idx.dtype, n_cell + 1, scheme)
            caretaker(cell_id, cell_idx, cell_start, idx)
            put(f"sort_{tag}_{scheme}", idx_in=idx0, cell_id=cid, cell_idx=cell_idx0,
                length=length, idx_out=idx.data[:length], cell_start=cell_start.data)
        sorted_idx = idx.data.copy()
        sorted_idx[length:] = idx0[length:]
        cs = cell_start.data.copy()

        # --- shuffle_local / shuffle_global
        u01 = rng.uniform(0, 1, n_sd)
        idx = Index.from_ndarray(sorted_idx.copy())
        idx.length = S.INT(length)
        idx.shuffle(S.from_ndarray(u01), parts=S.from_ndarray(cs))
        put(f"shuffle_local_{tag}", idx_in=sorted_idx, u01=u01, cell_start=cs, idx_out=idx.data)
        shuffled = idx.data.copy()
        idx = Index.from_ndarray(sorted_idx.copy())
        idx.length = S.INT(length)
        idx.shuffle(S.from_ndarray(u01))
        put(f"shuffle_global_{tag}", idx_in=sorted_idx, u01=u01, length=length,
            idx_out=idx.data)

        # --- find_pairs
        idx = Index.from_ndarray(shuffled.copy())
        idx.length = S.INT(length)
        cell_id = IndexedStorage.from_ndarray(idx, cid)
        cell_idx = Index.from_ndarray(cell_idx0.copy())
        flag = PairIndicator(n_sd)
        flag.indicator.data[:] = False
        flag.update(S.from_ndarray(cs), cell_idx, cell_id)
        put(f"find_pairs_{tag}", idx=shuffled, length=length, cell_id=cid, cell_idx=cell_idx0,
            cell_start=cs, flag=flag.indicator.data.copy())

        # --- sort_within_pair_by_attr (multiplicity)
        mult = rng.integers(1, 6, n_sd).astype(np.int64)
        n = IndexedStorage.from_ndarray(idx, mult)
        backend.sort_within_pair_by_attr(idx, flag, n)
        put(f"sort_within_pair_{tag}", idx_in=shuffled, mult=mult, idx_out=idx.data.copy())
        paired = idx.data.copy()

        # --- pairwise reductions
        vol = rng.exponential(1e-12, n_sd)
        v = IndexedStorage.from_ndarray(idx, vol)
        out = PairwiseStorage.empty(n_sd // 2, dtype=float)
        for op in ("distance", "max", "min", "sum", "multiply"):
            out.data[:] = np.nan
            getattr(out, op)(v, flag)
            put(f"pair_{op}_{tag}", out=out.data.copy())
        out.data[:] = np.nan
        out.max(n, flag)
        put(f"pair_max_int_{tag}", out=out.data.copy())
        outn = S.empty(n_sd, dtype=float)
        backend.sort_pair(outn, v, flag, idx)
        put(f"pair_ops_{tag}", idx=paired, vol=vol, mult=mult, flag=flag.indicator.data.copy(),
            length=length, sort_pair=outn.data.copy())

        # --- normalize (prob = max(n) * (v1+v2) * b)
        prob = PairwiseStorage.empty(n_sd // 2, dtype=float)
        prob.sum(v, flag)
        prob *= 1.5e3 * 2e9
        tmp = PairwiseStorage.empty(n_sd // 2, dtype=float)
        tmp.max(n, flag)
        prob *= tmp
        prob_in = prob.data.copy()
        norm_factor = S.empty(n_cell, dtype=float)
        backend.normalize(prob=prob, cell_id=cell_id, cell_idx=cell_idx,
                          cell_start=S.from_ndarray(cs), norm_factor=norm_factor, timestep=1.5,
                          dv=0.7)
        put(f"normalize_{tag}", prob_in=prob_in, prob_out=prob.data.copy(),
            norm_factor=norm_factor.data.copy(), timestep=1.5, dv=0.7)

        # --- scale_prob_for_adaptive_sdm_gamma on a copy of prob
        for first in (True, False):
            prob_a = PairwiseStorage.from_ndarray(prob.data.copy())
            dt_left = S.from_ndarray(rng.uniform(0.0, 2.0, n_cell))
            dt_left.data[0] = 0.0
            dt_left_in = dt_left.data.copy()
            n_substep = S.from_ndarray(rng.integers(0, 3, n_cell).astype(np.int64))
            n_substep_in = n_substep.data.copy()
            dt_min_stat = S.from_ndarray(
                np.full(n_cell, np.nan) if first else rng.uniform(0.001, 3.0, n_cell))
            dt_min_in = dt_min_stat.data.copy()
            backend.scale_prob_for_adaptive_sdm_gamma(
                prob=prob_a, multiplicity=n, cell_id=cell_id, dt_left=dt_left, dt=2.0,
                dt_range=(0.01, 1.25), is_first_in_pair=flag, stats_n_substep=n_substep,
                stats_dt_min=dt_min_stat)
            put(f"adaptive_{tag}_{int(first)}", prob_in=prob.data.copy(),
                prob_out=prob_a.data.copy(), dt_left_in=dt_left_in, dt_left_out=dt_left.data,
                n_substep_in=n_substep_in, n_substep_out=n_substep.data, dt_min_in=dt_min_in,
                dt_min_out=dt_min_stat.data, dt=2.0, dt_range=(0.01, 1.25))
            put(f"adaptive_end_{tag}_{int(first)}",
                end=backend.adaptive_sdm_end(dt_left, S.from_ndarray(cs)))

        # --- compute_gamma (in place: out aliases prob, as collision.py:203,217-219)
        rand = rng.uniform(0, 1, n_sd // 2)
        rate = S.from_ndarray(np.zeros(n_cell, dtype=np.int64))
        deficit = S.from_ndarray(np.zeros(n_cell, dtype=np.int64))
        backend.compute_gamma(prob=prob, rand=S.from_ndarray(rand), multiplicity=n,
                              cell_id=cell_id, collision_rate_deficit=deficit,
                              collision_rate=rate, is_first_in_pair=flag, out=prob)
        put(f"gamma_{tag}", rand=rand, gamma=prob.data.copy(), rate=rate.data, deficit=deficit.data)

        # --- collision_coalescence with A = 1 and 3 attribute rows
        for n_attr in (1, 3):
            attrs = rng.exponential(1e-9, (n_attr, n_sd))
            a = IndexedStorage.from_ndarray(idx, attrs.copy())
            nn = IndexedStorage.from_ndarray(idx, mult.copy())
            healthy = S.from_ndarray(np.full((1,), 1))
            coal = S.from_ndarray(np.zeros(n_cell, dtype=np.int64))
            backend.collision_coalescence(multiplicity=nn, idx=idx, attributes=a, gamma=prob,
                                          healthy=healthy, cell_id=cell_id,
                                          coalescence_rate=coal, is_first_in_pair=flag)
            put(f"coalescence_{tag}_{n_attr}", attrs_in=attrs, attrs_out=a.data, mult_out=nn.data,
                healthy=healthy.data, coal_rate=coal.data)
            # --- remove_zero_n_or_flagged on the result
            idx2 = Index.from_ndarray(paired.copy())
            idx2.length = S.INT(length)
            idx2.remove_zero_n_or_flagged(nn)
            put(f"remove_zero_{tag}_{n_attr}", idx_out=idx2.data.copy(), new_length=int(idx2.length))

        # --- sort_by_key (cell_idx <- argsort(dt_left)[::-1]); ties included
        keys = np.round(rng.uniform(0, 2, n_cell), 1)
        ci = Index.from_ndarray(np.arange(n_cell, dtype=np.int64))
        ci.sort_by_key(S.from_ndarray(keys))
        put(f"sort_by_key_{tag}", keys=keys, out=ci.data.copy())

        # --- cell_id from cell_origin
        grid = (5, 7)
        origin = np.stack([rng.integers(0, grid[0], n_sd), rng.integers(0, grid[1], n_sd)])
        strides = Mesh(grid, size=grid).strides
        cid_out = S.from_ndarray(np.zeros(n_sd, dtype=np.int64))
        backend.cell_id(cid_out, S.from_ndarray(origin), S.from_ndarray(strides))
        put(f"cell_id_{tag}", origin=origin, strides=strides, out=cid_out.data)

    os.makedirs(OUT, exist_ok=True)
    np.savez_compressed(os.path.join(OUT, "methods.npz"), **g)
    print("methods.npz:", len(g), "arrays")


def snapshot(particulator, dyn):
    attrs = particulator.attributes
    idx = attrs._ParticleAttributes__idx  # pylint: disable=protected-access
    snap = {
        "length": int(idx.length),
        "idx": idx.data.copy(),
        "multiplicity": attrs["multiplicity"].to_ndarray(raw=True),
        "attributes": attrs.get_extensive_attribute_storage().to_ndarray(raw=True),
        "flag": dyn.is_first_in_pair.indicator.data.copy(),
        "gamma": dyn.gamma.data.copy(),
        "collision_rate": dyn.collision_rate.data.copy(),
        "collision_rate_deficit": dyn.collision_rate_deficit.data.copy(),
        "coalescence_rate": dyn.coalescence_rate.data.copy(),
        "stats_n_substep": dyn.stats_n_substep.data.copy(),
        "stats_dt_min": dyn.stats_dt_min.data.copy(),
        "cell_start": attrs._ParticleAttributes__cell_start.data.copy(),  # pylint: disable=protected-access
    }
    if dyn.enable_breakup:
        snap["breakup_rate"] = dyn.breakup_rate.data.copy()
        snap["breakup_rate_deficit"] = dyn.breakup_rate_deficit.data.copy()
    return snap


def run_scenario(name, *, n_sd, grid, dv, dt, steps, seed, make_dynamic, init, formulae_kw=None):
    """drives the reference Builder/Particulator and records the state after
    every model step"""
    formulae = Formulae(fastmath=False, seed=seed, **(formulae_kw or {}))
    backend = CPU(formulae)
    env = Box(dv=dv, dt=dt)
    n_cell = int(np.prod(grid))
    if n_cell > 1:
        env.mesh = Mesh(grid, size=grid, dv=dv)
    builder = Builder(n_sd=n_sd, backend=backend, environment=env)
    builder.add_dynamic(make_dynamic())
    rng = np.random.default_rng(seed + 1)
    attributes = init(rng, n_sd, n_cell)
    initial = {k: np.array(v) for k, v in attributes.items()}
    if n_cell == 1:
        attributes.pop("cell id", None)
    particulator = builder.build(attributes=attributes, products=())
    dyn = particulator.dynamics["Collision"]
    out = {f"init/{k}": v for k, v in initial.items()}
    out["meta/n_sd"] = n_sd
    out["meta/n_cell"] = n_cell
    out["meta/dv"] = float(particulator.mesh.dv)
    out["meta/dt"] = dt
    out["meta/steps"] = steps
    out["meta/seed"] = seed
    out["meta/rho_w"] = float(formulae.constants.rho_w)
    out["init/signed water mass"] = particulator.attributes["signed water mass"].to_ndarray(raw=True)
    for step in range(steps):
        particulator.run(1)
        for k, v in snapshot(particulator, dyn).items():
            out[f"step{step}/{k}"] = np.asarray(v)
    os.makedirs(OUT, exist_ok=True)
    np.savez_compressed(os.path.join(OUT, f"scenario_{name}.npz"), **out)
    last = snapshot(particulator, dyn)
    print(f"scenario_{name}: length {n_sd} -> {last['length']},"
          f" coalescences {last['coalescence_rate'].sum()},"
          f" substeps {last['stats_n_substep'].tolist()[:8]}")


def init_multicell(rng, n_sd, n_cell):
    cid = rng.integers(0, n_cell, n_sd)
    if n_cell > 3:
        cid[cid == 1] = 0
        cid[cid == 2] = 3
        cid[0] = 2
    return {
        "multiplicity": rng.integers(1, 5, n_sd).astype(np.int64),
        "volume": rng.exponential(4.0 / 3 * np.pi * (30.531e-6) ** 3, n_sd),
        "cell id": cid.astype(np.int64),
    }


def gen_scenarios():
    b = 1.5e3
    common = {"n_sd": 301, "grid": (3, 3), "dv": 2e-8, "dt": 1.0, "steps": 24,
              "init": init_multicell}
    run_scenario("golovin_local", seed=44, **common,
                 make_dynamic=lambda: Coalescence(collision_kernel=Golovin(b=b), adaptive=False,
                                                  croupier="local"))
    run_scenario("golovin_global", seed=45, **common,
                 make_dynamic=lambda: Coalescence(collision_kernel=Golovin(b=b), adaptive=False,
                                                  croupier="global"))
    run_scenario("golovin_substeps", seed=46, **common,
                 make_dynamic=lambda: Coalescence(collision_kernel=Golovin(b=b), adaptive=False,
                                                  substeps=3))
    run_scenario("golovin_adaptive", seed=47, **{**common, "dv": 4e-9, "steps": 12},
                 make_dynamic=lambda: Coalescence(collision_kernel=Golovin(b=b), adaptive=True,
                                                  dt_coal_range=(0.01, 1.0)))
    run_scenario("golovin_adaptive_optrand", seed=48, **{**common, "dv": 4e-9, "steps": 12},
                 make_dynamic=lambda: Coalescence(collision_kernel=Golovin(b=b), adaptive=True,
                                                  dt_coal_range=(0.05, 1.0),
                                                  optimized_random=True))
    run_scenario("golovin_box", seed=49, **{**common, "grid": (1,), "n_sd": 256, "dv": 5e-8},
                 make_dynamic=lambda: Coalescence(collision_kernel=Golovin(b=b), adaptive=False))


if __name__ == "__main__":
    what = sys.argv[1] if len(sys.argv) > 1 else "all"
    if what in ("methods", "all"):
        gen_methods()
    if what in ("scenarios", "all"):
        gen_scenarios()
