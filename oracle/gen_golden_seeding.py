"""Generates tests/golden/seeding.npz by RUNNING THE REFERENCE (PySDM Numba backend): input /
output vectors of the backend method `seeding` (particle injection into empty slots,
PySDM/backends/impl_numba/methods/seeding_methods.py:12-65; caller
PySDM/particulator.py:447-490, dynamic PySDM/dynamics/seeding.py:91-117) -- SURVEY 8f rank 4.

TEST INFRASTRUCTURE (see gen_golden.py).  Usage:

    NUMBA_NUM_THREADS=1 python oracle/gen_golden_seeding.py
"""

import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)

import gen_golden as gg  # noqa: E402  (sets up the import path of the reference)
import numpy as np  # noqa: E402
from PySDM import Formulae  # noqa: E402
from PySDM.backends import CPU  # noqa: E402


def gen_seeding():
    rng = np.random.default_rng(20261106)
    g = {}
    backend = CPU(Formulae(fastmath=False))
    S = backend.Storage
    Index = gg.make_Index(backend)
    IndexedStorage = gg.make_IndexedStorage(backend)

    # (tag, n_sd, number of empty slots, A, number of seeds, number to inject)
    cases = (("few", 200, 30, 1, 12, 5), ("all_seeds", 1000, 400, 3, 64, 64),
             ("all_slots", 333, 17, 2, 40, 17), ("none", 50, 10, 1, 4, 0),
             ("one", 64, 1, 1, 1, 1), ("big", 20000, 3000, 3, 2500, 2049))
    for tag, n_sd, n_empty, n_attr, n_seeds, n_inject in cases:
        mult = rng.integers(1, 10 ** 6, n_sd).astype(np.int64)
        attrs = rng.exponential(1e-12, (n_attr, n_sd))
        empty = rng.choice(n_sd, n_empty, replace=False)
        mult[empty] = 0
        attrs[:, empty] = np.nan  # "particles with nan filled attributes" (particulator.py:458)
        idx0 = rng.permutation(n_sd).astype(np.int64)
        seed_index = rng.permutation(n_seeds).astype(np.int64)  # as after Index.shuffle
        seed_mult = rng.integers(1, 10 ** 4, n_seeds).astype(np.int64)
        seed_attrs = rng.exponential(1e-10, (n_attr, n_seeds))

        idx = Index.from_ndarray(idx0.copy())
        s_mult, s_attrs = S.from_ndarray(mult.copy()), S.from_ndarray(attrs.copy())
        s_index = Index.from_ndarray(seed_index.copy())
        backend.seeding(
            idx=idx, multiplicity=s_mult, extensive_attributes=s_attrs,
            seeded_particle_index=s_index,
            seeded_particle_multiplicity=IndexedStorage.from_ndarray(s_index, seed_mult.copy()),
            seeded_particle_extensive_attributes=IndexedStorage.from_ndarray(
                s_index, seed_attrs.copy()),
            number_of_super_particles_to_inject=n_inject)
        for k, v in dict(idx=idx0, multiplicity=mult, attributes=attrs, seed_index=seed_index,
                         seed_multiplicity=seed_mult, seed_attributes=seed_attrs,
                         n_inject=n_inject, idx_out=idx.data.copy(),
                         multiplicity_out=s_mult.data.copy(),
                         attributes_out=s_attrs.data.copy()).items():
            g[f"seeding_{tag}/{k}"] = np.asarray(v)

    os.makedirs(gg.OUT, exist_ok=True)
    np.savez_compressed(os.path.join(gg.OUT, "seeding.npz"), **g)
    print("seeding.npz:", len(g), "arrays")


if __name__ == "__main__":
    gen_seeding()
