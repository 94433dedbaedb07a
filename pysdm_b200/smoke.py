"""smoke test for __graft_entry__.smoke(): one small multi-cell Golovin collision run on
cuda:0 through the fused kernel and the method-level kernels, checked bit-for-bit against
the CPU oracle (the oracle is only the checker here)."""

import numpy as np


def run(n_cell=64, sd_per_cell=128, steps=3):
    from oracle.sdm_step import Golovin as OGolovin  # pylint: disable=import-outside-toplevel
    from oracle.sdm_step import OracleCollision  # pylint: disable=import-outside-toplevel

    from .collision import DeviceCollision, Golovin  # pylint: disable=import-outside-toplevel

    rng = np.random.default_rng(7)
    n_sd = n_cell * sd_per_cell
    pop = {
        "multiplicity": rng.integers(1, 5, n_sd).astype(np.int64),
        "water_mass": rng.exponential(1000 * 4 / 3 * np.pi * 30.531e-6 ** 3, n_sd),
        "cell_id": rng.permutation(np.arange(n_sd) // sd_per_cell),
        "n_cell": n_cell,
    }
    common = {"dt": 1.0, "dv": 2e-7, "seed": 5}
    ora = OracleCollision(kernel=OGolovin(1.5e3), **pop, **common)
    sims = [DeviceCollision(kernel=Golovin(1.5e3), host_rng=True, fused=f, **pop, **common)
            for f in (True, False)]
    for _ in range(steps):
        ora()
        for sim in sims:
            sim()
    for sim in sims:
        n = ora.length
        assert sim.length == n
        assert (sim.host("idx")[:n] == ora.idx[:n]).all()
        assert (sim.host("multiplicity") == ora.multiplicity).all()
        assert (sim.host("attributes").view(np.int64) == ora.attributes.view(np.int64)).all()
        assert (sim.host("coalescence_rate") == ora.coalescence_rate).all()
    assert sims[0].n_fused_steps == steps and ora.coalescence_rate.sum() > 0
    print(f"smoke ok: {n_sd} SDs, {steps} steps, {int(ora.coalescence_rate.sum())} coalesced, "
          f"length {ora.length}")
