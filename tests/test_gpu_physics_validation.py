"""GPU: physical validation of the whole device path with the DEVICE random generator, where
bit-parity with the reference no longer applies: the Golovin box of
Shima et al. 2009 (PySDM_examples.Shima_et_al_2009 settings: b = 1.5e3 / s, n_part = 2^23 / m^3,
X0 = volume of a 30.531 um droplet, dt = 1 s, dv = 1e6 m^3) as 64 independent cells of 512
super-droplets, fused one-launch steps with in-kernel Philox uniforms, droplet number per cell
from the `moments` kernel.  For the additive kernel K = b (x + y) the Smoluchowski equation
gives dN/dt = -b L N exactly (L = liquid water volume fraction, conserved), the same balance
the reference tests against through `Golovin.analytic_solution`
(PySDM/dynamics/collisions/collision_kernels/golovin.py:22-44,
tests/smoke_tests/box/shima_et_al_2009/test_convergence.py)."""

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_golovin_box_number_decay_with_device_rng():
    import torch  # pylint: disable=import-outside-toplevel

    from pysdm_b200 import methods as m  # pylint: disable=import-outside-toplevel
    from pysdm_b200.collision import DeviceCollision, Golovin  # pylint: disable=import-outside-toplevel

    n_cell, per_cell, b, dv, dt = 64, 512, 1.5e3, 1e6, 1.0
    n_sd = n_cell * per_cell
    x0 = 4 / 3 * np.pi * 30.531e-6 ** 3
    rng = np.random.default_rng(44)
    volume = rng.exponential(x0, n_sd)
    mult = np.full(n_sd, int(2 ** 23 * dv / per_cell), dtype=np.int64)
    sim = DeviceCollision(multiplicity=mult, water_mass=1000.0 * volume,
                          cell_id=np.arange(n_sd) // per_cell, n_cell=n_cell, dt=dt, dv=dv,
                          kernel=Golovin(b), seed=44, fused=True)
    dev = sim.multiplicity.device
    m0 = torch.empty(n_cell, dtype=torch.float64, device=dev)
    mk = torch.empty((1, n_cell), dtype=torch.float64, device=dev)

    def number_and_volume():
        """per cell: sum n (0th moment) and sum n v (1st moment of volume, not divided)"""
        m.moments(m0, mk, sim.multiplicity, sim.volume(), sim.cell_id, sim.idx, sim.length,
                  [1.0], -np.inf, np.inf, sim.attributes[0], sim.attributes[0], 0, True)
        return m0.cpu().numpy().copy(), mk.cpu().numpy()[0].copy()

    n_start, v_start = number_and_volume()
    liquid = v_start / dv
    np.testing.assert_allclose(liquid, 2 ** 23 * x0, rtol=0.15)  # ~1e-6, sampling noise
    for t_end in (300, 600):
        sim.run(300)
        n_now, v_now = number_and_volume()
        np.testing.assert_allclose(v_now, v_start, rtol=1e-12)  # water volume is conserved
        ratio = n_now / n_start
        expect = np.exp(-b * liquid * t_end)
        # cells are independent realisations: the ensemble mean must follow the analytic decay
        assert abs(ratio.mean() / expect.mean() - 1) < 0.03, (t_end, ratio.mean(), expect.mean())
        # ... and no single cell may be far off (stochastic scatter of 512 super-droplets)
        assert np.max(np.abs(ratio / expect - 1)) < 0.35, (t_end, ratio / expect)
    assert sim.n_fused_steps == 600
