"""Rows either side of the collision step: `moments`, `spectrum_moments`,
`calculate_displacement`, `flag_precipitated`, `flag_out_of_column` and the Storage operator
chain of `Displacement` -- the C oracle (CPU) and, with `-m gpu`, the CUDA kernels through the
C-ABI, both against tests/golden/transport.npz, vectors produced by RUNNING THE REFERENCE
(PySDM's Numba backend, one thread; generator: oracle/gen_golden_transport.py).

Bar: the oracle (serial sums in the reference's order) is bit-exact.  The CUDA kernels sum
per-cell contributions with fp64 atomics in a different order and use CUDA's pow: 1e-12
relative for sums of one sign (the tolerance north_star states for attributes), integer /
index outputs and the displacement arithmetic bit-exact."""

import os

import numpy as np
import pytest

from oracle import sdm_oracle as orc

from .golden_util import GOLDEN, assert_same_f64, case

RTOL = 1e-12


@pytest.fixture(scope="module")
def g():
    return np.load(os.path.join(GOLDEN, "transport.npz"))


def close(be, got, expected, what):
    if be is orc:
        assert_same_f64(got, expected, what)
    else:
        np.testing.assert_allclose(got, expected, rtol=RTOL, atol=0, equal_nan=True, err_msg=what)


def _rank(value):
    """scalars come back from the npz as 0-d arrays: restore the Python type the reference saw"""
    value = value[()]
    return int(value) if np.issubdtype(value.dtype, np.integer) else float(value)


@pytest.mark.parametrize("variant", ("plain", "range", "weighted", "norank"))
@pytest.mark.parametrize("tag", "abc")
def test_moments(be, g, tag, variant):
    pop, c = case(g, f"moments_{tag}"), case(g, f"moments_{tag}_{variant}")
    n_cell, ranks = int(pop["n_cell"]), c["ranks"]
    m0 = np.full(n_cell, np.nan)
    mk = np.full((len(ranks), n_cell), np.nan)
    be.moments(m0, mk, pop["multiplicity"], pop["volume"], pop["cell_id"], pop["idx"],
               int(pop["length"]), ranks, float(c["min_x"]), float(c["max_x"]), pop["mass"],
               np.abs(pop["mass"]), _rank(c["weighting_rank"]), bool(c["skip"]))
    close(be, m0, c["moment_0"], "moment_0")
    close(be, mk, c["moments"], "moments")
    assert np.count_nonzero(c["moment_0"]) > 0


@pytest.mark.parametrize("variant", ("r1", "w1r0", "r2_5"))
@pytest.mark.parametrize("tag", "abc")
def test_spectrum_moments(be, g, tag, variant):
    pop, c = case(g, f"moments_{tag}"), case(g, f"spectrum_{tag}_{variant}")
    n_cell, n_bins = int(pop["n_cell"]), len(c["bins"]) - 1
    m0, mk = np.full((n_bins, n_cell), np.nan), np.full((n_bins, n_cell), np.nan)
    be.spectrum_moments(m0, mk, pop["multiplicity"], pop["volume"], pop["cell_id"], pop["idx"],
                        int(pop["length"]), _rank(c["rank"]), c["bins"], pop["mass"],
                        np.abs(pop["mass"]), _rank(c["weighting_rank"]))
    close(be, m0, c["moment_0"], "moment_0")
    close(be, mk, c["moments"], "moments")


@pytest.mark.parametrize("n_substeps", (1, 2))
@pytest.mark.parametrize("n_dims", (1, 2, 3))
@pytest.mark.parametrize("scheme", ("ImplicitInSpace", "ExplicitInSpace"))
def test_calculate_displacement(be, g, scheme, n_dims, n_substeps):
    c = case(g, f"displacement_{scheme}_{n_dims}d_{n_substeps}")
    disp = np.full_like(c["displacement"], np.nan)
    for dim in range(n_dims):
        be.calculate_displacement(dim, disp, c[f"courant{dim}"], c["cell_origin"],
                                  c["position_in_cell"], int(c["n_substeps"]),
                                  implicit_in_space=scheme == "ImplicitInSpace")
    assert_same_f64(disp, c["displacement"], "displacement")  # bit-exact on both paths


def test_update_position_storage_chain(be, g):
    c = case(g, "update_position")
    origin, pos = c["cell_origin"].copy(), c["position_in_cell"].copy()
    tmp = np.zeros_like(origin)
    be.update_position(origin, pos, c["displacement"], tmp, c["grid"])
    np.testing.assert_array_equal(tmp, c["floor"])
    np.testing.assert_array_equal(origin, c["cell_origin_out"])
    assert_same_f64(pos, c["position_in_cell_out"], "position_in_cell")
    assert (c["cell_origin_unbounded"] < 0).any() and (c["floor"] != 0).any()


@pytest.mark.parametrize("tag", "ab")
def test_flag_precipitated_and_out_of_column(be, g, tag):
    c, o = case(g, f"precipitated_{tag}"), case(g, f"out_of_column_{tag}")
    idx, healthy = c["idx"].copy(), np.ones(1, dtype=np.int64)
    rain = be.flag_precipitated(c["cell_origin"], c["position_in_cell"], c["water_mass"],
                                c["multiplicity"], idx, int(c["length"]), healthy,
                                int(c["level"]), c["displacement"])
    np.testing.assert_array_equal(idx, c["idx_out"])
    np.testing.assert_array_equal(healthy, c["healthy"])
    if be is orc:
        assert rain == float(c["rainfall_mass"])
    else:
        np.testing.assert_allclose(rain, float(c["rainfall_mass"]), rtol=RTOL)
    assert float(c["rainfall_mass"]) > 0 and (c["idx_out"] == len(idx)).any()
    idx, healthy = c["idx"].copy(), np.ones(1, dtype=np.int64)
    be.flag_out_of_column(c["cell_origin"], c["position_in_cell"], idx, int(c["length"]), healthy,
                          int(o["top"]))
    np.testing.assert_array_equal(idx, o["idx_out"])
    np.testing.assert_array_equal(healthy, o["healthy"])
    assert (o["idx_out"] == len(idx)).any()


@pytest.mark.gpu
@pytest.mark.parametrize("n_substeps", (1, 2))
@pytest.mark.parametrize("n_dims", (1, 2, 3))
@pytest.mark.parametrize("scheme", ("ImplicitInSpace", "ExplicitInSpace"))
def test_displacement_step_one_launch(g, scheme, n_dims, n_substeps):
    """pysdm_b200_displacement_step (Displacement.__call__ without sedimentation in one launch,
    PySDM/dynamics/displacement.py:103-153) against the reference's displacement vectors and
    the Storage operator chain applied to them in numpy, bit for bit; Courant numbers scaled up
    so that positions leave their cells in both directions and wrap around the domain"""
    import torch  # pylint: disable=import-outside-toplevel

    from pysdm_b200 import methods as m  # pylint: disable=import-outside-toplevel

    from . import device_adapter  # pylint: disable=import-outside-toplevel

    c = case(g, f"displacement_{scheme}_{n_dims}d_{n_substeps}")
    origin, pos = c["cell_origin"], c["position_in_cell"]
    grid = [c["courant0"].shape[0] - 1] + list(c["courant0"].shape[1:])
    for scale in (1.0, -3.7):
        courant = [np.ascontiguousarray(c[f"courant{dim}"] * scale) for dim in range(n_dims)]
        if scale == 1.0:
            disp = c["displacement"]  # the reference's own output
        else:  # the method-level kernel (pinned on the reference's vectors above)
            disp = np.full_like(c["displacement"], np.nan)
            for dim in range(n_dims):
                device_adapter.calculate_displacement(
                    dim, disp, courant[dim], origin, pos, int(c["n_substeps"]),
                    implicit_in_space=scheme == "ImplicitInSpace")
        moved = pos + disp
        floor = np.floor(moved).astype(np.int64)
        expected_origin = (origin + floor) % np.asarray(grid, dtype=np.int64)[:, None]
        expected_pos = moved - floor
        dev = torch.device("cuda:0")
        d_disp = torch.full(disp.shape, float("nan"), dtype=torch.float64, device=dev)
        d_origin, d_pos = torch.from_numpy(origin.copy()).to(dev), torch.from_numpy(pos.copy()).to(dev)
        m.displacement_step(d_disp, [torch.from_numpy(x).to(dev) for x in courant], d_origin,
                            d_pos, grid, n_substeps=int(c["n_substeps"]),
                            implicit_in_space=scheme == "ImplicitInSpace")
        assert_same_f64(d_disp.cpu().numpy(), disp, "displacement")
        np.testing.assert_array_equal(d_origin.cpu().numpy(), expected_origin)
        assert_same_f64(d_pos.cpu().numpy(), expected_pos, "position_in_cell")
        if scale != 1.0:
            assert (floor < 0).any() and (floor > 0).any()
            assert ((origin + floor) < 0).any() or ((origin + floor) >= np.asarray(grid)[:, None]).any()
