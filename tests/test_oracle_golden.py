"""The C oracle (CPU) and, with `-m gpu`, the CUDA method-level kernels through the
C-ABI (tests/device_adapter.py), both against vectors produced by the reference itself
(tests/golden/*.npz, made by oracle/gen_golden.py from PySDM's Numba backend with
Formulae(fastmath=False)) and against the known-answer vectors quoted in the
reference's own unit tests.  Bit-exact everywhere."""

import numpy as np
import pytest

from oracle import sdm_oracle as orc
from oracle.sdm_step import Golovin, OracleCollision

from .golden_util import (GOLOVIN_B, SCENARIO_KW, SCENARIOS, TAGS, assert_same_f64, case,
                          load_scenario)


# ---- known-answer vectors quoted from the reference's tests -----------------------
NONZERO = 44.0


@pytest.mark.parametrize("i, idx, flag, gamma, expected", (
    # tests/unit_tests/backends/test_collisions_methods.py:17-60
    (0, (0, 1), (1, 0), (NONZERO,), (0, 1, False)),
    (0, (1, 0), (1, 0), (NONZERO,), (1, 0, False)),
    (0, (0, 1, 2), (0, 1, 0), (NONZERO,), (1, 2, False)),
    (1, range(8), (1, 0, 0, 1, 0, 0, 1, 0), (NONZERO, NONZERO), (3, 4, False)),
    (2, range(8), (1, 0, 0, 1, 0, 0, 1, 0), (NONZERO, NONZERO, 0), (-1, -1, True)),
    (3, range(8), (1, 0, 0, 1, 0, 0, 1, 0), (NONZERO, NONZERO, 0, NONZERO), (6, 7, False)),
))
def test_pair_indices_truth_table(i, idx, flag, gamma, expected):
    idx = np.asarray(list(idx) + [0], dtype=np.int64)
    flag = np.asarray(list(flag) + [0], dtype=np.bool_)
    j, k, skip = orc.pair_indices(idx, flag, np.asarray(gamma, dtype=np.float64))
    assert (j[i], k[i], skip[i]) == expected


@pytest.mark.parametrize("dt_left, cell_start, expected", (
    # test_collisions_methods.py:62-80
    ((4, 5, 4.5, 0, 0), (0, 1, 2, 3, 4, 5), 3),
    ((4, 5, 4.5, 3, 0.1), (0, 1, 2, 3, 4, 5), 5),
))
def test_adaptive_sdm_end_vectors(be, dt_left, cell_start, expected):
    assert be.adaptive_sdm_end(np.asarray(dt_left, dtype=np.float64),
                                np.asarray(cell_start, dtype=np.int64)) == expected


U01 = np.asarray([0.1, 0.4, 0.2, 0.5, 0.9, 0.1, 0.6, 0.3])


def test_shuffle_global_vector(be):
    # tests/unit_tests/impl/test_particle_attributes.py:149-171
    idx = np.arange(8, dtype=np.int64)
    be.shuffle_global(idx, 8, U01)
    np.testing.assert_array_equal(idx, [1, 3, 5, 7, 6, 0, 4, 2])


def test_shuffle_local_vector(be):
    # tests/unit_tests/impl/test_particle_attributes.py:173-201
    idx = np.arange(8, dtype=np.int64)
    be.shuffle_local(idx, U01, np.asarray([0, 0, 2, 5, 7, 8], dtype=np.int64))
    np.testing.assert_array_equal(idx, [1, 0, 2, 3, 4, 5, 6, 7])


def test_compute_gamma_grid(be):
    # tests/unit_tests/dynamics/collisions/test_sdm_single_cell.py:216-258:
    # gamma = floor(p) + (r < p - floor(p)) on a grid of (prob, rand)
    probs = np.linspace(0, 3, 87)
    rands = np.linspace(0, 1, 87, endpoint=False)
    for p in probs:
        n = len(rands)
        prob = np.full(n, p)
        idx = np.arange(2 * n, dtype=np.int64)
        flag = np.tile(np.asarray([True, False]), n)
        mult = np.tile(np.asarray([10, 1], dtype=np.int64), n)
        rate, deficit = np.zeros(1, np.int64), np.zeros(1, np.int64)
        be.compute_gamma(prob, rands.copy(), idx, 2 * n, mult, np.zeros(2 * n, np.int64),
                          deficit, rate, flag, prob)
        expected = np.floor(p) + (rands < p - np.floor(p))
        np.testing.assert_array_equal(prob, expected)


# ---- method-level vectors from the reference run ---------------------------------------
@pytest.mark.parametrize("tag", TAGS)
@pytest.mark.parametrize("scheme", ("counting_sort", "counting_sort_parallel"))
def test_counting_sort(be, golden_methods, tag, scheme):
    c = case(golden_methods, f"sort_{tag}_{scheme}")
    new_idx = np.full(len(c["idx_in"]), -1, dtype=np.int64)
    cell_start = np.zeros(len(c["cell_start"]), dtype=np.int64)
    length = int(c["length"])
    be.counting_sort_by_cell_id(new_idx, c["idx_in"].copy(), c["cell_id"], c["cell_idx"], length,
                                 cell_start)
    np.testing.assert_array_equal(new_idx[:length], c["idx_out"])
    np.testing.assert_array_equal(cell_start, c["cell_start"])


@pytest.mark.parametrize("tag", TAGS)
def test_shuffles(be, golden_methods, tag):
    c = case(golden_methods, f"shuffle_local_{tag}")
    idx = c["idx_in"].copy()
    be.shuffle_local(idx, c["u01"], c["cell_start"])
    np.testing.assert_array_equal(idx, c["idx_out"])
    c = case(golden_methods, f"shuffle_global_{tag}")
    idx = c["idx_in"].copy()
    be.shuffle_global(idx, int(c["length"]), c["u01"])
    np.testing.assert_array_equal(idx, c["idx_out"])


@pytest.mark.parametrize("tag", TAGS)
def test_pair_pipeline(be, golden_methods, tag):
    """find_pairs -> sort_within_pair -> pair reductions -> normalize -> adaptive
    scaling -> gamma -> coalescence -> remove_zero, each checked on its own vector"""
    g = golden_methods
    fp = case(g, f"find_pairs_{tag}")
    length = int(fp["length"])
    n_sd = len(fp["idx"])
    flag = np.zeros(n_sd, dtype=np.bool_)
    be.find_pairs(fp["cell_start"], flag, fp["cell_id"], fp["cell_idx"], fp["idx"], length)
    np.testing.assert_array_equal(flag[:length], fp["flag"][:length])

    sw = case(g, f"sort_within_pair_{tag}")
    idx = sw["idx_in"].copy()
    be.sort_within_pair_by_attr(idx, length, flag, sw["mult"])
    np.testing.assert_array_equal(idx, sw["idx_out"])

    po = case(g, f"pair_ops_{tag}")
    out = np.full(n_sd // 2, np.nan)
    for op in ("distance", "max", "min", "sum", "multiply"):
        out[:] = np.nan
        be.pair_op(op, out, po["vol"], flag, idx, length)
        assert_same_f64(out, g[f"pair_{op}_{tag}/out"], op)
    out[:] = np.nan
    be.pair_op("max", out, po["mult"], flag, idx, length)
    assert_same_f64(out, g[f"pair_max_int_{tag}/out"], "max int")
    outn = np.full(n_sd, np.nan)
    be.sort_pair(outn, po["vol"], flag, idx, length)
    assert_same_f64(outn, po["sort_pair"], "sort_pair")

    nm = case(g, f"normalize_{tag}")
    prob = nm["prob_in"].copy()
    norm_factor = np.full(len(fp["cell_idx"]), np.nan)
    be.normalize(prob, fp["cell_id"], fp["cell_idx"], fp["cell_start"], norm_factor,
                  float(nm["timestep"]), float(nm["dv"]))
    assert_same_f64(norm_factor, nm["norm_factor"], "norm_factor")
    assert_same_f64(prob, nm["prob_out"], "normalize")

    for first in (1, 0):
        ad = case(g, f"adaptive_{tag}_{first}")
        prob_a = ad["prob_in"].copy()
        dt_left = ad["dt_left_in"].copy()
        n_substep = ad["n_substep_in"].copy()
        dt_min = ad["dt_min_in"].copy()
        be.scale_prob_for_adaptive_sdm_gamma(prob_a, idx, length, po["mult"], fp["cell_id"],
                                              dt_left, float(ad["dt"]), tuple(ad["dt_range"]),
                                              flag, n_substep, dt_min)
        assert_same_f64(prob_a, ad["prob_out"], "adaptive prob")
        assert_same_f64(dt_left, ad["dt_left_out"], "dt_left")
        assert_same_f64(dt_min, ad["dt_min_out"], "stats_dt_min")
        np.testing.assert_array_equal(n_substep, ad["n_substep_out"])
        assert be.adaptive_sdm_end(dt_left, fp["cell_start"]) == int(
            g[f"adaptive_end_{tag}_{first}/end"])

    gm = case(g, f"gamma_{tag}")
    n_cell = len(fp["cell_idx"])
    rate, deficit = np.zeros(n_cell, np.int64), np.zeros(n_cell, np.int64)
    be.compute_gamma(prob, gm["rand"], idx, length, po["mult"], fp["cell_id"], deficit, rate,
                      flag, prob)
    assert_same_f64(prob, gm["gamma"], "gamma")
    np.testing.assert_array_equal(rate, gm["rate"])
    np.testing.assert_array_equal(deficit, gm["deficit"])

    for n_attr in (1, 3):
        co = case(g, f"coalescence_{tag}_{n_attr}")
        attrs = co["attrs_in"].copy()
        mult = po["mult"].copy()
        healthy = np.ones(1, np.int64)
        coal = np.zeros(n_cell, np.int64)
        be.collision_coalescence(mult, idx, length, attrs, prob, healthy, fp["cell_id"], coal,
                                  flag)
        np.testing.assert_array_equal(mult, co["mult_out"])
        assert_same_f64(attrs, co["attrs_out"], "attributes")
        np.testing.assert_array_equal(healthy, co["healthy"])
        np.testing.assert_array_equal(coal, co["coal_rate"])
        rz = case(g, f"remove_zero_{tag}_{n_attr}")
        idx2 = idx.copy()
        new_length = be.remove_zero_n_or_flagged(mult, idx2, length)
        assert new_length == int(rz["new_length"])
        np.testing.assert_array_equal(idx2, rz["idx_out"])


@pytest.mark.parametrize("tag", TAGS)
def test_sort_by_key_and_cell_id(be, golden_methods, tag):
    c = case(golden_methods, f"sort_by_key_{tag}")
    out = np.zeros(len(c["keys"]), np.int64)
    be.sort_by_key(out, c["keys"])
    np.testing.assert_array_equal(out, c["out"])
    c = case(golden_methods, f"cell_id_{tag}")
    out = np.zeros(c["origin"].shape[1], np.int64)
    be.cell_id(out, np.ascontiguousarray(c["origin"], dtype=np.int64), c["strides"])
    np.testing.assert_array_equal(out, c["out"])


@pytest.mark.parametrize("n", (1, 2, 33, 1000, 2048, 2049, 12345, 32768, 65536, 65537, 300000))
def test_sort_by_key_sizes(be, n):
    """`argsort(kind="stable")[::-1]` (index_methods.py:46-48) at the sizes where the CUDA path
    switches algorithm (one-CTA sort up to 65536 keys, multi-kernel radix sort above), with
    heavy ties (dt_left is mostly 0 or dt), signed zeros, negatives, infinities and NaN"""
    rng = np.random.default_rng(n)
    keys = rng.choice([0.0, 5.0, -0.0, 2.5, np.inf, -np.inf, np.nan], n)
    some = rng.random(n) < 0.3
    keys[some] = rng.normal(0, 3, int(some.sum()))
    out = np.full(n, -1, np.int64)
    be.sort_by_key(out, keys)
    np.testing.assert_array_equal(out, np.argsort(keys, kind="stable")[::-1])


@pytest.mark.parametrize("shape", ("all_equal", "mostly_zero", "two_values", "low_bytes_differ",
                                   "one_nonzero", "all_distinct"))
@pytest.mark.parametrize("n", (7, 4097, 20001, 32768))
def test_sort_by_key_adaptive_shapes(be, n, shape):
    """the key sets the adaptive loop produces (collision.py:181, `dt_left` per cell): all equal
    before the first sub-step, mostly 0 afterwards -- the one-CTA sort treats the minimal key and
    constant digits specially"""
    rng = np.random.default_rng(n + len(shape))
    if shape == "all_equal":
        keys = np.full(n, 5.0)
    elif shape == "mostly_zero":
        keys = np.zeros(n)
        some = rng.random(n) < 0.03
        keys[some] = 5.0 - rng.random(int(some.sum())) * 4.9
    elif shape == "two_values":
        keys = rng.choice([0.0, 5.0], n)
    elif shape == "low_bytes_differ":
        keys = (1.0 + rng.integers(0, 1 << 14, n) * 2.0 ** -52)
    elif shape == "one_nonzero":
        keys = np.zeros(n)
        keys[n // 3] = 0.25
    else:
        keys = rng.random(n)
    out = np.full(n, -1, np.int64)
    be.sort_by_key(out, keys)
    np.testing.assert_array_equal(out, np.argsort(keys, kind="stable")[::-1])


# ---- whole time steps ---------------------------------------------------------------------
def make_oracle_from_scenario(s, name, be=orc):
    kw = dict(SCENARIO_KW[name])
    return OracleCollision(
        multiplicity=s["init/multiplicity"], water_mass=s["init/signed water mass"],
        cell_id=s["init/cell id"] if "init/cell id" in s.files and int(s["meta/n_cell"]) > 1
        else None,
        n_cell=int(s["meta/n_cell"]), dt=float(s["meta/dt"]), dv=float(s["meta/dv"]),
        kernel=Golovin(GOLOVIN_B), seed=int(s["meta/seed"]), rho_w=float(s["meta/rho_w"]), be=be, **kw)


def check_against_snapshot(sut, s, step):
    p = f"step{step}/"
    length = int(s[p + "length"])
    assert sut.length == length, f"step {step}: length"
    np.testing.assert_array_equal(sut.idx[:length], s[p + "idx"][:length], f"step {step} idx")
    np.testing.assert_array_equal(sut.multiplicity, s[p + "multiplicity"])
    assert_same_f64(sut.attributes, s[p + "attributes"], f"step {step} attributes")
    assert_same_f64(sut.gamma, s[p + "gamma"], f"step {step} gamma")
    for k in ("collision_rate", "collision_rate_deficit", "coalescence_rate", "stats_n_substep"):
        np.testing.assert_array_equal(getattr(sut, k), s[p + k], f"step {step} {k}")
    assert_same_f64(sut.stats_dt_min, s[p + "stats_dt_min"], "stats_dt_min")


@pytest.mark.parametrize("name", SCENARIOS)
def test_scenarios(be, name):
    s = load_scenario(name)
    sut = make_oracle_from_scenario(s, name, be)
    for step in range(int(s["meta/steps"])):
        sut()
        check_against_snapshot(sut, s, step)
