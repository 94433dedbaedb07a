"""`seeding` (particle injection into empty slots): the C oracle (CPU) and,
with `-m gpu`, the CUDA kernels through the C-ABI, against tests/golden/seeding.npz -- vectors
produced by RUNNING THE REFERENCE (`SeedingMethods.seeding` of PySDM's Numba backend,
backends/impl_numba/methods/seeding_methods.py:12-65; generator oracle/gen_golden_seeding.py).

Bar: bit-exact (index / integer work and verbatim copies of fp64 attributes, NaN included)."""

import os

import numpy as np
import pytest

from .golden_util import GOLDEN, assert_same_f64, case

CASES = ("few", "all_seeds", "all_slots", "none", "one", "big")


@pytest.fixture(scope="module")
def g():
    return np.load(os.path.join(GOLDEN, "seeding.npz"))


def run(be, c, n_inject):
    idx, mult, attrs = c["idx"].copy(), c["multiplicity"].copy(), c["attributes"].copy()
    be.seeding(idx, mult, attrs, c["seed_index"], c["seed_multiplicity"], c["seed_attributes"],
               n_inject)
    return idx, mult, attrs


@pytest.mark.parametrize("tag", CASES)
def test_seeding(be, g, tag):
    c = case(g, f"seeding_{tag}")
    n_inject = int(c["n_inject"])
    idx, mult, attrs = run(be, c, n_inject)
    np.testing.assert_array_equal(idx, c["idx_out"])
    np.testing.assert_array_equal(mult, c["multiplicity_out"])
    assert_same_f64(attrs, c["attributes_out"], "extensive attributes")
    # the vectors do exercise the method: injected slots are marked -1, the rest untouched
    assert int((c["idx_out"] == -1).sum()) == n_inject
    assert int((c["multiplicity"] == 0).sum()) - int((c["multiplicity_out"] == 0).sum()) == n_inject


def test_seeding_more_than_empty_slots_raises(be, g):
    """the reference asserts that as many super-droplets were injected as asked for
    (seeding_methods.py:40-43)"""
    c = case(g, "seeding_one")
    with pytest.raises(AssertionError):
        run(be, c, 1 + int((c["multiplicity"] == 0).sum()))
