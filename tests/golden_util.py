"""helpers shared by the golden-vector tests"""
import os

import numpy as np

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
TAGS = ("a", "b", "c", "d")
SCENARIOS = ("golovin_local", "golovin_global", "golovin_substeps", "golovin_adaptive",
             "golovin_adaptive_optrand", "golovin_box")
SCENARIO_KW = {
    "golovin_local": {"adaptive": False, "croupier": "local"},
    "golovin_global": {"adaptive": False, "croupier": "global"},
    "golovin_substeps": {"adaptive": False, "substeps": 3},
    "golovin_adaptive": {"adaptive": True, "dt_coal_range": (0.01, 1.0)},
    "golovin_adaptive_optrand": {"adaptive": True, "dt_coal_range": (0.05, 1.0),
                                 "optimized_random": True},
    "golovin_box": {"adaptive": False},
}
GOLOVIN_B = 1.5e3


def case(g, name):
    """dict of the arrays stored under '<name>/' in methods.npz"""
    prefix = name + "/"
    return {k[len(prefix):]: g[k] for k in g.files if k.startswith(prefix)}


def load_scenario(name):
    return np.load(os.path.join(GOLDEN, f"scenario_{name}.npz"))


def assert_same_f64(actual, expected, what=""):
    """bit-exact fp64 comparison (NaN == NaN)"""
    a = np.ascontiguousarray(actual, dtype=np.float64).view(np.int64)
    e = np.ascontiguousarray(expected, dtype=np.float64).view(np.int64)
    nan_ok = np.isnan(np.asarray(actual, dtype=np.float64)) & np.isnan(
        np.asarray(expected, dtype=np.float64))
    bad = (a != e) & ~nan_ok
    assert not bad.any(), f"{what}: {int(bad.sum())} of {bad.size} differ, first at {np.argmax(bad)}"


def assert_attributes_close(actual, expected, what="", rtol=1e-12):
    """extensive attributes: north_star tolerance 1e-12 relative.  The reference's
    `coalesce`/`break_up` helpers are compiled with fastmath=True regardless of
    Formulae(fastmath=False) (impl_numba/conf.py:5-10 via collisions_methods.py:44),
    so LLVM contracts some a*b+c into FMA depending on loop unrolling: 1-ulp
    differences are compiler noise of the reference, not of the port."""
    np.testing.assert_allclose(actual, expected, rtol=rtol, atol=0, err_msg=what)
