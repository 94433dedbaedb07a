"""Floating-point "physics" rows of the collision path:
the C oracle (CPU) and, with `-m gpu`, the CUDA kernels through the C-ABI, both against
tests/golden/physics.npz -- vectors produced by RUNNING THE REFERENCE (PySDM's Numba backend,
Formulae(fastmath=False); generator: oracle/gen_golden_physics.py).

Bar: the oracle reproduces the reference bit for bit (exceptions are listed where they
occur: three scratch arrays written by helpers the reference always compiles with LLVM
fast-math).  The CUDA path uses CUDA's libm (pow/exp/log/erf/sinh/asinh/atanh are within
1-2 ulp of glibc's but not bit-identical), so it is held to the north_star's floating-point
tolerance, 1e-12 relative; integer outputs (multiplicities, rates, n_fragment of SLAMS) are
bit-exact for both."""

import numpy as np
import pytest

from oracle import sdm_oracle as orc

from .golden_util import GOLDEN, assert_same_f64, case

RTOL = 1e-12
BERRY = (1, 1, -27, 1.65, -58, 1.9, 15, 1.13, 16.7, 1, 0.004, 4, 8)  # hydrodynamic.py:11-13


@pytest.fixture(scope="module")
def g():
    import os  # pylint: disable=import-outside-toplevel

    return np.load(os.path.join(GOLDEN, "physics.npz"))


def same(be, got, expected, what, rtol=RTOL, atol=0.0, exact=True):
    """bit-exact for the oracle (unless exact=False), `rtol` for the CUDA path"""
    if be is orc and exact:
        assert_same_f64(got, expected, what)
    else:
        np.testing.assert_allclose(got, expected, rtol=rtol, atol=atol, equal_nan=True,
                                   err_msg=what)


@pytest.mark.parametrize("tag", ("third", "two_thirds", "two", "minus_two", "half", "p2_83"))
def test_signed_power(be, g, tag):
    c = case(g, "pow_" + tag)
    x = c["x"].copy()
    be.power(x, float(c["e"]))
    same(be, x, c["out"], "**= " + tag)


def test_radius_of_volume(be, g):
    # attributes/physics/radius.py:15-17 (quirk A12: pow(., 1/3), not cbrt)
    c = case(g, "radius")
    out = np.empty_like(c["volume"])
    be.product_scalar(out, c["volume"], float(c["factor"]))
    be.power(out, 1 / 3)
    same(be, out, c["out"], "radius")


def test_interpolation_gunn_kinzer(be, g):
    c = case(g, "interpolation")
    out = np.full_like(c["radius"], np.nan)
    be.interpolation(out, c["radius"], float(c["factor"]), c["a"], c["b"])
    same(be, out, c["out"], "interpolation")


def test_terminal_velocity_rogers_yau(be, g):
    c = case(g, "rogers_yau")
    out = np.full_like(c["radius"], np.nan)
    be.terminal_velocity_rogers_yau(out, c["radius"], c["params"])
    same(be, out, c["out"], "Rogers-Yau")


def test_power_series(be, g):
    c = case(g, "power_series")
    out = np.full_like(c["radius"], np.nan)
    be.power_series(out, c["radius"], c["prefactors"], c["powers"])
    same(be, out, c["out"], "power_series")


@pytest.mark.parametrize("tag", "abc")
def test_linear_collection_efficiency(be, g, tag):
    c = case(g, "lce_" + tag)
    out = np.full_like(c["out"], np.nan)
    be.linear_collection_efficiency(c["params"], orc.int_mask_of(BERRY), out, c["radii"],
                                    c["flag"], c["idx"], len(c["idx"]), float(c["unit"]))
    # E is a sum of terms of both signs clamped at 0: absolute floor for the CUDA libm
    same(be, out, c["out"], "linear_collection_efficiency", rtol=1e-11, atol=1e-13)


def _limits(c):
    return float(c["vmin"]), (None if np.isnan(c["nfmax"]) else int(c["nfmax"]))


@pytest.mark.parametrize("tag", ("lim", "nolim"))
def test_simple_fragmentations(be, g, tag):
    """exponential, Gaussian, Feingold 1988, SLAMS (+ limiters)"""
    com = case(g, "frag_common")
    x_plus_y, rand = com["x_plus_y"], com["rand"]
    n = len(rand)
    c = case(g, "frag_exp_" + tag)
    vmin, nfmax = _limits(c)

    def fresh():
        return np.full(n, np.nan), np.full(n, np.nan)

    nf, fv = fresh()
    be.exp_fragmentation(nf, float(c["scale"]), fv, x_plus_y, rand, vmin, nfmax)
    same(be, fv, c["frag_volume"], "exp frag_volume")
    same(be, nf, c["n_fragment"], "exp n_fragment")

    c = case(g, "frag_gauss_" + tag)
    nf, fv = fresh()
    be.gauss_fragmentation(nf, float(c["mu"]), float(c["sigma"]), fv, x_plus_y, rand, vmin, nfmax)
    same(be, fv, c["frag_volume"], "gauss frag_volume")
    same(be, nf, c["n_fragment"], "gauss n_fragment")

    c = case(g, "frag_feingold_" + tag)
    nf, fv = fresh()
    be.feingold1988_fragmentation(nf, float(c["scale"]), fv, x_plus_y, rand, float(c["fragtol"]),
                                  vmin, nfmax)
    same(be, fv, c["frag_volume"], "feingold frag_volume")
    same(be, nf, c["n_fragment"], "feingold n_fragment")

    c = case(g, "frag_slams_" + tag)
    nf, fv = fresh()
    probs = np.full(n, np.nan)
    be.slams_fragmentation(nf, fv, x_plus_y, probs, rand, vmin, nfmax)
    same(be, probs, c["probs"], "slams probs")
    same(be, fv, c["frag_volume"], "slams frag_volume")
    same(be, nf, c["n_fragment"], "slams n_fragment")


@pytest.mark.parametrize("tag", ("lim", "nolim"))
def test_straub_fragmentation(be, g, tag):
    c = case(g, "frag_straub_" + tag)
    rand = case(g, "frag_common")["rand"]
    n = len(rand)
    vmin, nfmax = _limits(case(g, "frag_exp_" + tag))
    nf, fv = np.full(n, np.nan), np.full(n, np.nan)
    tmp = {k: np.zeros(n) for k in ("Nr1", "Nr2", "Nr3", "Nr4", "Nrt")}
    tmp["d34"] = np.full(n, np.nan)
    be.straub_fragmentation(nf, c["CW"], c["gam"], c["ds"], fv, c["v_max"], c["x_plus_y"], rand,
                            vmin, nfmax, **tmp)
    same(be, fv, c["frag_volume"], "straub frag_volume", rtol=1e-11)
    same(be, nf, c["n_fragment"], "straub n_fragment", rtol=1e-11)
    same(be, tmp["Nr1"], c["Nr1"], "Nr1")
    same(be, tmp["Nr2"], c["Nr2"], "Nr2")
    # scratch arrays of the reference's fast-math helpers (straub_mass_remainder,
    # conf.JIT_FLAGS): ~1 % of the entries differ from the reference in the last bit
    for k in ("Nr3", "Nr4", "Nrt", "d34"):
        same(be, tmp[k], c[k], k, rtol=1e-11, atol=1e-25, exact=False)
        if be is orc:
            exact = np.mean(tmp[k].view(np.int64) == c[k].view(np.int64))
            assert exact > 0.98, (k, exact)


@pytest.mark.parametrize("tag", ("lim", "nolim"))
def test_ll82_fragmentation(be, g, tag):
    c = case(g, "frag_ll82_" + tag)
    rand = case(g, "frag_common")["rand"].copy()
    n = len(rand)
    vmin, nfmax = _limits(case(g, "frag_exp_" + tag))
    nf, fv = np.full(n, np.nan), np.full(n, np.nan)
    tmp = {k: np.zeros(n) for k in ("Rf", "Rs", "Rd")}
    be.ll82_fragmentation(nf, c["CKE"], c["W"], c["W2"], c["St"], c["ds"], c["dl"], c["dcoal"], fv,
                          c["x_plus_y"], rand, vmin, nfmax, **tmp)
    # the inverse-CDF sampling amplifies libm differences (erfinv near +-1, 10 fixed-point
    # iterations on erf): looser bound for the CUDA path
    same(be, fv, c["frag_volume"], "ll82 frag_volume", rtol=1e-9)
    same(be, nf, c["n_fragment"], "ll82 n_fragment", rtol=1e-9)
    same(be, rand, c["rand_out"], "ll82 rand (updated in place)", rtol=1e-9, atol=1e-15)
    same(be, tmp["Rf"], c["Rf"], "Rf")
    same(be, tmp["Rs"], c["Rs"], "Rs", rtol=1e-10, atol=1e-15)
    same(be, tmp["Rd"], c["Rd"], "Rd", rtol=1e-10, atol=1e-15)


def test_ll82_coalescence_check(be, g):
    c = case(g, "ll82_check")
    Ec = c["Ec_in"].copy()
    be.ll82_coalescence_check(Ec, c["dl"])
    assert_same_f64(Ec, c["Ec_out"], "Ec")


@pytest.mark.parametrize("tag", "abc")
@pytest.mark.parametrize("handle_all_breakups", (0, 1))
def test_collision_coalescence_breakup(be, g, handle_all_breakups, tag):
    """break_up / break_up_while (collisions_methods.py:62-311): multiplicities, rates,
    deficits and `healthy` bit-exact; attributes within 1e-12 (the reference's helpers are
    LLVM fast-math code, see golden_util.assert_attributes_close; the oracle matches all
    but ~0.5 % of them to the bit)"""
    c = case(g, f"breakup_{handle_all_breakups}_{tag}")
    mult, attrs = c["mult_in"].copy(), c["attrs_in"].copy()
    n_cell = len(c["coalescence_rate"])
    rates = [np.zeros(n_cell, np.int64) for _ in range(3)]
    healthy = np.ones(1, np.int64)
    be.collision_coalescence_breakup(
        mult, c["idx"], len(c["idx"]), attrs, c["gamma"], c["rand"], c["Ec"], c["Eb"],
        c["fragment_mass"], healthy, c["cell_id"], rates[0], rates[1], rates[2], c["flag"],
        float(c["max_multiplicity"]), c["particle_mass"], bool(handle_all_breakups))
    np.testing.assert_array_equal(mult, c["mult_out"])
    np.testing.assert_array_equal(rates[0], c["coalescence_rate"])
    np.testing.assert_array_equal(rates[1], c["breakup_rate"])
    np.testing.assert_array_equal(rates[2], c["breakup_rate_deficit"])
    np.testing.assert_array_equal(healthy, c["healthy"])
    np.testing.assert_allclose(attrs, c["attrs_out"], rtol=RTOL, atol=0)
    if be is orc:
        assert np.mean(attrs.view(np.int64) == c["attrs_out"].view(np.int64)) > 0.98
