"""GPU drop-in test: the UNMODIFIED reference orchestration (Builder / Particulator /
Coalescence / ParticleAttributes / Index / PairIndicator ... from PySDM itself) driven
once by the reference's Numba backend and once by `pysdm_b200.B200`, same attributes, same
uniform random numbers (HostReplayRandom = the Numba backend's PCG64 stream), compared
after every step: permutation, pair flags, gamma, multiplicities, counters bit-exact,
attributes bit-exact too for coalescence.

Needs the reference importable: baseline/_ref (pip --target install of /root/reference,
git-ignored, travels to the GPU box) or /root/reference; skipped otherwise.
"""

import os
import sys

import numpy as np
import pytest

from .golden_util import assert_same_f64

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_REF = next((p for p in (os.path.join(ROOT, "baseline", "_ref"), "/root/reference")
             if os.path.isdir(os.path.join(p, "PySDM"))), None)


@pytest.fixture(scope="module")
def pysdm():
    if _REF is None:
        pytest.skip("reference PySDM not available (baseline/_ref)")
    os.environ.setdefault("NUMBA_NUM_THREADS", "1")
    os.environ.setdefault("CI", "1")  # the reference's own switch: no ThrustRTC import (not installed)
    sys.path.insert(0, os.path.join(ROOT, "oracle", "refshims"))
    sys.path.insert(1, _REF)
    np.NaN = np.nan  # the reference predates NumPy 2
    import PySDM  # pylint: disable=import-outside-toplevel

    return PySDM


def build(pysdm, backend, *, n_sd, grid, dv, dt, attributes, dynamic_kw, kernel_b=1.5e3,
          fused_dynamic=False):
    from PySDM import Builder  # pylint: disable=import-outside-toplevel
    from PySDM.dynamics import Coalescence  # pylint: disable=import-outside-toplevel
    if fused_dynamic:
        from pysdm_b200 import FusedCoalescence as Coalescence  # pylint: disable=import-outside-toplevel
    from PySDM.dynamics.collisions.collision_kernels import Golovin  # pylint: disable=import-outside-toplevel
    from PySDM.environments import Box  # pylint: disable=import-outside-toplevel
    from PySDM.impl.mesh import Mesh  # pylint: disable=import-outside-toplevel

    env = Box(dv=dv, dt=dt)
    n_cell = int(np.prod(grid))
    if n_cell > 1:
        env.mesh = Mesh(grid, size=grid, dv=dv)
    builder = Builder(n_sd=n_sd, backend=backend, environment=env)
    builder.add_dynamic(Coalescence(collision_kernel=Golovin(b=kernel_b), **dynamic_kw))
    attrs = {k: np.array(v) for k, v in attributes.items()}
    if n_cell == 1:
        attrs.pop("cell id", None)
    return builder.build(attributes=attrs, products=())


def snapshot(particulator):
    dyn = particulator.dynamics["Collision"]
    attrs = particulator.attributes
    idx = attrs._ParticleAttributes__idx  # pylint: disable=protected-access

    def host(storage):
        return np.asarray(storage.to_ndarray())

    return {
        "length": int(idx.length),
        "idx": host(idx),
        "multiplicity": attrs["multiplicity"].to_ndarray(raw=True),
        "attributes": attrs.get_extensive_attribute_storage().to_ndarray(raw=True),
        "flag": host(dyn.is_first_in_pair.indicator),
        "gamma": host(dyn.gamma),
        "collision_rate": host(dyn.collision_rate),
        "collision_rate_deficit": host(dyn.collision_rate_deficit),
        "coalescence_rate": host(dyn.coalescence_rate),
        "stats_n_substep": host(dyn.stats_n_substep),
        "stats_dt_min": host(dyn.stats_dt_min),
    }


CASES = {
    "local": dict(grid=(4, 5), n_sd=1500, dv=2e-8, kw=dict(adaptive=False, croupier="local")),
    "global": dict(grid=(3, 3), n_sd=700, dv=2e-8, kw=dict(adaptive=False, croupier="global")),
    "substeps": dict(grid=(3, 3), n_sd=700, dv=2e-8, kw=dict(adaptive=False, substeps=2)),
    "adaptive": dict(grid=(3, 3), n_sd=700, dv=4e-9,
                     kw=dict(adaptive=True, dt_coal_range=(0.01, 1.0))),
    "adaptive_optrand": dict(grid=(3, 3), n_sd=700, dv=4e-9,
                             kw=dict(adaptive=True, dt_coal_range=(0.05, 1.0),
                                     optimized_random=True)),
    "box": dict(grid=(1,), n_sd=1024, dv=1e-7, kw=dict(adaptive=False)),
    # one cell beyond the shared-memory batches of the fused step: shuffle replay + streaming kernel
    "box_stream": dict(grid=(1,), n_sd=20000, dv=2e-5, kw=dict(adaptive=False)),
    "box_stream_adaptive": dict(grid=(1,), n_sd=20000, dv=4e-6,
                                kw=dict(adaptive=True, dt_coal_range=(0.02, 1.0))),
}


@pytest.mark.parametrize("case", sorted(CASES))
def test_b200_backend_equals_numba_backend(pysdm, case):
    from PySDM import Formulae  # pylint: disable=import-outside-toplevel
    from PySDM.backends import CPU  # pylint: disable=import-outside-toplevel

    from pysdm_b200 import B200  # pylint: disable=import-outside-toplevel

    c = CASES[case]
    n_sd, n_cell = c["n_sd"], int(np.prod(c["grid"]))
    rng = np.random.default_rng(99)
    attributes = {
        "multiplicity": rng.integers(1, 5, n_sd).astype(np.int64),
        "volume": rng.exponential(4.0 / 3 * np.pi * (30.531e-6) ** 3, n_sd),
        "cell id": rng.integers(0, n_cell, n_sd).astype(np.int64),
    }
    seed = 1234
    common = dict(n_sd=n_sd, grid=c["grid"], dv=c["dv"], dt=1.0, attributes=attributes,
                  dynamic_kw=c["kw"])
    from pysdm_b200.recorder import RECORDER  # pylint: disable=import-outside-toplevel

    ref = build(pysdm, CPU(Formulae(fastmath=False, seed=seed)), **common)
    sut = build(pysdm, B200(Formulae(fastmath=False, seed=seed), host_rng=True), **common)
    fused0, replayed0 = RECORDER.n_fused, RECORDER.n_replayed
    collided = 0
    for step in range(8):
        ref.run(1)
        sut.run(1)
        a, b = snapshot(sut), snapshot(ref)
        n = b["length"]
        assert a["length"] == n, f"step {step}: length"
        np.testing.assert_array_equal(a["idx"][:n], b["idx"][:n], f"step {step}: idx")
        np.testing.assert_array_equal(a["flag"][:n], b["flag"][:n], f"step {step}: flag")
        np.testing.assert_array_equal(a["multiplicity"], b["multiplicity"])
        assert_same_f64(a["attributes"], b["attributes"], f"step {step}: attributes")
        assert_same_f64(a["gamma"], b["gamma"], f"step {step}: gamma")
        assert_same_f64(a["stats_dt_min"], b["stats_dt_min"], f"step {step}: stats_dt_min")
        for k in ("collision_rate", "collision_rate_deficit", "coalescence_rate",
                  "stats_n_substep"):
            np.testing.assert_array_equal(a[k], b[k], f"step {step}: {k}")
        collided = int(b["coalescence_rate"].sum())
    assert collided > 0
    # the stock dynamic's backend calls are recorded and run as one fused launch per step when
    # they form a step the fused kernel covers (pysdm_b200/recorder.py); the adaptive sequence
    # (a host read of stats_dt_min in its middle) and the global croupier are replayed / eager
    if case in ("local", "substeps", "box", "box_stream"):
        assert RECORDER.n_fused - fused0 == 8 * c["kw"].get("substeps", 1)
        assert RECORDER.n_replayed == replayed0
    else:
        assert RECORDER.n_fused == fused0


@pytest.mark.parametrize("host_rng", (True, False))
@pytest.mark.parametrize("case", ("local", "substeps", "box", "global", "adaptive",
                                  "adaptive_optrand", "box_stream", "box_stream_adaptive"))
def test_fused_coalescence_dynamic(pysdm, case, host_rng):
    """`pysdm_b200.FusedCoalescence` (PySDM's Coalescence with step() on the one-launch fused
    kernel) inside the unmodified PySDM orchestration.  host_rng=True: against the reference's
    Numba backend with the stock dynamic, same uniforms.  host_rng=False: against the stock
    dynamic on the B200 backend (method-level kernels), same device Philox stream.  Everything
    bit-identical after every step; the "global" croupier is not covered by the fused kernel
    and must fall through to the reference's step()."""
    from PySDM import Formulae  # pylint: disable=import-outside-toplevel
    from PySDM.backends import CPU  # pylint: disable=import-outside-toplevel

    from pysdm_b200 import B200  # pylint: disable=import-outside-toplevel

    c = CASES[case]
    n_sd, n_cell = c["n_sd"], int(np.prod(c["grid"]))
    rng = np.random.default_rng(98)
    attributes = {
        "multiplicity": rng.integers(1, 5, n_sd).astype(np.int64),
        "volume": rng.exponential(4.0 / 3 * np.pi * (30.531e-6) ** 3, n_sd),
        "cell id": rng.integers(0, n_cell, n_sd).astype(np.int64),
    }
    common = dict(n_sd=n_sd, grid=c["grid"], dv=c["dv"], dt=1.0, attributes=attributes,
                  dynamic_kw=c["kw"])
    formulae = dict(fastmath=False, seed=4321)
    ref_backend = CPU(Formulae(**formulae)) if host_rng else B200(Formulae(**formulae))
    ref = build(pysdm, ref_backend, **common)
    sut = build(pysdm, B200(Formulae(**formulae), host_rng=host_rng), fused_dynamic=True, **common)
    for step in range(8):
        ref.run(1)
        sut.run(1)
        a, b = snapshot(sut), snapshot(ref)
        n = b["length"]
        assert a["length"] == n, f"step {step}: length"
        np.testing.assert_array_equal(a["idx"][:n], b["idx"][:n], f"step {step}: idx")
        np.testing.assert_array_equal(a["flag"][:n], b["flag"][:n], f"step {step}: flag")
        np.testing.assert_array_equal(a["multiplicity"], b["multiplicity"])
        assert_same_f64(a["attributes"], b["attributes"], f"step {step}: attributes")
        assert_same_f64(a["gamma"], b["gamma"], f"step {step}: gamma")
        for k in ("collision_rate", "collision_rate_deficit", "coalescence_rate",
                  "stats_n_substep"):
            np.testing.assert_array_equal(a[k], b[k], f"step {step}: {k}")
        assert_same_f64(a["stats_dt_min"], b["stats_dt_min"], f"step {step}: stats_dt_min")
        # derived attributes are recomputed from what the fused step wrote
        np.testing.assert_array_equal(sut.attributes["volume"].to_ndarray(),
                                      ref.attributes["volume"].to_ndarray())
    dyn = sut.dynamics["Collision"]
    if c["kw"].get("adaptive"):  # one fused launch per sub-step of the adaptive loop
        assert dyn.n_fused_steps >= int(snapshot(ref)["stats_n_substep"].max()) > 8
    else:
        expected = 0 if case == "global" else 8 * c["kw"].get("substeps", 1)
        assert dyn.n_fused_steps == expected
    assert int(snapshot(ref)["coalescence_rate"].sum()) > 0


# ---- configs 2 and 3 of BASELINE.json in miniature: gravitational kernels with hydrodynamic
# efficiency, adaptive time step, breakup with the fragmentation functions ---------------------
def build_physics(pysdm, backend, *, n_sd, dv, dt, attributes, dynamic):
    from PySDM import Builder  # pylint: disable=import-outside-toplevel
    from PySDM.environments import Box  # pylint: disable=import-outside-toplevel

    builder = Builder(n_sd=n_sd, backend=backend, environment=Box(dv=dv, dt=dt))
    builder.add_dynamic(dynamic())
    return builder.build(attributes={k: np.array(v) for k, v in attributes.items()}, products=())


def _dynamics(name, fused=False):
    # pylint: disable=import-outside-toplevel
    from PySDM.dynamics import Coalescence, Collision
    if fused:
        from pysdm_b200 import FusedCoalescence as Coalescence, FusedCollision as Collision
    from PySDM.dynamics.collisions.breakup_efficiencies import ConstEb
    from PySDM.dynamics.collisions.breakup_fragmentations import (AlwaysN, Exponential,
                                                                  Feingold1988, Gaussian,
                                                                  LowList1982Nf, SLAMS,
                                                                  Straub2010Nf)
    from PySDM.dynamics.collisions.coalescence_efficiencies import (Berry1967, ConstEc,
                                                                    LowList1982Ec, Straub2010Ec)
    from PySDM.dynamics.collisions.collision_kernels import Geometric, Hydrodynamic
    from PySDM.physics import si

    vmin = (0.01 * si.mm) ** 3 * np.pi / 6
    return {
        # config 2: Berry 1967 box
        "hydrodynamic_adaptive": lambda: Coalescence(
            collision_kernel=Hydrodynamic(), adaptive=True, dt_coal_range=(0.1, 100.0)),
        "geometric": lambda: Coalescence(collision_kernel=Geometric(), adaptive=False),
        # config 3: de Jong et al. breakup box
        "breakup_straub": lambda: Collision(
            collision_kernel=Geometric(), coalescence_efficiency=Straub2010Ec(),
            breakup_efficiency=ConstEb(1.0),
            fragmentation_function=Straub2010Nf(vmin=vmin, nfmax=10000), adaptive=True,
            warn_overflows=False),
        "breakup_berry_straub": lambda: Collision(
            collision_kernel=Geometric(), coalescence_efficiency=Berry1967(),
            breakup_efficiency=ConstEb(1.0),
            fragmentation_function=Straub2010Nf(vmin=vmin, nfmax=10000), adaptive=True,
            warn_overflows=False),
        "breakup_ll82": lambda: Collision(
            collision_kernel=Geometric(), coalescence_efficiency=LowList1982Ec(),
            breakup_efficiency=ConstEb(1.0),
            fragmentation_function=LowList1982Nf(vmin=vmin, nfmax=10000), adaptive=True,
            warn_overflows=False),
        "breakup_always_n": lambda: Collision(
            collision_kernel=Geometric(), coalescence_efficiency=ConstEc(0.5),
            breakup_efficiency=ConstEb(1.0), fragmentation_function=AlwaysN(n=4),
            adaptive=False, warn_overflows=False),
        "breakup_exponential": lambda: Collision(
            collision_kernel=Geometric(), coalescence_efficiency=ConstEc(0.5),
            breakup_efficiency=ConstEb(0.9),
            fragmentation_function=Exponential(scale=4 / 3 * np.pi * (100 * si.um) ** 3,
                                               vmin=vmin, nfmax=100),
            adaptive=False, warn_overflows=False),
        "breakup_gaussian": lambda: Collision(
            collision_kernel=Geometric(), coalescence_efficiency=ConstEc(0.5),
            breakup_efficiency=ConstEb(1.0),
            fragmentation_function=Gaussian(mu=4 / 3 * np.pi * (100 * si.um) ** 3,
                                            sigma=4 / 3 * np.pi * (50 * si.um) ** 3,
                                            vmin=vmin, nfmax=100),
            adaptive=False, warn_overflows=False),
        "breakup_feingold": lambda: Collision(
            collision_kernel=Geometric(), coalescence_efficiency=ConstEc(0.5),
            breakup_efficiency=ConstEb(1.0),
            fragmentation_function=Feingold1988(scale=4 / 3 * np.pi * (100 * si.um) ** 3,
                                                fragtol=1e-3, vmin=vmin, nfmax=100),
            adaptive=False, warn_overflows=False),
        "breakup_slams": lambda: Collision(
            collision_kernel=Geometric(), coalescence_efficiency=ConstEc(0.5),
            breakup_efficiency=ConstEb(1.0), fragmentation_function=SLAMS(vmin=vmin, nfmax=100),
            adaptive=False, warn_overflows=False),
    }[name]


FRAG_FUNCTION = {"breakup_straub": "Straub2010Nf", "breakup_berry_straub": "Straub2010Nf",
                 "breakup_ll82": "LowList1982Nf", "breakup_feingold": "Feingold1988"}
PHYSICS_CASES = ("hydrodynamic_adaptive", "geometric", "breakup_straub", "breakup_berry_straub",
                 "breakup_ll82", "breakup_always_n", "breakup_exponential", "breakup_gaussian",
                 "breakup_feingold", "breakup_slams")


FUSED_BREAKUP = ("breakup_straub", "breakup_berry_straub", "breakup_always_n")


@pytest.mark.parametrize("fused", (False, True))
@pytest.mark.parametrize("case", PHYSICS_CASES)
def test_b200_backend_physics_cases(pysdm, case, fused):
    """Same protocol as above for the gravitational kernels / breakup.  The collision
    probability now goes through pow/exp/log of CUDA's libm, which is not bit-identical to
    glibc's: multiplicities, pair flags, permutation, gamma and the rate counters must still
    be identical (a difference would be an fp64 gamma within an ulp of an integer boundary,
    north_star; it is counted and reported as the test failure), attributes to 1e-12."""
    from PySDM import Formulae  # pylint: disable=import-outside-toplevel
    from PySDM.backends import CPU  # pylint: disable=import-outside-toplevel

    from pysdm_b200 import B200  # pylint: disable=import-outside-toplevel

    n_sd = 2048
    rng = np.random.default_rng(7)
    radius = rng.exponential(0.12e-3, n_sd) + 20e-6 if case.startswith("breakup") else \
        rng.exponential(12e-6, n_sd) + 2e-6
    attributes = {
        "multiplicity": rng.integers(*((5, 200) if case.startswith("breakup") else (200, 4000)), n_sd).astype(np.int64),
        "volume": 4 / 3 * np.pi * radius ** 3,
    }
    fkw = dict(fastmath=False, seed=4321)
    if case in FRAG_FUNCTION:
        fkw["fragmentation_function"] = FRAG_FUNCTION[case]
    dv, dt = (1.0, 1.0) if case.startswith("breakup") else (0.05, 10.0)
    common = dict(n_sd=n_sd, dv=dv, dt=dt, attributes=attributes)
    ref = build_physics(pysdm, CPU(Formulae(**fkw)), dynamic=_dynamics(case), **common)
    # fused=True: pysdm_b200.FusedCoalescence / FusedCollision -- the step (and, for the
    # efficiencies / fragmentation functions the fused breakup kernel covers, its breakup half)
    # in one launch each; anything else must fall through to the reference's step()
    sut = build_physics(pysdm, B200(Formulae(**fkw), host_rng=True),
                        dynamic=_dynamics(case, fused), **common)
    events = 0
    for step in range(6):
        ref.run(1)
        sut.run(1)
        a, b = snapshot(sut), snapshot(ref)
        n = b["length"]
        assert a["length"] == n, f"step {step}: length"
        np.testing.assert_array_equal(a["idx"][:n], b["idx"][:n], f"step {step}: idx")
        np.testing.assert_array_equal(a["flag"][:n], b["flag"][:n], f"step {step}: flag")
        flips = int((a["gamma"] != b["gamma"]).sum())
        assert flips == 0, f"step {step}: {flips} gamma values differ (integer-boundary flips)"
        np.testing.assert_array_equal(a["multiplicity"], b["multiplicity"], f"step {step}")
        np.testing.assert_allclose(a["attributes"], b["attributes"], rtol=1e-12, atol=0,
                                   err_msg=f"step {step}: attributes")
        for k in ("collision_rate", "collision_rate_deficit", "coalescence_rate",
                  "stats_n_substep"):
            np.testing.assert_array_equal(a[k], b[k], f"step {step}: {k}")
        dyn_a, dyn_b = sut.dynamics["Collision"], ref.dynamics["Collision"]
        if case.startswith("breakup"):
            for k in ("breakup_rate", "breakup_rate_deficit"):
                np.testing.assert_array_equal(getattr(dyn_a, k).to_ndarray(),
                                              getattr(dyn_b, k).to_ndarray(), f"step {step}: {k}")
            events += int(dyn_b.breakup_rate.to_ndarray().sum())
        events += int(b["coalescence_rate"].sum())
    assert events > 0
    if fused:
        n_fused = sut.dynamics["Collision"].n_fused_steps
        if case in FUSED_BREAKUP or not case.startswith("breakup"):
            assert n_fused >= 6, f"{case}: the fused step was expected to run"
        else:
            assert n_fused == 0, f"{case}: not covered by the fused breakup kernel"


# ---- the dynamics next to the collision step that touch the same SoA state --
def test_seeding_dynamic_with_coalescence(pysdm):
    """`PySDM.dynamics.Seeding` (dynamics/seeding.py:91-117 -> particulator.py:447-490 ->
    backend.seeding) next to `Coalescence` in a box padded with empty slots, after the
    reference's own tests/unit_tests/dynamics/test_seeding.py: Numba backend vs B200 backend,
    same uniforms; the injected slots, the shuffled seed index and the whole state are
    bit-identical after every step."""
    from PySDM import Builder, Formulae  # pylint: disable=import-outside-toplevel
    from PySDM.backends import CPU  # pylint: disable=import-outside-toplevel
    from PySDM.dynamics import Coalescence, Seeding  # pylint: disable=import-outside-toplevel
    from PySDM.dynamics.collisions.collision_kernels import Golovin  # pylint: disable=import-outside-toplevel
    from PySDM.environments import Box  # pylint: disable=import-outside-toplevel

    from pysdm_b200 import B200  # pylint: disable=import-outside-toplevel

    n_initial, n_pad, n_seeds = 600, 200, 16
    rng = np.random.default_rng(5)
    volume = np.pad(rng.exponential(4.0 / 3 * np.pi * (30.531e-6) ** 3, n_initial), (0, n_pad))
    multiplicity = np.pad(rng.integers(1, 5, n_initial).astype(float), (0, n_pad),
                          constant_values=np.nan)
    seed_mass = list(rng.exponential(1e-9, n_seeds))
    seed_mult = list(rng.integers(1, 9, n_seeds))

    def make(backend):
        builder = Builder(n_sd=n_initial + n_pad, backend=backend, environment=Box(dt=1.0, dv=1e-5))
        builder.add_dynamic(Coalescence(collision_kernel=Golovin(b=1.5e3), adaptive=False))
        builder.add_dynamic(Seeding(
            super_droplet_injection_rate=lambda time: 3 if 2 <= time < 7 else 0,
            seeded_particle_multiplicity=seed_mult,
            seeded_particle_extensive_attributes={"signed water mass": seed_mass}))
        return builder.build(attributes={"volume": volume.copy(),
                                         "multiplicity": multiplicity.copy()}, products=())

    fkw = dict(fastmath=False, seed=77)
    ref, sut = make(CPU(Formulae(**fkw))), make(B200(Formulae(**fkw), host_rng=True))
    counts = []
    for step in range(10):
        ref.run(1)
        sut.run(1)
        a, b = snapshot(sut), snapshot(ref)
        n = b["length"]
        assert a["length"] == n, f"step {step}: length"
        np.testing.assert_array_equal(a["idx"][:n], b["idx"][:n], f"step {step}: idx")
        np.testing.assert_array_equal(a["multiplicity"], b["multiplicity"], f"step {step}")
        assert_same_f64(a["attributes"], b["attributes"], f"step {step}: attributes")
        np.testing.assert_array_equal(sut.dynamics["Seeding"].index.to_ndarray(),
                                      ref.dynamics["Seeding"].index.to_ndarray())
        counts.append(int(ref.attributes.super_droplet_count))
        assert int(sut.attributes.super_droplet_count) == counts[-1]
    assert max(np.diff(counts)) > 0  # injections did happen
    assert int(snapshot(ref)["coalescence_rate"].sum()) > 0


@pytest.mark.parametrize("constant", (True, False))
def test_relaxed_velocity_dynamic(pysdm, constant):
    """`PySDM.dynamics.RelaxedVelocity` (dynamics/relaxed_velocity.py:60-90: Storage `product`,
    `*=`, `+=`, `**=`, `exp` on the SoA state + the `relative fall momentum` / `terminal
    velocity` attributes) on the B200 backend vs the Numba backend, after the reference's
    tests/unit_tests/dynamics/test_relaxed_velocity.py; 1e-12 relative (CUDA pow/exp)."""
    from PySDM import Builder, Formulae  # pylint: disable=import-outside-toplevel
    from PySDM.backends import CPU  # pylint: disable=import-outside-toplevel
    from PySDM.dynamics import RelaxedVelocity  # pylint: disable=import-outside-toplevel
    from PySDM.environments import Box  # pylint: disable=import-outside-toplevel

    from pysdm_b200 import B200  # pylint: disable=import-outside-toplevel

    n_sd = 512
    rng = np.random.default_rng(11)
    attributes = {
        "multiplicity": rng.integers(1, 100, n_sd).astype(np.int64),
        "volume": 4 / 3 * np.pi * (rng.exponential(0.2e-3, n_sd) + 5e-6) ** 3,
        "relative fall momentum": np.zeros(n_sd),
    }

    def make(backend):
        builder = Builder(n_sd=n_sd, backend=backend, environment=Box(dt=1.0, dv=1.0))
        builder.add_dynamic(RelaxedVelocity(c=2.0, constant=constant))
        builder.request_attribute("relative fall velocity")
        builder.request_attribute("terminal velocity")
        return builder.build(attributes={k: v.copy() for k, v in attributes.items()}, products=())

    ref, sut = make(CPU(Formulae(fastmath=False))), make(B200(Formulae(fastmath=False)))
    for step in range(4):
        ref.run(1)
        sut.run(1)
        for name in ("relative fall momentum", "relative fall velocity", "terminal velocity"):
            expected = ref.attributes[name].to_ndarray()
            np.testing.assert_allclose(sut.attributes[name].to_ndarray(), expected, rtol=1e-12,
                                       atol=0, err_msg=f"step {step}: {name}")
            assert np.abs(expected).max() > 0


@pytest.mark.parametrize("host_rng", (True, False))
@pytest.mark.parametrize("grid", ((1,), (4, 5)))
@pytest.mark.parametrize("kernel", ("Geometric", "Hydrodynamic"))
def test_fused_coalescence_geometric(pysdm, kernel, grid, host_rng):
    """`FusedCoalescence` with the gravitational kernels -- Geometric, collision_kernels/
    geometric.py:15-22, and Hydrodynamic (Berry's collection efficiency, impl/parameterized.py:
    13-30): the fused kernel gathers the orchestration's own "radius" and "relative
    fall velocity" attributes.  host_rng=False: against the stock dynamic on the B200 backend,
    same device Philox stream -- everything bit-identical.  host_rng=True: against the Numba
    backend with the same uniforms -- integers / permutation / gamma identical, attributes to
    1e-12 (radius goes through CUDA's pow there, glibc's here)."""
    from PySDM import Builder, Formulae  # pylint: disable=import-outside-toplevel
    from PySDM.backends import CPU  # pylint: disable=import-outside-toplevel
    from PySDM.dynamics import Coalescence  # pylint: disable=import-outside-toplevel
    from PySDM.dynamics.collisions.collision_kernels import Geometric, Hydrodynamic  # pylint: disable=import-outside-toplevel
    from PySDM.environments import Box  # pylint: disable=import-outside-toplevel
    from PySDM.impl.mesh import Mesh  # pylint: disable=import-outside-toplevel

    from pysdm_b200 import B200, FusedCoalescence  # pylint: disable=import-outside-toplevel

    n_sd, n_cell = 2048, int(np.prod(grid))
    rng = np.random.default_rng(17)
    attributes = {
        "multiplicity": rng.integers(200, 4000, n_sd).astype(np.int64),
        "volume": 4 / 3 * np.pi * (rng.exponential(12e-6, n_sd) + 2e-6) ** 3,
    }
    if n_cell > 1:
        attributes["cell id"] = rng.integers(0, n_cell, n_sd).astype(np.int64)

    def make(backend, dynamic):
        env = Box(dv=0.05 / n_cell, dt=10.0)
        if n_cell > 1:
            env.mesh = Mesh(grid, size=grid, dv=0.05 / n_cell)
        builder = Builder(n_sd=n_sd, backend=backend, environment=env)
        collision_kernel = Geometric(collection_efficiency=0.9) if kernel == "Geometric" \
            else Hydrodynamic()
        builder.add_dynamic(dynamic(collision_kernel=collision_kernel, adaptive=False,
                                    croupier="local"))
        return builder.build(attributes={k: v.copy() for k, v in attributes.items()}, products=())

    fkw = dict(fastmath=False, seed=31)
    ref = make(CPU(Formulae(**fkw)) if host_rng else B200(Formulae(**fkw)), Coalescence)
    sut = make(B200(Formulae(**fkw), host_rng=host_rng), FusedCoalescence)
    for step in range(6):
        ref.run(1)
        sut.run(1)
        a, b = snapshot(sut), snapshot(ref)
        n = b["length"]
        assert a["length"] == n, f"step {step}: length"
        np.testing.assert_array_equal(a["idx"][:n], b["idx"][:n], f"step {step}: idx")
        np.testing.assert_array_equal(a["flag"][:n], b["flag"][:n], f"step {step}: flag")
        flips = int((a["gamma"] != b["gamma"]).sum())
        assert flips == 0, f"step {step}: {flips} gamma values differ"
        np.testing.assert_array_equal(a["multiplicity"], b["multiplicity"], f"step {step}")
        if host_rng:
            np.testing.assert_allclose(a["attributes"], b["attributes"], rtol=1e-12, atol=0)
        else:
            assert_same_f64(a["attributes"], b["attributes"], f"step {step}: attributes")
        for k in ("collision_rate", "collision_rate_deficit", "coalescence_rate"):
            np.testing.assert_array_equal(a[k], b[k], f"step {step}: {k}")
        # the derived attributes: the fused step refreshes them for the super-droplets a
        # coalescence changed and stamps them as current; the stock path recomputes all of them
        for name in ("volume", "radius", "relative fall velocity"):
            x, y = sut.attributes[name].to_ndarray(), ref.attributes[name].to_ndarray()
            if host_rng:
                np.testing.assert_allclose(x, y, rtol=1e-12, atol=0, err_msg=name)
            else:
                assert_same_f64(x, y, f"step {step}: {name}")
    assert sut.dynamics["Collision"].n_fused_steps == 6
    assert sut.dynamics["Collision"]._dependents is not None  # pylint: disable=protected-access
    assert int(snapshot(ref)["coalescence_rate"].sum()) > 0


# ---- the remaining collision kernels through the stock dynamic -------------------
def _kernel(name):
    # pylint: disable=import-outside-toplevel
    from PySDM.dynamics.collisions.collision_kernels import (ConstantK, Electric, Linear,
                                                             SimpleGeometric)

    return {"ConstantK": lambda: ConstantK(a=2e-11), "Linear": lambda: Linear(a=1e-10, b=1.5e3),
            "Electric": Electric, "SimpleGeometric": lambda: SimpleGeometric(C=3e5)}[name]()


@pytest.mark.parametrize("fused", (False, True))
@pytest.mark.parametrize("kernel", ("ConstantK", "Electric", "SimpleGeometric", "Linear"))
def test_other_collision_kernels_equal_numba(pysdm, kernel, fused):
    """collision_kernels/{constantK,electric,simple_geometric,linear}.py through the unmodified
    orchestration: B200 (stock `Coalescence`, and `FusedCoalescence`, which covers ConstantK and
    -- as a Parameterized kernel -- Electric) against the Numba backend, same uniforms.
    The reference's `Linear` is broken upstream (`output.sum_pair` does not exist on
    PairwiseStorage, linear.py:13; its own header says TODO #744): both backends must fail the
    same way."""
    from PySDM import Builder, Formulae  # pylint: disable=import-outside-toplevel
    from PySDM.backends import CPU  # pylint: disable=import-outside-toplevel
    from PySDM.dynamics import Coalescence  # pylint: disable=import-outside-toplevel
    from PySDM.environments import Box  # pylint: disable=import-outside-toplevel
    from PySDM.impl.mesh import Mesh  # pylint: disable=import-outside-toplevel

    from pysdm_b200 import B200, FusedCoalescence  # pylint: disable=import-outside-toplevel

    n_sd, grid = 4096, (4, 4)
    rng = np.random.default_rng(23)
    attributes = {
        "multiplicity": rng.integers(200, 4000, n_sd).astype(np.int64),
        "volume": 4 / 3 * np.pi * (rng.exponential(12e-6, n_sd) + 2e-6) ** 3,
        "cell id": rng.integers(0, 16, n_sd).astype(np.int64),
    }

    def make(backend, dynamic):
        dv = {"ConstantK": 1e-5, "Electric": 0.04}.get(kernel, 0.004)
        env = Box(dv=dv, dt=10.0)
        env.mesh = Mesh(grid, size=grid, dv=dv)
        builder = Builder(n_sd=n_sd, backend=backend, environment=env)
        builder.add_dynamic(dynamic(collision_kernel=_kernel(kernel), adaptive=False))
        return builder.build(attributes={k: v.copy() for k, v in attributes.items()}, products=())

    fkw = dict(fastmath=False, seed=57)
    ref = make(CPU(Formulae(**fkw)), Coalescence)
    sut = make(B200(Formulae(**fkw), host_rng=True), FusedCoalescence if fused else Coalescence)
    if kernel == "Linear":
        with pytest.raises(AttributeError, match="sum_pair"):
            ref.run(1)
        if fused:  # the fused step evaluates the kernel itself (sum, *= b, += a; pinned on the
            # oracle in test_gpu_step.py::test_other_kernels_against_oracle) and does run
            sut.run(5)
            assert sut.dynamics["Collision"].n_fused_steps == 5
            assert int(snapshot(sut)["coalescence_rate"].sum()) > 0
        else:
            with pytest.raises(AttributeError, match="sum_pair"):
                sut.run(1)
        return
    for step in range(5):
        ref.run(1)
        sut.run(1)
        a, b = snapshot(sut), snapshot(ref)
        n = b["length"]
        assert a["length"] == n, f"step {step}: length"
        np.testing.assert_array_equal(a["idx"][:n], b["idx"][:n], f"step {step}: idx")
        np.testing.assert_array_equal(a["flag"][:n], b["flag"][:n], f"step {step}: flag")
        flips = int((a["gamma"] != b["gamma"]).sum())
        assert flips == 0, f"step {step}: {flips} gamma values differ (integer-boundary flips)"
        np.testing.assert_array_equal(a["multiplicity"], b["multiplicity"], f"step {step}")
        np.testing.assert_allclose(a["attributes"], b["attributes"], rtol=1e-12, atol=0)
        for k in ("collision_rate", "collision_rate_deficit", "coalescence_rate"):
            np.testing.assert_array_equal(a[k], b[k], f"step {step}: {k}")
    assert int(snapshot(ref)["coalescence_rate"].sum()) > 0
    if fused:
        expected = 5 if kernel in ("ConstantK", "Electric") else 0
        assert sut.dynamics["Collision"].n_fused_steps == expected


# ---- BASELINE.json configs at (or in the kind of) their stated sizes ------------------------------
def _compare_runs(ref, suts, steps, *, breakup=False, fused_expected=()):
    events = 0
    for step in range(steps):
        ref.run(1)
        b = snapshot(ref)
        n = b["length"]
        for name, sut in suts.items():
            sut.run(1)
            a = snapshot(sut)
            what = f"{name} step {step}"
            assert a["length"] == n, f"{what}: length"
            np.testing.assert_array_equal(a["idx"][:n], b["idx"][:n], f"{what}: idx")
            np.testing.assert_array_equal(a["flag"][:n], b["flag"][:n], f"{what}: flag")
            flips = int((a["gamma"] != b["gamma"]).sum())
            print(f"{what}: gamma integer-boundary flips = {flips} of {a['gamma'].size}")
            assert flips == 0, f"{what}: {flips} gamma values differ (integer-boundary flips)"
            np.testing.assert_array_equal(a["multiplicity"], b["multiplicity"], what)
            np.testing.assert_allclose(a["attributes"], b["attributes"], rtol=1e-12, atol=0,
                                       err_msg=what)
            for k in ("collision_rate", "collision_rate_deficit", "coalescence_rate",
                      "stats_n_substep"):
                np.testing.assert_array_equal(a[k], b[k], f"{what}: {k}")
            assert_same_f64(a["stats_dt_min"], b["stats_dt_min"], f"{what}: stats_dt_min")
            if breakup:
                for k in ("breakup_rate", "breakup_rate_deficit"):
                    np.testing.assert_array_equal(
                        getattr(sut.dynamics["Collision"], k).to_ndarray(),
                        getattr(ref.dynamics["Collision"], k).to_ndarray(), f"{what}: {k}")
        events += int(b["coalescence_rate"].sum())
    assert events > 0
    for name in fused_expected:
        assert suts[name].dynamics["Collision"].n_fused_steps > 0, name


def test_config4_in_kind(pysdm):
    """BASELINE configs[3] (Sc-rain, Arabas et al. 2015) in kind: a 2-D mesh of 32 x 32 cells
    with 128 super-droplets each, `Coalescence(Geometric(1), adaptive=True,
    optimized_random=True)` (Morrison_and_Grabowski_2007/common.py:30-35), A = 3 extensive
    attributes ("dry volume", "kappa times dry volume" next to the water mass,
    environments/kinematic_2d.py:63-64,84), dt = 5 s, dv of one 1.5 km / 512 cell: Numba
    backend vs B200 with the stock dynamic and with `FusedCoalescence` (fused adaptive
    sub-steps), same uniforms"""
    from PySDM import Builder, Formulae  # pylint: disable=import-outside-toplevel
    from PySDM.backends import CPU  # pylint: disable=import-outside-toplevel
    from PySDM.dynamics import Coalescence  # pylint: disable=import-outside-toplevel
    from PySDM.dynamics.collisions.collision_kernels import Geometric  # pylint: disable=import-outside-toplevel
    from PySDM.environments import Box  # pylint: disable=import-outside-toplevel
    from PySDM.impl.mesh import Mesh  # pylint: disable=import-outside-toplevel

    from pysdm_b200 import B200, FusedCoalescence  # pylint: disable=import-outside-toplevel

    grid, per_cell = (32, 32), 128
    n_cell = grid[0] * grid[1]
    n_sd = n_cell * per_cell
    dv = (1500 / 512) ** 2
    rng = np.random.default_rng(2015)
    volume = 4 / 3 * np.pi * (10e-6 * np.exp(0.47 * rng.standard_normal(n_sd))) ** 3
    volume[rng.random(n_sd) < 0.01] *= 60.0  # a few drizzle drops: cells that need sub-steps
    attributes = {
        "volume": volume, "cell id": np.arange(n_sd, dtype=np.int64) // per_cell,
        "multiplicity": rng.integers(1, 2 * int(100e6 * dv / per_cell), n_sd).astype(np.int64),
        "dry volume": volume * 1e-3, "kappa times dry volume": volume * 0.6e-3}

    def make(backend, dynamic):
        env = Box(dv=dv, dt=5.0)
        env.mesh = Mesh(grid, size=grid, dv=dv)
        builder = Builder(n_sd=n_sd, backend=backend, environment=env)
        builder.add_dynamic(dynamic(collision_kernel=Geometric(collection_efficiency=1),
                                    adaptive=True, optimized_random=True))
        builder.request_attribute("dry volume")
        builder.request_attribute("kappa times dry volume")
        return builder.build(attributes={k: v.copy() for k, v in attributes.items()}, products=())

    fkw = dict(fastmath=False, seed=15)
    ref = make(CPU(Formulae(**fkw)), Coalescence)
    assert len(ref.attributes.get_extensive_attribute_keys()) == 3
    suts = {"stock": make(B200(Formulae(**fkw), host_rng=True), Coalescence),
            "fused": make(B200(Formulae(**fkw), host_rng=True), FusedCoalescence)}
    _compare_runs(ref, suts, 4, fused_expected=("fused",))
    assert int(snapshot(ref)["stats_n_substep"].max()) > 4, "no cell needed a sub-step"


def test_config2_full_size(pysdm):
    """BASELINE configs[1] at its size: Berry 1967 box, 2^20 super-droplets in one cell,
    `Coalescence(Hydrodynamic(), adaptive=True, dt_coal_range=(0.1, 100))`, dt = 10 s, dv = 10 m^3
    (Berry_1967/settings.py:25-38), B200 vs Numba, same uniforms, 3 steps"""
    from PySDM import Formulae  # pylint: disable=import-outside-toplevel
    from PySDM.backends import CPU  # pylint: disable=import-outside-toplevel

    from pysdm_b200 import B200  # pylint: disable=import-outside-toplevel

    n_sd = 1 << 20
    rng = np.random.default_rng(1967)
    volume = rng.exponential(4 / 3 * np.pi * 10e-6 ** 3, n_sd)
    attributes = {"volume": volume,
                  "multiplicity": np.full(n_sd, int(239e6 * 10 / n_sd), dtype=np.int64)}
    fkw = dict(fastmath=False, seed=67)
    common = dict(n_sd=n_sd, dv=10.0, dt=10.0, attributes=attributes)
    ref = build_physics(pysdm, CPU(Formulae(**fkw)), dynamic=_dynamics("hydrodynamic_adaptive"),
                        **common)
    suts = {"stock": build_physics(pysdm, B200(Formulae(**fkw), host_rng=True),
                                   dynamic=_dynamics("hydrodynamic_adaptive"), **common),
            "fused": build_physics(pysdm, B200(Formulae(**fkw), host_rng=True),
                                   dynamic=_dynamics("hydrodynamic_adaptive", fused=True), **common)}
    _compare_runs(ref, suts, 3, fused_expected=("fused",))


def test_config3_full_size(pysdm):
    """BASELINE configs[2] at its size: de Jong et al. breakup box, 2^22 super-droplets,
    `Collision(Geometric(), Straub2010Ec(), ConstEb(1), Straub2010Nf(...), adaptive=True)`,
    dt = 1 s, dv = 10^6 m^3 (deJong_Mackay_et_al_2023/settings_0D.py:27-52), B200 vs Numba,
    same uniforms, 2 steps"""
    from PySDM import Formulae  # pylint: disable=import-outside-toplevel
    from PySDM.backends import CPU  # pylint: disable=import-outside-toplevel

    from pysdm_b200 import B200  # pylint: disable=import-outside-toplevel

    n_sd = 1 << 22
    rng = np.random.default_rng(2023)
    volume = rng.exponential(4 / 3 * np.pi * 0.2e-3 ** 3, n_sd)
    attributes = {"volume": volume,
                  "multiplicity": np.full(n_sd, max(1, int(1e3 * 1e6 / n_sd)), dtype=np.int64)}
    fkw = dict(fastmath=False, seed=23, fragmentation_function="Straub2010Nf")
    common = dict(n_sd=n_sd, dv=1e6, dt=1.0, attributes=attributes)
    ref = build_physics(pysdm, CPU(Formulae(**fkw)), dynamic=_dynamics("breakup_straub"), **common)
    suts = {"stock": build_physics(pysdm, B200(Formulae(**fkw), host_rng=True),
                                   dynamic=_dynamics("breakup_straub"), **common),
            "fused": build_physics(pysdm, B200(Formulae(**fkw), host_rng=True),
                                   dynamic=_dynamics("breakup_straub", fused=True), **common)}
    _compare_runs(ref, suts, 2, breakup=True, fused_expected=("fused",))


@pytest.mark.parametrize("kernel", ("Golovin", "ConstantK", "Geometric", "Hydrodynamic",
                                    "SimpleGeometric"))
def test_recorded_stock_dynamic_equals_eager(pysdm, kernel):
    """`Builder(backend=B200())` with the STOCK `Coalescence`: the recording backend
    (pysdm_b200/recorder.py: one fused launch per step, uniforms generated in the kernel, the
    two `Random` fills elided) against the same backend executing every call at once
    (`B200(defer=False)`), same device Philox stream -- everything bit-identical after every
    step.  SimpleGeometric is a sequence the fused kernel does not cover: the record must be
    replayed call by call."""
    from PySDM import Builder, Formulae  # pylint: disable=import-outside-toplevel
    from PySDM.dynamics import Coalescence  # pylint: disable=import-outside-toplevel
    from PySDM.dynamics.collisions.collision_kernels import (  # pylint: disable=import-outside-toplevel
        ConstantK, Geometric, Golovin, Hydrodynamic, SimpleGeometric)
    from PySDM.environments import Box  # pylint: disable=import-outside-toplevel
    from PySDM.impl.mesh import Mesh  # pylint: disable=import-outside-toplevel

    from pysdm_b200 import B200  # pylint: disable=import-outside-toplevel
    from pysdm_b200.recorder import RECORDER  # pylint: disable=import-outside-toplevel

    n_sd, grid = 6000, (5, 6)
    rng = np.random.default_rng(41)
    attributes = {
        "multiplicity": rng.integers(1, 40, n_sd).astype(np.int64),
        "volume": 4 / 3 * np.pi * (rng.exponential(12e-6, n_sd) + 2e-6) ** 3,
        "cell id": rng.integers(0, 30, n_sd).astype(np.int64),
    }
    make_kernel = {"Golovin": lambda: Golovin(b=1.5e3), "ConstantK": lambda: ConstantK(a=3e-11),
                   "Geometric": lambda: Geometric(collection_efficiency=0.9),
                   "Hydrodynamic": Hydrodynamic, "SimpleGeometric": lambda: SimpleGeometric(C=3e5)}
    dv = {"Golovin": 1e-7, "ConstantK": 1e-6}.get(kernel, 2e-5)

    def make(backend):
        env = Box(dv=dv, dt=10.0)
        env.mesh = Mesh(grid, size=grid, dv=dv)
        builder = Builder(n_sd=n_sd, backend=backend, environment=env)
        builder.add_dynamic(Coalescence(collision_kernel=make_kernel[kernel](), adaptive=False,
                                        substeps=2))
        return builder.build(attributes={k: v.copy() for k, v in attributes.items()}, products=())

    fkw = dict(fastmath=False, seed=91)
    ref = make(B200(Formulae(**fkw), defer=False))
    sut = make(B200(Formulae(**fkw)))
    fused0, replayed0 = RECORDER.n_fused, RECORDER.n_replayed
    for step in range(5):
        ref.run(1)
        sut.run(1)
        a, b = snapshot(sut), snapshot(ref)
        n = b["length"]
        assert a["length"] == n, f"step {step}: length"
        np.testing.assert_array_equal(a["idx"][:n], b["idx"][:n], f"step {step}: idx")
        np.testing.assert_array_equal(a["flag"][:n], b["flag"][:n], f"step {step}: flag")
        np.testing.assert_array_equal(a["multiplicity"], b["multiplicity"], f"step {step}")
        assert_same_f64(a["attributes"], b["attributes"], f"step {step}: attributes")
        assert_same_f64(a["gamma"], b["gamma"], f"step {step}: gamma")
        for k in ("collision_rate", "collision_rate_deficit", "coalescence_rate"):
            np.testing.assert_array_equal(a[k], b[k], f"step {step}: {k}")
        # the uniform arrays of a fused step were never written; reading them fills them
        dyn_a, dyn_b = sut.dynamics["Collision"], ref.dynamics["Collision"]
        if step == 2:
            assert_same_f64(dyn_a.rnd_opt_coll.pairs_rand.to_ndarray(),
                            dyn_b.rnd_opt_coll.pairs_rand.to_ndarray(), "pairs_rand")
            assert_same_f64(dyn_a.rnd_opt_coll.rand.to_ndarray(),
                            dyn_b.rnd_opt_coll.rand.to_ndarray(), "rand")
    assert int(snapshot(ref)["coalescence_rate"].sum()) > 0
    if kernel == "SimpleGeometric":
        assert RECORDER.n_fused == fused0 and RECORDER.n_replayed > replayed0
    else:
        assert RECORDER.n_fused - fused0 == 10 and RECORDER.n_replayed == replayed0


def test_vtk_exporter_attributes_equal_numba(pysdm, tmp_path, monkeypatch):
    """SURVEY 8(f)4, exporters: the reference's `VTKExporter.export_attributes`
    (PySDM/exporters/vtk_exporter.py:62-110) run UNCHANGED over a particulator on the B200
    backend -- it downloads every attribute with `to_ndarray(raw=True)` -- hands the writer
    the same payload as over the Numba backend after the same coalescence steps (pyevtk is not
    in the image: its `pointsToVTK` is replaced by a recorder of what it is given).
    `export_products` / the netCDF exporter write products, which are outside the path (8f)."""
    from PySDM import Builder, Formulae  # pylint: disable=import-outside-toplevel
    from PySDM.backends import CPU  # pylint: disable=import-outside-toplevel
    from PySDM.dynamics import Coalescence  # pylint: disable=import-outside-toplevel
    from PySDM.dynamics.collisions.collision_kernels import Golovin  # pylint: disable=import-outside-toplevel
    from PySDM.environments import Box  # pylint: disable=import-outside-toplevel
    from PySDM.exporters import vtk_exporter  # pylint: disable=import-outside-toplevel
    from PySDM.impl.mesh import Mesh  # pylint: disable=import-outside-toplevel

    from pysdm_b200 import B200, FusedCoalescence  # pylint: disable=import-outside-toplevel

    grid, per_cell = (6, 5), 64
    n_sd = grid[0] * grid[1] * per_cell
    rng = np.random.default_rng(62)
    cell = np.arange(n_sd, dtype=np.int64) // per_cell
    attributes = {
        "volume": rng.exponential(4 / 3 * np.pi * 30.531e-6 ** 3, n_sd),
        "multiplicity": rng.integers(1, 2 ** 22, n_sd).astype(np.int64),
        "cell id": cell, "cell origin": np.stack([cell // grid[1], cell % grid[1]]),
        "position in cell": rng.uniform(0, 1, (2, n_sd))}

    written = []
    monkeypatch.setattr(vtk_exporter, "pointsToVTK",
                        lambda path, x, y, z, data: written.append((path, x, y, z, dict(data))))

    def export(backend, dynamic):
        env = Box(dv=1e6, dt=1.0)
        env.mesh = Mesh(grid, size=(60.0, 50.0), dv=1e6)
        builder = Builder(n_sd=n_sd, backend=backend, environment=env)
        builder.add_dynamic(dynamic(collision_kernel=Golovin(b=1.5e3)))
        particulator = builder.build(attributes={k: v.copy() for k, v in attributes.items()},
                                     products=())
        exporter = vtk_exporter.VTKExporter(path=str(tmp_path))
        out = []
        for _ in range(2):
            particulator.run(2)
            exporter.export_attributes(particulator)
            out.append(written.pop())
        assert not written
        return out

    fkw = dict(fastmath=False, seed=62)
    ref = export(CPU(Formulae(**fkw)), Coalescence)
    for dynamic in (Coalescence, FusedCoalescence):
        sut = export(B200(Formulae(**fkw), host_rng=True), dynamic)
        for (path_r, *xyz_r, data_r), (path_s, *xyz_s, data_s) in zip(ref, sut):
            assert path_s == path_r and sorted(data_s) == sorted(data_r)
            assert {"multiplicity", "cell id", "cell origin[0]", "position in cell[1]",
                    "water mass"} <= set(data_s)
            for a, b in zip(xyz_r, xyz_s):
                np.testing.assert_array_equal(a, b)
            for key, expected in data_r.items():
                got = data_s[key]
                assert got.dtype == expected.dtype and got.shape == expected.shape, key
                assert got.flags["C_CONTIGUOUS"] or got.flags["F_CONTIGUOUS"], key
                if expected.dtype.kind == "f":  # north_star: attributes within 1e-12 relative
                    np.testing.assert_allclose(got, expected, rtol=1e-12, atol=0, err_msg=key)
                else:
                    np.testing.assert_array_equal(got, expected, err_msg=key)
