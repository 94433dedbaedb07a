"""GPU: the device-resident collision step (pysdm_b200.collision.DeviceCollision), in its
method-level and fused modes, against (a) the state snapshots the reference itself
produced (tests/golden/scenario_*.npz) and (b) the CPU oracle on larger seeded inputs.
Bit-exact: permutation, pair flags, gamma (sign of zero included), multiplicities,
counters and -- for coalescence -- the attributes too."""

import os

import numpy as np
import pytest

from oracle.sdm_step import Golovin as OGolovin
from oracle.sdm_step import OracleCollision

from .golden_util import (GOLOVIN_B, SCENARIO_KW, SCENARIOS, assert_same_f64, load_scenario)

pytestmark = pytest.mark.gpu


def device_sim(**kw):
    from pysdm_b200.collision import DeviceCollision  # pylint: disable=import-outside-toplevel

    return DeviceCollision(**kw)


def device_kernel(name="Golovin", **kw):
    from pysdm_b200 import collision  # pylint: disable=import-outside-toplevel

    return getattr(collision, name)(**kw)


def compare_states(dev, ora, what):
    assert dev.length == ora.length, f"{what}: length"
    n = ora.length
    np.testing.assert_array_equal(dev.host("idx")[:n], ora.idx[:n], f"{what}: idx")
    np.testing.assert_array_equal(dev.host("flag")[:n], ora.flag[:n], f"{what}: flag")
    np.testing.assert_array_equal(dev.host("multiplicity"), ora.multiplicity, f"{what}: n")
    assert_same_f64(dev.host("attributes"), ora.attributes, f"{what}: attributes")
    assert_same_f64(dev.host("gamma"), ora.gamma, f"{what}: gamma")
    assert_same_f64(dev.host("norm_factor"), ora.norm_factor, f"{what}: norm_factor")
    for k in ("collision_rate", "collision_rate_deficit", "coalescence_rate", "stats_n_substep",
              "cell_start"):
        np.testing.assert_array_equal(dev.host(k), getattr(ora, k), f"{what}: {k}")
    assert_same_f64(dev.host("stats_dt_min"), ora.stats_dt_min, f"{what}: stats_dt_min")


@pytest.mark.parametrize("fused", ("auto", False))
@pytest.mark.parametrize("name", SCENARIOS)
def test_reference_scenarios(name, fused):
    s = load_scenario(name)
    kw = dict(SCENARIO_KW[name])
    multi = "init/cell id" in s.files and int(s["meta/n_cell"]) > 1
    sut = device_sim(
        multiplicity=s["init/multiplicity"], water_mass=s["init/signed water mass"],
        cell_id=s["init/cell id"] if multi else None, n_cell=int(s["meta/n_cell"]),
        dt=float(s["meta/dt"]), dv=float(s["meta/dv"]), kernel=device_kernel(b=GOLOVIN_B),
        seed=int(s["meta/seed"]), rho_w=float(s["meta/rho_w"]), host_rng=True, fused=fused, **kw)
    for step in range(int(s["meta/steps"])):
        sut()
        p = f"step{step}/"
        length = int(s[p + "length"])
        assert sut.length == length, f"step {step}: length"
        np.testing.assert_array_equal(sut.host("idx")[:length], s[p + "idx"][:length])
        np.testing.assert_array_equal(sut.host("flag")[:length], s[p + "flag"][:length])
        np.testing.assert_array_equal(sut.host("multiplicity"), s[p + "multiplicity"])
        assert_same_f64(sut.host("attributes"), s[p + "attributes"], f"step {step} attributes")
        assert_same_f64(sut.host("gamma"), s[p + "gamma"], f"step {step} gamma")
        for k in ("collision_rate", "collision_rate_deficit", "coalescence_rate",
                  "stats_n_substep"):
            np.testing.assert_array_equal(sut.host(k), s[p + k], f"step {step} {k}")
    if fused == "auto" and name in ("golovin_local", "golovin_substeps", "golovin_box"):
        assert sut.n_fused_steps > 0, "the fused kernel was expected to run"


def ragged_population(seed, n_cell, mean, n_attr=1, max_mult=4):
    rng = np.random.default_rng(seed)
    sizes = rng.poisson(mean, n_cell)
    if n_cell >= 5:
        sizes[rng.integers(0, n_cell, max(1, n_cell // 10))] = 0   # empty cells
        sizes[rng.integers(0, n_cell, max(1, n_cell // 10))] = 1   # lonely SDs
        sizes[rng.integers(0, n_cell, max(1, n_cell // 10))] = 3   # odd cells
    cell_id = np.repeat(np.arange(n_cell), sizes)
    n_sd = len(cell_id)
    perm = rng.permutation(n_sd)  # SDs are NOT stored cell-contiguously
    cell_id = cell_id[perm]
    mult = rng.integers(1, max_mult + 1, n_sd).astype(np.int64)
    mass = rng.exponential(1000 * 4 / 3 * np.pi * (30.531e-6) ** 3, n_sd)
    extra = [rng.exponential(1e-18, n_sd) for _ in range(n_attr - 1)]
    return {"multiplicity": mult, "water_mass": mass, "cell_id": cell_id, "n_cell": n_cell,
            "extra_attributes": extra}


CASES = {
    "ragged_small_cells": dict(seed=1, n_cell=700, mean=14, n_attr=1, dv=2e-8, steps=6),
    "ragged_128": dict(seed=2, n_cell=300, mean=128, n_attr=3, dv=4e-7, steps=5),
    "big_cells": dict(seed=3, n_cell=5, mean=3000, n_attr=1, dv=2e-5, steps=3),
    "substeps": dict(seed=4, n_cell=64, mean=40, n_attr=2, dv=1e-7, steps=4, substeps=3),
    "one_cell": dict(seed=5, n_cell=1, mean=5000, n_attr=1, dv=2e-6, steps=3),
    # cells beyond the shared-memory batches (> 8192 SDs; boxes): shuffle replay + streaming kernel
    "box_stream": dict(seed=12, n_cell=1, mean=20000, n_attr=2, dv=3e-5, steps=3),
    "big_cells_stream": dict(seed=13, n_cell=6, mean=11000, n_attr=1, dv=2e-5, steps=3),
    "box_stream_substeps": dict(seed=14, n_cell=1, mean=9000, n_attr=1, dv=1e-5, steps=2,
                                substeps=2),
    "adaptive_box_stream": dict(seed=15, n_cell=1, mean=20000, n_attr=1, dv=3e-6, steps=3,
                                adaptive=True, dt_coal_range=(0.02, 1.0)),
    "adaptive_big_stream": dict(seed=16, n_cell=6, mean=11000, n_attr=2, dv=2e-6, steps=3,
                                adaptive=True, dt_coal_range=(0.05, 1.0), optimized_random=True,
                                max_mult=2),
    # adaptive time step (one fused launch per sub-step): ragged cells so that the re-sorted
    # cell_idx shifts the pair parity (quirk A2) and `normalize` picks other cells' factors (A1)
    "adaptive_ragged": dict(seed=6, n_cell=300, mean=40, n_attr=3, dv=1e-8, steps=4,
                            adaptive=True, dt_coal_range=(0.01, 1.0)),
    "adaptive_optrand": dict(seed=7, n_cell=120, mean=60, n_attr=2, dv=2e-8, steps=4,
                             adaptive=True, dt_coal_range=(0.05, 1.0), optimized_random=True),
    # small multiplicities: SDs are removed inside the sub-step loop, the re-sort by cell then
    # uses the dt_left-sorted cell_idx (ranks are no longer cell ids)
    "adaptive_removals": dict(seed=8, n_cell=200, mean=50, n_attr=1, dv=1e-8, steps=5,
                              adaptive=True, dt_coal_range=(0.02, 1.0), max_mult=2),
    "adaptive_big_cells": dict(seed=9, n_cell=6, mean=2500, n_attr=1, dv=3e-7, steps=3,
                               adaptive=True, dt_coal_range=(0.05, 1.0), max_mult=3),
    # stats_dt_min reset to a finite value (as the CollisionTimestepMin product does,
    # products/collision/collision_timestep_min.py:21-24): finished cells keep updating it
    "adaptive_stats_reset": dict(seed=11, n_cell=150, mean=40, n_attr=1, dv=1e-8, steps=4,
                                 adaptive=True, dt_coal_range=(0.01, 1.0), stats_reset=True),
    "adaptive_dt_min": dict(seed=10, n_cell=50, mean=30, n_attr=1, dv=2e-10, steps=2,
                            adaptive=True, dt_coal_range=(0.25, 1.0)),
}


@pytest.mark.parametrize("case", sorted(CASES))
@pytest.mark.parametrize("fused", (True, False))
def test_against_oracle(case, fused):
    c = dict(CASES[case])
    steps, dv = c.pop("steps"), c.pop("dv")
    substeps = c.pop("substeps", 1)
    extra = {k: c.pop(k) for k in ("adaptive", "dt_coal_range", "optimized_random") if k in c}
    stats_reset = c.pop("stats_reset", False)
    pop = ragged_population(**c)
    if c["n_cell"] == 1:
        pop["cell_id"] = np.zeros_like(pop["cell_id"])
    common = dict(dt=1.0, dv=dv, seed=100 + c["seed"], substeps=substeps, **extra)
    ora = OracleCollision(kernel=OGolovin(GOLOVIN_B), **pop, **common)
    dev = device_sim(kernel=device_kernel(b=GOLOVIN_B), host_rng=True, fused=fused, **pop,
                     **common)
    for step in range(steps):
        if stats_reset:
            ora.stats_dt_min[:] = 1.0
            dev.stats_dt_min.fill_(1.0)
        ora()
        dev()
        compare_states(dev, ora, f"{case} step {step}")
    assert ora.coalescence_rate.sum() > 0, "test population never collided"
    if stats_reset:
        assert np.isfinite(ora.stats_dt_min).all() and ora.stats_dt_min.min() < 1.0
    if extra.get("adaptive"):
        assert ora.stats_n_substep.max() > steps, "the time step was never cut"
        np.testing.assert_array_equal(dev.host("dt_left"), ora.dt_left)
        if fused:
            assert dev.n_fused_steps >= ora.stats_n_substep.max()
    elif fused:
        assert dev.n_fused_steps == steps * substeps


@pytest.mark.parametrize("fused", (True, False))
@pytest.mark.parametrize("kernel", ("ConstantK", "Linear"))
def test_other_kernels_against_oracle(kernel, fused):
    """collision_kernels/constantK.py:11-12 and linear.py:12-15, fused and method-level, against
    the oracle's restatement of the same operation sequences (`fill`; `sum_pair`, `*=`, `+=`)"""
    from oracle import sdm_step  # pylint: disable=import-outside-toplevel

    pop = ragged_population(seed=7, n_cell=200, mean=60, n_attr=2)
    kw = {"ConstantK": {"a": 3e-4}, "Linear": {"a": 1e-4, "b": GOLOVIN_B}}[kernel]
    common = dict(dt=1.0, dv=1e-3, seed=3, **pop)
    ora = OracleCollision(kernel=getattr(sdm_step, kernel)(**kw), **common)
    dev = device_sim(kernel=device_kernel(kernel, **kw), host_rng=True, fused=fused, **common)
    for step in range(4):
        ora()
        dev()
        compare_states(dev, ora, f"{kernel} step {step}")
    assert ora.coalescence_rate.sum() > 0
    assert dev.n_fused_steps == (4 if fused else 0)


def test_fused_philox_equals_methods_philox():
    """in-kernel Philox uniforms == the Random kernel's stream consumed by the method path"""
    pop = ragged_population(seed=8, n_cell=500, mean=100, n_attr=1)
    sims = [device_sim(kernel=device_kernel(b=GOLOVIN_B), dt=1.0, dv=2e-7, seed=11, fused=f, **pop)
            for f in (True, False)]
    for step in range(5):
        for s in sims:
            s()
        compare_states(sims[0], _Host(sims[1]), f"philox step {step}")
    assert sims[0].host("coalescence_rate").sum() > 0


@pytest.mark.parametrize("optimized_random", (False, True))
def test_adaptive_speculation_equals_methods(optimized_random):
    """the adaptive sub-step loop with sub-steps launched ahead of the host's read-back (device
    working length, gate, roll-back of launches that did nothing) against the method-level
    kernels driven sub-step by sub-step, same device Philox stream; small multiplicities make
    sub-steps leave zero-multiplicity SDs in the middle of the loop"""
    pop = ragged_population(seed=21, n_cell=300, mean=60, n_attr=2, max_mult=2)
    common = dict(dt=1.0, dv=1e-8, seed=13, adaptive=True, dt_coal_range=(0.02, 1.0),
                  optimized_random=optimized_random)
    sims = [device_sim(kernel=device_kernel(b=GOLOVIN_B), fused=f, **pop, **common)
            for f in (True, False)]
    assert sims[0].substep_speculation > 0
    lengths = set()
    for step in range(5):
        for s in sims:
            s()
        compare_states(sims[0], _Host(sims[1]), f"speculative adaptive step {step}")
        np.testing.assert_array_equal(sims[0].host("dt_left"), sims[1].host("dt_left"))
        np.testing.assert_array_equal(sims[0].host("cell_idx"), sims[1].host("cell_idx"))
        lengths.add(sims[0].length)
    assert len(lengths) > 2, "super-droplets were expected to be removed in several steps"
    assert sims[0].host("stats_n_substep").max() > 5
    assert sims[0].n_fused_steps >= sims[0].host("stats_n_substep").max()
    assert sims[0].rnd.call == sims[1].rnd.call


def test_speculative_stepping_equals_step_by_step():
    """run() queues fused step k+1 before reading `healthy` of step k (gate word of the fused
    kernel); small multiplicities make steps leave zero-multiplicity SDs, so the gate closes,
    the queued launch must do nothing and be redone after the removal: the final state has to
    equal the one of step-by-step execution (which test_fused_philox_equals_methods_philox
    pins to the method-level path)"""
    pop = ragged_population(seed=9, n_cell=400, mean=90, n_attr=2, max_mult=2)
    sims = [device_sim(kernel=device_kernel(b=GOLOVIN_B), dt=1.0, dv=1e-7, seed=12, fused=True,
                       **pop) for _ in range(2)]
    lengths = set()
    for _ in range(8):
        sims[0]()
        lengths.add(sims[0].length)
    sims[1].run(8)
    assert len(lengths) > 2, "super-droplets were expected to be removed in several steps"
    assert sims[1].n_fused_steps == 8 and sims[1].n_launches > sims[0].n_launches
    compare_states(sims[1], _Host(sims[0]), "speculative vs step by step")


def test_host_step_stripes_equal_separate_runs():
    """HostCollisionStep (host arrays in/out, cells cut into stripes, three streams) gives what
    the stripes give when run one after another as plain DeviceCollisions"""
    from pysdm_b200.host_step import HostCollisionStep  # pylint: disable=import-outside-toplevel
    from pysdm_b200.sharding import shard_population  # pylint: disable=import-outside-toplevel

    pop = ragged_population(seed=10, n_cell=333, mean=70, n_attr=2)
    order = np.argsort(pop["cell_id"], kind="stable")
    mult, cid = pop["multiplicity"][order], pop["cell_id"][order]
    attrs = np.stack([pop["water_mass"][order]] + [e[order] for e in pop["extra_attributes"]])
    kw = dict(dt=1.0, dv=2e-7)
    n_stripes, steps = 3, 3
    host = HostCollisionStep(multiplicity=mult, attributes=attrs, cell_id=cid, n_cell=333,
                             n_stripes=n_stripes, seed=21, kernel=device_kernel(b=GOLOVIN_B), **kw)
    refs = []
    for s in range(n_stripes):
        n_s, a_s, c_s, n_cell_s, mine = shard_population(mult, attrs, cid, 333, s, n_stripes)
        refs.append((device_sim(multiplicity=n_s, water_mass=a_s[0], extra_attributes=[a_s[1]],
                                cell_id=c_s, n_cell=n_cell_s, seed=21 + s,
                                kernel=device_kernel(b=GOLOVIN_B), **kw), mine))
    for step in range(steps):
        host.step()
        for sim, mine in refs:
            if step:
                sim.reset_permutation()
            sim()
            np.testing.assert_array_equal(host.out_multiplicity.numpy()[mine],
                                          sim.host("multiplicity"), f"step {step}")
            assert_same_f64(host.out_attributes.numpy()[:, mine], sim.host("attributes"),
                            f"step {step} attributes")
        # the host model carries the state on: outputs become the next step's inputs
        host.multiplicity.copy_(host.out_multiplicity)
        host.attributes.copy_(host.out_attributes)
    rate = host.out_coalescence_rate.numpy()
    np.testing.assert_array_equal(rate, np.concatenate([s.host("coalescence_rate")
                                                        for s, _ in refs]))
    assert rate.sum() > 0 and host.h2d_bytes == mult.size * 8 * 4


def test_host_step_keeps_permutation_when_cells_are_unchanged():
    """HostCollisionStep.step(cell_id_changed=False): cell ids are not re-sent and the
    permutation with its sort by cell is carried from step to step, as a DeviceCollision that
    is simply called again carries it"""
    from pysdm_b200.host_step import HostCollisionStep  # pylint: disable=import-outside-toplevel
    from pysdm_b200.sharding import shard_population  # pylint: disable=import-outside-toplevel

    pop = ragged_population(seed=12, n_cell=200, mean=60, n_attr=1)
    order = np.argsort(pop["cell_id"], kind="stable")
    mult, cid = pop["multiplicity"][order], pop["cell_id"][order]
    attrs = pop["water_mass"][order].reshape(1, -1)
    kw = dict(dt=1.0, dv=2e-7)
    n_stripes = 2
    host = HostCollisionStep(multiplicity=mult, attributes=attrs, cell_id=cid, n_cell=200,
                             n_stripes=n_stripes, seed=5, kernel=device_kernel(b=GOLOVIN_B), **kw)
    refs = []
    for s in range(n_stripes):
        n_s, a_s, c_s, n_cell_s, mine = shard_population(mult, attrs, cid, 200, s, n_stripes)
        refs.append((device_sim(multiplicity=n_s, water_mass=a_s[0], cell_id=c_s, n_cell=n_cell_s,
                                seed=5 + s, kernel=device_kernel(b=GOLOVIN_B), **kw), mine))
    removed = 0
    for step in range(5):
        host.step(cell_id_changed=step == 0)
        assert host.h2d_bytes == mult.size * 8 * (3 if step == 0 else 2)
        for sim, mine in refs:
            sim()
            np.testing.assert_array_equal(host.out_multiplicity.numpy()[mine],
                                          sim.host("multiplicity"), f"step {step}")
            assert_same_f64(host.out_attributes.numpy()[:, mine], sim.host("attributes"),
                            f"step {step} attributes")
        removed = sum(sim.n_sd - sim.length for sim, _ in refs)
        host.multiplicity.copy_(host.out_multiplicity)
        host.attributes.copy_(host.out_attributes)
    assert removed > 0, "super-droplets were expected to be removed along the way"


@pytest.mark.parametrize("adaptive", (False, True))
def test_full_size_fused_equals_oracle(adaptive):
    """BASELINE.json's headline size -- 2^24 super-droplets in 2^17 cells of 128, the bench.py
    workload (configs[4]) -- fused step against the CPU oracle, same uniforms (the oracle's
    PCG64 stream replayed on the device), two time steps, everything bit for bit; also with the
    adaptive time step (one fused launch per sub-step)"""
    n_sd, per_cell = 1 << 24, 128
    rng = np.random.default_rng(4321)
    pop = {"multiplicity": rng.integers(1, 5, n_sd).astype(np.int64),
           "water_mass": rng.exponential(1000 * 4 / 3 * np.pi * 30.531e-6 ** 3, n_sd),
           "cell_id": np.arange(n_sd, dtype=np.int64) // per_cell, "n_cell": n_sd // per_cell}
    common = dict(dt=1.0, dv=2e-7 if not adaptive else 4e-8, seed=7)
    if adaptive:
        common.update(adaptive=True, dt_coal_range=(0.25, 1.0))
    ora = OracleCollision(kernel=OGolovin(GOLOVIN_B), **pop, **common)
    dev = device_sim(kernel=device_kernel(b=GOLOVIN_B), host_rng=True, fused=True, **pop, **common)
    for step in range(2):
        ora()
        dev()
        compare_states(dev, ora, f"2^24 step {step}")
    assert ora.coalescence_rate.sum() > 0
    assert dev.n_fused_steps >= (int(ora.stats_n_substep.max()) if adaptive else 2)
    if adaptive:
        assert ora.stats_n_substep.max() > 2, "the time step was never cut"


def test_full_size_properties_and_fused_equals_methods():
    """BASELINE.json's headline size (2^24 super-droplets, 2^17 cells of 128, the bench.py
    workload) with the device Philox stream (which the oracle cannot replay; the test above
    pins the same kernel on the oracle's stream): the fused step must equal the method-level
    kernels bit for bit (same Philox stream), and the size-independent invariants of a
    coalescence step must hold: idx stays a permutation, every cell keeps its droplets, the
    number of pairs is sum(n_c // 2), the drop in total multiplicity equals the coalescence
    counters exactly, and total water mass sum(n * m) is conserved to 1e-12"""
    import torch  # pylint: disable=import-outside-toplevel

    n_sd, per_cell = 1 << 24, 128
    rng = np.random.default_rng(1234)
    pop = {"multiplicity": rng.integers(1, 5, n_sd).astype(np.int64),
           "water_mass": rng.exponential(1000 * 4 / 3 * np.pi * 30.531e-6 ** 3, n_sd),
           "cell_id": np.arange(n_sd, dtype=np.int64) // per_cell, "n_cell": n_sd // per_cell}
    sims = [device_sim(kernel=device_kernel(b=GOLOVIN_B), dt=1.0, dv=2e-7, seed=5, fused=f, **pop)
            for f in (True, False)]
    n0 = int(pop["multiplicity"].sum())
    mass0 = float((pop["multiplicity"] * pop["water_mass"]).sum())
    for step in range(2):
        for s in sims:
            s()
        a, b = sims
        assert a.length == b.length
        for name in ("idx", "flag", "multiplicity", "attributes", "gamma", "coalescence_rate",
                     "collision_rate", "collision_rate_deficit", "cell_start"):
            x, y = getattr(a, name), getattr(b, name)
            if name in ("idx", "flag"):
                x, y = x[:a.length], y[:b.length]
            if x.dtype == torch.float64:  # bit patterns (sign of zero, NaN fill beyond the length)
                x, y = x.view(torch.int64), y.view(torch.int64)
            assert torch.equal(x, y), f"step {step}: {name}"
        idx = a.idx[:a.length]
        assert int(torch.unique(idx).numel()) == a.length, "idx is not a permutation"
        if a.sorted:  # (the removal of zero-multiplicity SDs leaves idx unsorted until the next step)
            cells = a.cell_id[idx]
            assert bool((cells[1:] >= cells[:-1]).all()), "idx no longer sorted by cell"
        sizes = a.cell_start[1:] - a.cell_start[:-1]  # of the sort this step ran on
        assert int(sizes.sum()) == a.sorted_length
        assert int(a.flag[:a.sorted_length].sum()) == int((sizes // 2).sum())
        n_now = int(a.multiplicity[idx].sum())
        assert n0 - n_now == int(a.coalescence_rate.sum()) > 0
        mass = float((a.multiplicity[idx].double() * a.attributes[0][idx]).sum())
        assert abs(mass - mass0) <= 1e-12 * mass0
    assert sims[0].n_fused_steps == 2 and sims[1].n_fused_steps == 0


@pytest.mark.parametrize("mean, n_cell", ((3, 2000), (128, 300), (3000, 4), (700, 1)))
def test_batched_shuffle_local_equals_oracle(mean, n_cell):
    """pysdm_b200_shuffle_local_batched (the shuffle phases of the fused kernel alone, what the
    backend uses when the sort told it that every cell is small) against the oracle's serial
    shuffle_local, ragged cells incl. empty ones, unaligned u01 view as `optimized_random` makes"""
    import torch  # pylint: disable=import-outside-toplevel

    from oracle import sdm_oracle as orc  # pylint: disable=import-outside-toplevel
    from pysdm_b200 import methods as m  # pylint: disable=import-outside-toplevel

    rng = np.random.default_rng(mean)
    sizes = rng.poisson(mean, n_cell)
    sizes[rng.integers(0, n_cell, max(1, n_cell // 10))] = 0
    cell_start = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int64)
    n = int(cell_start[-1])
    n_sd = n + 5  # positions beyond the working length stay untouched
    idx = rng.permutation(n_sd).astype(np.int64)
    for shift in (0, 1):
        u_all = rng.random(n_sd + shift)
        expect = idx.copy()
        orc.shuffle_local(expect, np.ascontiguousarray(u_all[shift:]), cell_start)
        d_idx = torch.from_numpy(idx.copy()).cuda()
        d_u = torch.from_numpy(u_all).cuda()[shift:]
        m.shuffle_local(d_idx, d_u, torch.from_numpy(cell_start).cuda(), n, int(sizes.max()))
        np.testing.assert_array_equal(d_idx.cpu().numpy(), expect)


class _Host:  # pylint: disable=too-few-public-methods
    """numpy face of a DeviceCollision, shaped like OracleCollision for compare_states"""

    def __init__(self, sim):
        self.sim = sim
        self.length = sim.length

    def __getattr__(self, name):
        return self.sim.host(name)


@pytest.mark.parametrize("n_cell", (96, 1))
@pytest.mark.parametrize("mode", ("fixed", "adaptive_optrand"))
@pytest.mark.parametrize("fused", (True, False))
def test_geometric_standalone_against_oracle(fused, mode, n_cell):
    """BASELINE configs[3] in kind without PySDM in the process: `DeviceCollision` with the
    Geometric kernel (radius and Gunn-Kinzer fall velocity derived from the water mass; the
    fused step refreshes them only for the super-droplets a coalescence changed, the
    method-level path recomputes them in full before every step), A = 3 attributes, cells of
    ~128, adaptive time step with `optimized_random` -- against the oracle, which recomputes
    the derived attributes in full like the orchestration.  Radii go through CUDA's pow here
    and glibc's there: integers, permutation and gamma must still agree (flips are counted),
    attributes to 1e-12."""
    from oracle.sdm_step import Geometric as OGeometric  # pylint: disable=import-outside-toplevel

    table = np.load(os.path.join(os.path.dirname(__file__), "..", "pysdm_b200", "data",
                                 "gunn_kinzer_1949.npz"))
    rng = np.random.default_rng(31)
    sizes = rng.poisson(12288 // n_cell, n_cell)  # one cell: beyond the shared-memory batches
    cell_id = np.repeat(np.arange(n_cell), sizes)
    n_sd = len(cell_id)
    radius = 10e-6 * np.exp(0.47 * rng.standard_normal(n_sd))
    radius[rng.random(n_sd) < 0.02] *= 4.0
    volume = 4 / 3 * np.pi * radius ** 3
    dv = (1500 / 512) ** 2
    pop = {"multiplicity": rng.integers(1, 2 * int(100e6 * dv / 128), n_sd).astype(np.int64),
           "water_mass": 1000.0 * volume, "cell_id": cell_id, "n_cell": n_cell,
           "extra_attributes": [volume * 1e-3, volume * 0.6e-3]}
    # (one cell of 12288: every droplet meets ~96 times more partners; same droplets, 96 x the volume)
    common = dict(dt=5.0, dv=dv * (1 if n_cell > 1 else 96), seed=77)
    if mode == "adaptive_optrand":
        common.update(adaptive=True, optimized_random=True, dt_coal_range=(0.1, 5.0))
    ora = OracleCollision(kernel=OGeometric(1.0, (table["a"], table["b"], float(table["factor"]),
                                                  float(table["maximum_radius"]))),
                          **pop, **common)
    dev = device_sim(kernel=device_kernel("Geometric", collection_efficiency=1.0), host_rng=True,
                     fused=fused, **pop, **common)
    for step in range(4):
        ora()
        dev()
        what = f"geometric {mode} step {step}"
        n = ora.length
        assert dev.length == n, what
        np.testing.assert_array_equal(dev.host("idx")[:n], ora.idx[:n], what)
        np.testing.assert_array_equal(dev.host("flag")[:n], ora.flag[:n], what)
        flips = int((dev.host("gamma") != ora.gamma).sum())
        assert flips == 0, f"{what}: {flips} gamma values differ (integer-boundary flips)"
        np.testing.assert_array_equal(dev.host("multiplicity"), ora.multiplicity, what)
        np.testing.assert_allclose(dev.host("attributes"), ora.attributes, rtol=1e-12, atol=0)
        for k in ("collision_rate", "collision_rate_deficit", "coalescence_rate",
                  "stats_n_substep"):
            np.testing.assert_array_equal(dev.host(k), getattr(ora, k), f"{what}: {k}")
        # the derived attributes the fused step maintains == a full recomputation
        expect = np.empty(n_sd)
        ora.be.product_scalar(expect, ora.volume(), 1 / ora.kernel.PI_4_3)
        ora.be.power(expect, 1 / 3)
        if not fused:  # the method-level path recomputes them before a step, like the orchestration
            dev.refresh_dependents()
        np.testing.assert_allclose(dev.host("radius")[ora.idx[:n]], expect[ora.idx[:n]],
                                   rtol=1e-14, atol=0, err_msg=what)
        kept = {k: dev.host(k).copy() for k in ("radius", "fall_velocity", "_volume")}
        dev.refresh_dependents()
        for k, v in kept.items():
            np.testing.assert_array_equal(dev.host(k).view(np.int64), v.view(np.int64),
                                          f"{what}: {k} kept current by the step != recomputed")
    assert ora.coalescence_rate.sum() > 0
    if mode != "fixed":
        assert ora.stats_n_substep.max() > 4, "the time step was never cut"
    if fused:
        assert dev.n_fused_steps >= 4
