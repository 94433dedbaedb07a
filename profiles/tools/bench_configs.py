"""Supplementary timings (NOT the bench.py contract): BASELINE.json configs[0..2] -- the
single-cell boxes -- at their stated sizes, run through the UNMODIFIED PySDM orchestration
(baseline/_ref) once with `pysdm_b200.B200` and once with the reference's Numba CPU backend.
These configurations use the method-level kernels (one giant cell does not fit the fused
step): what is measured is the drop-in backend as a PySDM user sees it, Python orchestration
included.

    python profiles/tools/bench_configs.py [--steps 10] [--cpu-steps 2] [--configs 0,1,2]

One JSON line per (config, backend)."""

import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle", "refshims"))
_REF = os.path.join(ROOT, "baseline", "_ref")  # the unmodified reference (travels to the GPU box)
if not os.path.isdir(os.path.join(_REF, "PySDM")):
    raise ImportError("no PySDM install under baseline/_ref")
sys.path.insert(1, _REF)
os.environ.setdefault("CI", "1")
np.NaN = np.nan


def dynamics(config, fused=False):
    from PySDM.dynamics import Coalescence, Collision
    if fused:
        from pysdm_b200 import FusedCoalescence as Coalescence, FusedCollision as Collision
    from PySDM.dynamics.collisions.breakup_efficiencies import ConstEb
    from PySDM.dynamics.collisions.breakup_fragmentations import Straub2010Nf
    from PySDM.dynamics.collisions.coalescence_efficiencies import Straub2010Ec
    from PySDM.dynamics.collisions.collision_kernels import Geometric, Golovin, Hydrodynamic
    from PySDM.physics import si

    if config == 0:  # Shima et al. 2009 box
        return Coalescence(collision_kernel=Golovin(b=1.5e3), adaptive=False)
    if config == 4:  # bench.py's workload (configs[4]) through the PySDM orchestration
        return Coalescence(collision_kernel=Golovin(b=1.5e3), adaptive=False, croupier="local")
    if config == 5:  # the same multi-cell population with the gravitational (Geometric) kernel
        return Coalescence(collision_kernel=Geometric(), adaptive=False, croupier="local")
    if config == 6:  # ... and with Berry's hydrodynamic collection efficiency
        return Coalescence(collision_kernel=Hydrodynamic(), adaptive=False, croupier="local")
    if config == 1:  # Berry 1967 box
        return Coalescence(collision_kernel=Hydrodynamic(), adaptive=True,
                           dt_coal_range=(0.1, 100.0))
    if config == 3:  # the collision step of the Sc-rain case (Morrison_and_Grabowski_2007/common.py:30-35)
        return Coalescence(collision_kernel=Geometric(collection_efficiency=1), adaptive=True,
                           optimized_random=True)
    vmin = (0.01 * si.mm) ** 3 * np.pi / 6  # de Jong et al. breakup box
    return Collision(collision_kernel=Geometric(), coalescence_efficiency=Straub2010Ec(),
                     breakup_efficiency=ConstEb(1.0),
                     fragmentation_function=Straub2010Nf(vmin=vmin, nfmax=10000), adaptive=True,
                     warn_overflows=False)


SETUP = {
    0: dict(name="Shima 2009 box, Golovin, 2^17 SDs", log2=17, dt=1.0, dv=1e6, r0=30.531e-6,
            norm=2 ** 23 * 1e6, fkw={}),
    1: dict(name="Berry 1967 box, Hydrodynamic + adaptive, 2^20 SDs", log2=20, dt=10.0, dv=10.0,
            r0=10e-6, norm=239e6 * 10, fkw={}),
    2: dict(name="de Jong 2022 breakup box, Geometric + Straub, 2^22 SDs", log2=22, dt=1.0,
            dv=1e6, r0=0.2e-3, norm=1e3 * 1e6,
            fkw={"fragmentation_function": "Straub2010Nf"}),
    # per-GPU share of configs[3]: 32768 cells of 128 SDs, dv of one cell of the 1.5 km / 512 grid
    3: dict(name="Sc-rain collision step, Geometric + adaptive + optimized_random, A=3, "
                 "2^22 SDs in 32768 cells", log2=22,
            dt=5.0, dv=(1500 / 512) ** 2, r0=10e-6, norm=None, fkw={}, sd_per_cell=128, n_attr=3),
    4: dict(name="Golovin multi-cell coalescence, 2^24 SDs in 2^17 cells (bench.py workload)",
            log2=24, dt=1.0, dv=1.0, r0=30.531e-6, norm=None, fkw={}, sd_per_cell=128,
            multiplicity=2 ** 20),
    6: dict(name="Hydrodynamic multi-cell coalescence, non-adaptive, 2^24 SDs in 2^17 cells",
            log2=24, dt=1.0, dv=1.0, r0=30.531e-6, norm=None, fkw={}, sd_per_cell=128,
            multiplicity=2 ** 20),
    5: dict(name="Geometric multi-cell coalescence, non-adaptive, 2^24 SDs in 2^17 cells",
            log2=24, dt=1.0, dv=1.0, r0=30.531e-6, norm=None, fkw={}, sd_per_cell=128,
            multiplicity=2 ** 20),
}


BYTES_PER_SD = {0: 73, 1: 73, 2: 81, 3: 105, 4: 73, 5: 73, 6: 73}  # SURVEY.md 8d (A, breakup)


class Spy:
    """counts the (sub-)steps a run executes and the super-droplets each one works on (the
    working length), and times the fused launches with CUDA events: wraps the two C-ABI entry
    points every (sub-)step goes through -- pysdm_b200.methods.fused_collision_step (fused path)
    or compute_gamma (method-level path)"""

    def __init__(self):
        self.sd_substeps = self.substeps = 0
        self.events = []

    def __enter__(self):
        import torch
        from pysdm_b200 import methods as m

        self.m, self.real_fused, self.real_gamma = m, m.fused_collision_step, m.compute_gamma

        def fused(**kw):
            self.substeps += 1
            self.sd_substeps += int(kw["length"])
            beg, end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            beg.record()
            err = self.real_fused(**kw)
            end.record()
            self.events.append((beg, end))
            return err

        def gamma(prob, rand, idx, length, *args):
            self.substeps += 1
            self.sd_substeps += int(length)
            return self.real_gamma(prob, rand, idx, length, *args)

        m.fused_collision_step, m.compute_gamma = fused, gamma
        return self

    def __exit__(self, *exc):
        self.m.fused_collision_step, self.m.compute_gamma = self.real_fused, self.real_gamma

    def fused_ms(self):
        return sum(a.elapsed_time(b) for a, b in self.events)


def run(config, backend_name, steps, log2=None, peak_gbs=None):
    from PySDM import Builder, Formulae
    from PySDM.environments import Box

    c = SETUP[config]
    n_sd = 1 << (log2 or c["log2"])
    rng = np.random.default_rng(44)
    x0 = 4 / 3 * np.pi * c["r0"] ** 3
    volume = rng.exponential(x0, n_sd)
    env = Box(dv=c["dv"], dt=c["dt"])
    if "sd_per_cell" in c:  # multi-cell: mesh set as tests/unit_tests/dynamics/collisions/test_sdm_multi_cell.py:25-27
        from PySDM.impl.mesh import Mesh

        n_cell = n_sd // c["sd_per_cell"]
        env.mesh = Mesh((n_cell,), size=(n_cell,), dv=c["dv"])
        # ~100 droplets / cm^3 in every cell, lognormal-ish radii around r0
        if "multiplicity" not in c:
            volume = 4 / 3 * np.pi * (c["r0"] * np.exp(0.47 * rng.standard_normal(n_sd))) ** 3
        attributes = {"volume": volume, "cell id": np.arange(n_sd, dtype=np.int64) // c["sd_per_cell"],
                      "multiplicity": np.full(
                          n_sd, c.get("multiplicity", int(100e6 * c["dv"] / c["sd_per_cell"])),
                          dtype=np.int64)}
        if c.get("n_attr", 1) == 3:  # environments/kinematic_2d.py:63-64,84
            attributes["dry volume"] = volume * 1e-3
            attributes["kappa times dry volume"] = volume * 0.6e-3
    else:
        attributes = {"volume": volume, "multiplicity": np.full(
            n_sd, max(1, int(c["norm"] / n_sd)), dtype=np.int64)}
    formulae = Formulae(seed=44, **c["fkw"])
    gpu = backend_name.startswith("B200")
    if gpu:  # "B200F": FusedCoalescence instead of Coalescence
        from pysdm_b200 import B200

        backend = B200(formulae)
    else:
        from PySDM.backends import CPU

        backend = CPU(formulae)
    builder = Builder(n_sd=n_sd, backend=backend, environment=env)
    builder.add_dynamic(dynamics(config, fused=backend_name == "B200F"))
    for name in ("dry volume", "kappa times dry volume"):
        if name in attributes:
            builder.request_attribute(name)
    particulator = builder.build(attributes=attributes, products=())
    particulator.run(2)  # warm-up (JIT / workspaces)
    dyn = particulator.dynamics["Collision"]
    out = {"config": c["name"], "backend": backend_name, "n_sd": n_sd, "steps": steps}
    if gpu:
        import torch

        beg, end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        with Spy() as spy:
            t0 = time.perf_counter()
            beg.record()
            particulator.run(steps)
            end.record()
            torch.cuda.synchronize()
            dt = time.perf_counter() - t0
        ms = beg.elapsed_time(end) / steps  # CUDA events around the steps (host overhead included)
        algorithmic = BYTES_PER_SD[config] * spy.sd_substeps / steps
        out.update({
            "ms_per_step": ms, "wall_ms_per_step": dt / steps * 1e3,
            "sd_steps_per_s": n_sd / (ms * 1e-3),
            "substeps_per_step": spy.substeps / steps,
            "fused_steps": int(getattr(dyn, "n_fused_steps", 0)),
            "roofline": {
                "bound": "hbm", "unit": "GB/s", "algorithmic_bytes_per_step": algorithmic,
                "bytes_per_sd_substep": BYTES_PER_SD[config],
                "achieved": algorithmic / (ms * 1e-3) / 1e9, "peak": peak_gbs,
                "frac": None if not peak_gbs else algorithmic / (ms * 1e-3) / 1e9 / peak_gbs,
                "fused_kernel_ms_per_step": spy.fused_ms() / steps if spy.events else None,
                "frac_fused_kernels_only": None if not (peak_gbs and spy.events) else
                algorithmic / (spy.fused_ms() / steps * 1e-3) / 1e9 / peak_gbs}})
    else:
        t0 = time.perf_counter()
        particulator.run(steps)
        dt = time.perf_counter() - t0
        out.update({"ms_per_step": dt / steps * 1e3, "sd_steps_per_s": n_sd * steps / dt,
                    "threads": os.environ.get("NUMBA_NUM_THREADS", "default")})
    if gpu and backend_name == "B200F" and KERNEL_ONLY:
        kernel_only(particulator, c, n_sd)
    out["super_droplets_left"] = int(particulator.attributes.super_droplet_count)
    return out


def all_configs(steps, peak_gbs, numba=True, cpu_log2=20):
    """BASELINE configs[0..3] through `Builder(backend=B200())` (the plug-in call) at their
    stated sizes, with the fused dynamic where it applies; the reference's Numba backend next
    to each on a bounded sample (2^cpu_log2 SDs at most; SD-steps/s is size-normalised)"""
    rows = []
    for config in (0, 1, 2, 3):
        row = run(config, "B200F", steps, peak_gbs=peak_gbs)
        row["baseline_config"] = f"configs[{config}]"
        if numba:
            try:
                cpu = run(config, "CPU", 1, log2=min(cpu_log2, SETUP[config]["log2"]))
                row["numba"] = {k: cpu[k] for k in ("n_sd", "ms_per_step", "sd_steps_per_s",
                                                    "threads")}
                row["speedup_vs_numba"] = row["sd_steps_per_s"] / cpu["sd_steps_per_s"]
            except Exception as exc:  # pylint: disable=broad-except
                row["numba"] = {"unavailable": f"{type(exc).__name__}: {exc}"[:200]}
        rows.append(row)
    return rows


def pysdm_path(steps, peak_gbs):
    """BASELINE configs[4] (the bench.py workload) through the unmodified PySDM orchestration on
    the B200 backend: with the stock `Coalescence` (method-level kernels) and with the fused
    dynamic"""
    return {"stock_coalescence": run(4, "B200", steps, peak_gbs=peak_gbs),
            "fused_coalescence": run(4, "B200F", steps, peak_gbs=peak_gbs)}


KERNEL_ONLY = False


def kernel_only(particulator, c, n_sd, launches=50):
    """the fused launch alone (no orchestration, no dependent-attribute refresh between
    launches: same memory traffic, radius / velocity simply not updated), CUDA events on the
    launch stream; roofline by the algorithmic bytes of DESIGN.md 3.2 (+16 B/SD for the two
    gathered attributes of the Geometric kernel)"""
    import torch

    from pysdm_b200 import methods as m

    dyn = particulator.dynamics["Collision"]
    calls = []
    real = m.fused_collision_step

    def spy(**kw):
        calls.append(kw)
        return real(**kw)

    m.fused_collision_step = spy
    try:
        particulator.run(1)
    finally:
        m.fused_collision_step = real
    kw = dict(calls[-1])
    a, b = kw["idx_in"], kw["idx_out"]
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    for i in range(launches + 5):
        if i == 5:
            ev[0].record()
        kw["idx_in"], kw["idx_out"] = (a, b) if i % 2 == 0 else (b, a)
        kw["call_u01"], kw["call_rand"] = 1000 + 2 * i, 1001 + 2 * i
        kw["refresh_nf"] = False
        assert real(**kw) == 0
    ev[1].record()
    torch.cuda.synchronize()
    ms = ev[0].elapsed_time(ev[1]) / launches
    geometric = kw["kernel"] in ("Geometric", "Hydrodynamic")
    bytes_per_sd = 57 + 16 * kw["attributes"].shape[0] + (16 if geometric else 0)
    peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
    gbs = bytes_per_sd * n_sd / ms / 1e6
    print(json.dumps({"config": c["name"], "what": "fused launch only (CUDA events)",
                      "kernel": kw["kernel"], "launches": launches, "kernel_ms": ms,
                      "algorithmic_bytes_per_sd": bytes_per_sd, "achieved_gbs": gbs,
                      "peak_gbs_measured": peak, "frac": gbs / peak}), flush=True)
    assert dyn.n_fused_steps > 0


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--cpu-steps", type=int, default=2)
    ap.add_argument("--configs", default="0,1,2")
    ap.add_argument("--backends", default="B200,CPU")
    ap.add_argument("--log2", type=int, default=None)
    ap.add_argument("--kernel-only", action="store_true",
                    help="B200F: also time the fused launch alone with CUDA events")
    a = ap.parse_args()
    KERNEL_ONLY = a.kernel_only
    _peak = None
    if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")):
        _peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
    for cfg in (int(x) for x in a.configs.split(",")):
        for be in a.backends.split(","):
            print(json.dumps(run(cfg, be, a.steps if be.startswith("B200") else a.cpu_steps,
                                 a.log2, _peak)), flush=True)
