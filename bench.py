"""bench.py -- SD-steps/s of the SDM collision step (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]
    python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N ...

Workload (BASELINE.json configs[4], the one the metric is quoted
on): multi-cell Golovin coalescence, 2^24 super-droplets per GPU, 128 per cell (2^17 cells
per GPU), non-adaptive, "local" croupier, A = 1 extensive attribute, dt = 1 s, dv = 1 m^3.
A "step" is one model time step of the Collision dynamic on the whole population of every
rank.  Cells are sharded over ranks (weak scaling); the collision step needs no
communication -- the only collective is the timing barrier/max.

One JSON line on stdout (rank 0):
  value / ms_per_step   device-resident state, K fused steps between barriers (CUDA events)
  roofline              the fused launch: algorithmic bytes / mean launch time (CUDA events per
                        launch) against MEASURED_PEAKS.json; `traffic` = DRAM bytes of the
                        committed ncu capture, `achieved_dram` = traffic / launch time
  e2e                   the same step through the plug-in (Builder(backend=B200()) +
                        Particulator.run(1)), HOST state in / out every step (pinned memory); the
                        cells in 4 stripes, one Particulator and CUDA stream each, so that one
                        stripe's upload overlaps another's read-back (measured: 4 stripes 6.9 ms,
                        8: 8.2 ms, 16: 9.2 ms per step -- the host side of a stripe costs more
                        than the shorter pipeline tail saves)
  e2e_single_particulator   the same with one Particulator (no overlap of the two directions)
  e2e_striped           the same through HostCollisionStep (cells in stripes over 3 streams)
  cpu_baseline          the oracle port (and the reference's own Numba backend) on the host cores
  per_rank              elapsed time of every rank (N > 1)
  configs               BASELINE configs[0..3] through Builder(backend=B200()) (N = 1 only)
  pysdm_path            configs[4] through PySDM: stock Coalescence and the fused path (N = 1 only)
  transport             displacement -> NCCL migration -> sort -> adaptive collision loop with its
                        invariants (configs[3] in kind, cells sharded over the N ranks)

`--impl reference` times the CPU restatement of the reference's Numba path (oracle/, kind
"port") -- and next to it the reference's own Numba backend when it is importable -- on the
host cores, at the same configuration.
"""

import argparse
import ctypes
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

X0 = 4.0 / 3.0 * np.pi * 30.531e-6 ** 3  # mean droplet volume (Shima et al. 2009), m^3
RHO_W = 1000.0
GOLOVIN_B = 1.5e3
BYTES_PER_SD = 73  # algorithmic HBM bytes per super-droplet per step, A = 1


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="b200", choices=("b200", "reference"))
    ap.add_argument("--log2-n-sd", type=int, default=24, help="super-droplets per GPU (log2)")
    ap.add_argument("--sd-per-cell", type=int, default=128)
    ap.add_argument("--cpu-log2-n-sd", type=int, default=22,
                    help="bounded CPU sample of the GPU arm's cpu_baseline leg (log2 SDs)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-numba", action="store_true",
                    help="skip timing the reference's own Numba backend next to the port")
    ap.add_argument("--no-configs", action="store_true",
                    help="skip BASELINE configs[0..3] and the PySDM path of configs[4]")
    ap.add_argument("--no-transport", action="store_true")
    ap.add_argument("--e2e-steps", type=int, default=10)
    ap.add_argument("--e2e-stripes", type=int, default=4)
    ap.add_argument("--e2e-plugin-stripes", type=int, default=4,
                    help="Particulators (stripes of cells, one CUDA stream each) of the e2e leg")
    ap.add_argument("--config-steps", type=int, default=10)
    ap.add_argument("--transport-steps", type=int, default=20)
    return ap.parse_args()


def population(n_sd, sd_per_cell, rank):
    """seeded synthetic droplet population, cell-contiguous placement"""
    rng = np.random.default_rng(1234 + rank)
    volume = rng.exponential(X0, n_sd)
    return {
        "multiplicity": np.full(n_sd, 2 ** 20, dtype=np.int64),
        "water_mass": volume * RHO_W,
        "cell_id": (np.arange(n_sd, dtype=np.int64) // sd_per_cell),
        "n_cell": (n_sd + sd_per_cell - 1) // sd_per_cell,
    }


def workload_name(args):
    return (f"Golovin multi-cell coalescence, 2^{args.log2_n_sd} SD/GPU, {args.sd_per_cell} SD/cell, "
            "non-adaptive, local croupier, A=1 (BASELINE configs[4])")


def pysdm_paths():
    """makes the unmodified reference (baseline/_ref, travels to the GPU box) importable;
    returns False when it is not there"""
    ref = os.path.join(ROOT, "baseline", "_ref")
    if not os.path.isdir(os.path.join(ref, "PySDM")):
        return False
    os.environ.setdefault("CI", "1")  # PySDM's own switch: do not import ThrustRTC
    for path in (ref, os.path.join(ROOT, "oracle", "refshims")):  # pint / chempy stand-ins
        if path not in sys.path:
            sys.path.insert(0, path)
    np.NaN = np.nan  # the reference predates NumPy 2
    return True


# ---- CPU arm: the oracle port on the host cores -------------------------------------------
def cpu_run(args, steps, warmup, log2_n_sd):
    """times OracleCollision (C restatement of the Numba path, OpenMP where the reference
    uses prange) on 2^log2_n_sd SDs of the workload per step.  The OpenMP thread count is swept
    (1, 2, 4, ... all cores) on one step each and the best is used: on shared/virtualised
    hosts "all threads" can be far slower than a few."""
    from oracle import sdm_oracle as orc  # pylint: disable=import-outside-toplevel
    from oracle.sdm_step import Golovin, OracleCollision  # pylint: disable=import-outside-toplevel

    n_sd = 1 << log2_n_sd
    pop = population(n_sd, args.sd_per_cell, 0)
    sim = OracleCollision(kernel=Golovin(GOLOVIN_B), dt=1.0, dv=1.0, seed=44, **pop)
    ncpu = os.cpu_count() or 1
    counts = sorted({min(1 << k, ncpu) for k in range(0, 12)} | {ncpu})
    sweep = {}
    sim()
    for c in counts:
        orc.set_num_threads(c)
        t0 = time.perf_counter()
        sim()
        sweep[c] = time.perf_counter() - t0
        if sweep[c] > 4 * min(sweep.values()):
            break  # oversubscribed: more threads only get slower
    cores = min(sweep, key=sweep.get)
    orc.set_num_threads(cores)
    # keep the whole arm under about a minute: the step count the driver asked for, unless
    # that would take longer (then as many steps as fit, and the line says so)
    budget = 60.0
    fit = max(1, int(budget / max(sweep[cores], 1e-9)))
    clamped = steps + warmup > fit
    if clamped:
        warmup = min(warmup, max(1, fit // 5))
        steps = max(1, min(steps, fit - warmup))
    for _ in range(warmup):
        sim()
    t0 = time.perf_counter()
    for _ in range(steps):
        sim()
    dt = time.perf_counter() - t0
    return {
        "value": n_sd * steps / dt, "unit": "SD-steps/s", "cores": cores, "kind": "port",
        "sample": f"{steps} steps of 2^{log2_n_sd} SDs ({args.sd_per_cell}/cell) of the "
                  f"same workload; C port of the Numba path (oracle/), numpy PCG64 RNG included; "
                  f"best of OpenMP thread sweep {sorted(sweep)} = {cores} threads on "
                  f"{ncpu} host cores" + ("; step count cut to fit 60 s" if clamped else ""),
        "ms_per_step": dt / steps * 1e3, "steps": steps, "warmup": warmup,
    }


def numba_run(args, log2_n_sd, steps=2):
    """the reference's OWN Numba CPU backend (unmodified PySDM from baseline/_ref, its stock
    Builder / Particulator / Coalescence path) on the same workload, when it is importable
    on this box (numba + the install under baseline/_ref); reported next to the port, never used
    as the denominator of anything.  Returns a dict, or {"unavailable": why}."""
    if not pysdm_paths():
        return {"unavailable": "no PySDM install under baseline/_ref"}
    try:
        import numba  # pylint: disable=import-outside-toplevel
        from PySDM import Builder, Formulae  # pylint: disable=import-outside-toplevel
        from PySDM.backends import CPU  # pylint: disable=import-outside-toplevel
        from PySDM.dynamics import Coalescence  # pylint: disable=import-outside-toplevel
        from PySDM.dynamics.collisions.collision_kernels import Golovin  # pylint: disable=import-outside-toplevel
        from PySDM.environments import Box  # pylint: disable=import-outside-toplevel
        from PySDM.impl.mesh import Mesh  # pylint: disable=import-outside-toplevel

        n_sd = 1 << log2_n_sd
        pop = population(n_sd, args.sd_per_cell, 0)
        env = Box(dv=1.0, dt=1.0)
        env.mesh = Mesh((pop["n_cell"],), size=(pop["n_cell"],), dv=1.0)
        builder = Builder(n_sd=n_sd, backend=CPU(Formulae(seed=44)), environment=env)
        builder.add_dynamic(Coalescence(collision_kernel=Golovin(b=GOLOVIN_B), adaptive=False,
                                        croupier="local"))
        particulator = builder.build(products=(), attributes={
            "multiplicity": pop["multiplicity"], "volume": pop["water_mass"] / RHO_W,
            "cell id": pop["cell_id"]})
        particulator.run(1)  # JIT
        ncpu, best = numba.config.NUMBA_NUM_THREADS, None
        for threads in sorted({1, ncpu}):
            numba.set_num_threads(threads)
            particulator.run(1)
            t0 = time.perf_counter()
            particulator.run(steps)
            dt = (time.perf_counter() - t0) / steps
            if best is None or dt < best[1]:
                best = (threads, dt)
        return {"value": n_sd / best[1], "unit": "SD-steps/s", "cores": best[0],
                "kind": "reference", "ms_per_step": best[1] * 1e3,
                "sample": f"{steps} steps of 2^{log2_n_sd} SDs ({args.sd_per_cell}/cell), "
                          f"PySDM Numba backend (baseline/_ref, stock Builder/Coalescence path), "
                          f"best of {sorted({1, ncpu})} numba threads"}
    except Exception as exc:  # pylint: disable=broad-except
        return {"unavailable": f"{type(exc).__name__}: {exc}"[:200]}


def reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    # the stated configuration (2^24 SDs per step), the driver's step counts
    r = cpu_run(args, max(1, args.steps), max(1, args.warmup), args.log2_n_sd)
    line = {
        "impl": "reference", "metric": "SD-steps/s, SDM collision step", "value": r["value"],
        "unit": "SD-steps/s", "n_gpus": args.gpus, "steps": r["steps"], "warmup": r["warmup"],
        "ms_per_step": r["ms_per_step"], "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": workload_name(args), "n_sd_per_gpu": 1 << args.log2_n_sd,
                   "timed": r["sample"]},
        "cpu_baseline": {k: r[k] for k in ("value", "unit", "cores", "kind", "sample")},
        "e2e": {"value": r["value"], "unit": "SD-steps/s", "h2d_bytes_per_step": 0,
                "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    if not args.no_numba:
        line["cpu_baseline"]["numba_backend"] = numba_run(args, min(args.log2_n_sd, 22))
    print(json.dumps(line))


# ---- clocks: NVML through ctypes (no fork, no extra dependency) ------------------------------
class ClockSampler(threading.Thread):
    """samples SM clock and throttle reasons of one GPU every 0.5 ms while the timed region runs,
    straight from libnvidia-ml.so.1 (a query takes ~0.1 ms; nothing is forked)"""

    REASONS = (("sw_power_cap", 0x4), ("hw_slowdown", 0x8), ("sw_thermal_slowdown", 0x20),
               ("hw_thermal_slowdown", 0x40))

    def __init__(self, uuid, index):
        super().__init__(daemon=True)
        self.samples, self.stop_flag, self.error = [], threading.Event(), None
        self.max_mhz = None
        self.nvml = self.handle = None
        try:
            nvml = ctypes.CDLL("libnvidia-ml.so.1")
            if nvml.nvmlInit_v2() != 0:
                raise OSError("nvmlInit_v2 failed")
            handle = ctypes.c_void_p()
            rc = 1
            if uuid:
                rc = nvml.nvmlDeviceGetHandleByUUID(str(uuid).encode(), ctypes.byref(handle))
                if rc != 0 and not str(uuid).startswith("GPU-"):
                    rc = nvml.nvmlDeviceGetHandleByUUID(("GPU-" + str(uuid)).encode(),
                                                        ctypes.byref(handle))
            if rc != 0:
                visible = os.environ.get("CUDA_VISIBLE_DEVICES")
                if visible and all(v.strip().isdigit() for v in visible.split(",")):
                    index = int(visible.split(",")[index])
                rc = nvml.nvmlDeviceGetHandleByIndex_v2(ctypes.c_uint(index), ctypes.byref(handle))
            if rc != 0:
                raise OSError(f"no NVML handle (rc {rc})")
            mhz = ctypes.c_uint()
            if nvml.nvmlDeviceGetMaxClockInfo(handle, 1, ctypes.byref(mhz)) == 0:  # NVML_CLOCK_SM
                self.max_mhz = float(mhz.value)
            self.nvml, self.handle = nvml, handle
        except (OSError, AttributeError) as exc:
            self.error = str(exc)[:120]

    def sample(self):
        mhz, mask = ctypes.c_uint(), ctypes.c_ulonglong()
        if self.nvml.nvmlDeviceGetClockInfo(self.handle, 1, ctypes.byref(mhz)) != 0:
            return
        try:
            rc = self.nvml.nvmlDeviceGetCurrentClocksEventReasons(self.handle, ctypes.byref(mask))
        except AttributeError:
            rc = self.nvml.nvmlDeviceGetCurrentClocksThrottleReasons(self.handle,
                                                                     ctypes.byref(mask))
        self.samples.append((mhz.value, mask.value if rc == 0 else 0))

    def run(self):
        if self.nvml is None:
            return
        while not self.stop_flag.is_set():
            self.sample()
            self.stop_flag.wait(0.0005)

    def summary(self):
        self.stop_flag.set()
        if self.is_alive():
            self.join(timeout=2)
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz,
                    "reasons": ["unsampled: " + (self.error or "no samples")]}
        reasons = [name for name, bit in self.REASONS if any(m & bit for _, m in self.samples)]
        return {"sm_mhz": float(np.median([c for c, _ in self.samples])),
                "sm_max_mhz": self.max_mhz, "reasons": reasons, "samples": len(self.samples)}


def hbm_peak():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path, encoding="utf-8") as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def ncu_traffic(name="fused_step_traffic.json"):
    """dram bytes per launch of the fused kernel from the committed ncu capture, if any"""
    path = os.path.join(ROOT, "profiles", name)
    if os.path.exists(path):
        with open(path, encoding="utf-8") as f:
            return json.load(f).get("dram_bytes_per_launch")
    return None


# ---- GPU arm ---------------------------------------------------------------------------------
def gpu_arm(args):  # pylint: disable=too-many-locals,too-many-statements
    import torch  # pylint: disable=import-outside-toplevel
    import torch.distributed as dist  # pylint: disable=import-outside-toplevel

    from pysdm_b200.collision import DeviceCollision, Golovin  # pylint: disable=import-outside-toplevel

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    # stdout carries exactly one JSON line: whatever native libraries write to fd 1 during the
    # run (NCCL prints its version banner there) goes to stderr instead
    sys.stdout.flush()
    stdout_fd = os.dup(1)
    os.dup2(2, 1)
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the product has no CPU fallback); "
                         "use --impl reference for the CPU arm")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def gather(value):
        """the value of every rank, on every rank"""
        if world == 1:
            return [float(value)]
        t = torch.zeros(world, dtype=torch.float64, device="cuda")
        t[rank] = value
        dist.all_reduce(t)
        return [float(x) for x in t.tolist()]

    n_sd = 1 << args.log2_n_sd
    pop = population(n_sd, args.sd_per_cell, rank)
    sim = DeviceCollision(kernel=Golovin(GOLOVIN_B), dt=1.0, dv=1.0, seed=44 + rank, fused=True,
                          **pop)
    warmup = max(args.warmup, 3)
    for _ in range(warmup):
        sim.step_async()
    sim.drain()
    assert sim.n_fused_steps > 0

    # ---- timed region: K steps, device-resident state, CUDA events on the launch stream ----
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
          for _ in range(args.steps)]
    t_beg, t_end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    sampler = ClockSampler(getattr(torch.cuda.get_device_properties(local), "uuid", None), local)
    launches0 = sim.n_launches
    sampler.start()
    barrier()
    t_beg.record()
    for k in range(args.steps):
        ev[k][0].record()
        sim.step_async(mark=ev[k][1])  # the launch; `healthy` flags are read behind it
    sim.drain()              # `healthy` of the last steps (the reference reads it every step)
    t_end.record()
    barrier()
    n_clock_samples = len(sampler.samples)  # (an NVML query takes milliseconds on this box: the
    # sampler keeps running through the e2e and transport regions below, all under load)
    my_ms = t_beg.elapsed_time(t_end)
    kernel_ms = float(np.mean([a.elapsed_time(b) for a, b in ev]))
    launches = sim.n_launches - launches0
    per_rank = gather(my_ms)
    per_rank_kernel = gather(kernel_ms)
    elapsed_ms = max(per_rank)
    value = world * n_sd * args.steps / (elapsed_ms * 1e-3)

    peak, peak_src = hbm_peak()
    achieved = BYTES_PER_SD * n_sd / (kernel_ms * 1e-3) / 1e9
    traffic = ncu_traffic()
    line = {
        "metric": "SD-steps/s, SDM collision step", "value": value, "unit": "SD-steps/s",
        "n_gpus": world, "steps": args.steps, "warmup": warmup,
        "ms_per_step": elapsed_ms / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": workload_name(args), "n_sd_per_gpu": n_sd,
                   "n_cell_per_gpu": pop["n_cell"], "rng": "in-kernel Philox4x32-10",
                   "l2": "working set ~0.8 GB per GPU >> 126 MB L2, no flush needed",
                   "sharding": "cells sharded over ranks, no data-path collective",
                   "speculation_depth": sim.speculation_depth},
        "clocks": None, "gpu_launches": launches,
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                     "frac": achieved / peak, "traffic": traffic,
                     "kernel": "k_fused_step<Golovin, Philox>", "kernel_ms": kernel_ms,
                     "algorithmic_bytes_per_launch": BYTES_PER_SD * n_sd, "peak_source": peak_src,
                     "frac_of_nominal_8TBs": achieved / 8000.0,
                     "achieved_dram": None if traffic is None else traffic / (kernel_ms * 1e-3) / 1e9,
                     "frac_dram": None if traffic is None else traffic / (kernel_ms * 1e-3) / 1e9 / peak,
                     "limiter": "instruction issue / latency, not the DRAM pipe (ncu, profiles/): "
                                "`frac` counts the contract's algorithmic bytes, `frac_dram` the "
                                "bytes that actually crossed the HBM interface"},
        "per_rank": {"elapsed_ms": per_rank, "min_ms": min(per_rank),
                     "median_ms": float(np.median(per_rank)), "max_ms": max(per_rank),
                     "kernel_ms": per_rank_kernel},
    }
    del sim
    torch.cuda.empty_cache()

    # ---- e2e: host buffers in, host buffers out, every step (see DESIGN.md "measurement") ----
    if not args.no_e2e:
        e2e = e2e_plugin(args, pop, rank, barrier, gather, stripes=args.e2e_plugin_stripes)
        single = (e2e_plugin(args, pop, rank, barrier, gather, stripes=1)
                  if e2e is not None and args.e2e_plugin_stripes > 1 else None)
        striped = e2e_striped(args, pop, rank, barrier, gather)
        if e2e is None:  # no PySDM on this box: the repo's own host-facing API is the e2e path
            e2e, striped, single = striped, None, None
        line["e2e"] = e2e
        if single is not None:
            line["e2e_single_particulator"] = single
        if striped is not None:
            line["e2e_striped"] = striped
    if not args.no_transport:  # (N = 1: the same loop with one slab, the weak-scaling base line)
        line["transport"] = transport_run(args, rank, world, barrier, gather)
    line["clocks"] = sampler.summary()
    line["clocks"]["samples_in_first_timed_region"] = n_clock_samples
    line["clocks"]["window"] = ("value + e2e + transport timed regions (GPU under load "
                                "throughout); sampled every 0.5 ms through NVML")
    if rank == 0 and world == 1:
        if not args.no_cpu_baseline:
            r = cpu_run(args, 3, 1, min(args.cpu_log2_n_sd, args.log2_n_sd))
            line["cpu_baseline"] = {k: r[k] for k in ("value", "unit", "cores", "kind", "sample")}
            if not args.no_numba:
                line["cpu_baseline"]["numba_backend"] = numba_run(
                    args, min(args.cpu_log2_n_sd, args.log2_n_sd))
        if not args.no_configs:
            sys.path.insert(0, os.path.join(ROOT, "profiles", "tools"))
            import bench_configs  # pylint: disable=import-outside-toplevel,import-error

            try:
                line["configs"] = bench_configs.all_configs(args.config_steps, peak,
                                                            numba=not args.no_numba)
                line["pysdm_path"] = bench_configs.pysdm_path(args.config_steps, peak)
            except Exception as exc:  # pylint: disable=broad-except
                line["configs"] = {"unavailable": f"{type(exc).__name__}: {exc}"[:300]}
    if rank == 0:
        sys.stdout.flush()
        os.dup2(stdout_fd, 1)
        print(json.dumps(line), flush=True)
        os.dup2(2, 1)
    if world > 1:
        dist.destroy_process_group()


def e2e_plugin(args, pop, rank, barrier, gather, stripes=1):
    """the metric end to end through the reference-facing plug-in call: `Particulator`s built by
    the unmodified `Builder(n_sd, backend=B200())` with the collision dynamic; every step the
    super-droplet state (multiplicity, water mass) is uploaded from pinned host memory with
    `Storage.upload`, `particulator.run(1)` executes the step, and multiplicity, water mass and
    the per-cell coalescence counters are read back into pinned host memory with
    `Storage.download`.  (`cell id` is not re-sent: it does not change in this workload.)
    `stripes` > 1: the cells are cut into that many contiguous stripes, one Particulator each
    (cells do not interact), each driven under its own CUDA stream -- the upload of one stripe
    then overlaps the read-back of another (PCIe is full duplex); same bytes, same work."""
    if not pysdm_paths():
        return None
    import torch  # pylint: disable=import-outside-toplevel
    from PySDM import Builder, Formulae  # pylint: disable=import-outside-toplevel
    from PySDM.dynamics.collisions.collision_kernels import Golovin  # pylint: disable=import-outside-toplevel
    from PySDM.environments import Box  # pylint: disable=import-outside-toplevel
    from PySDM.impl.mesh import Mesh  # pylint: disable=import-outside-toplevel

    from pysdm_b200 import B200, FusedCoalescence  # pylint: disable=import-outside-toplevel

    n_sd, n_cell = len(pop["multiplicity"]), pop["n_cell"]
    sd_per_cell = n_sd // n_cell
    parts = []
    for s in range(stripes):
        c0, c1 = n_cell * s // stripes, n_cell * (s + 1) // stripes
        lo, hi = c0 * sd_per_cell, (n_sd if s == stripes - 1 else c1 * sd_per_cell)
        env = Box(dv=1.0, dt=1.0)
        env.mesh = Mesh((c1 - c0,), size=(c1 - c0,), dv=1.0)
        builder = Builder(n_sd=hi - lo, backend=B200(Formulae(seed=1000 * (rank + 1) + s)),
                          environment=env)
        builder.add_dynamic(FusedCoalescence(collision_kernel=Golovin(b=GOLOVIN_B),
                                             adaptive=False, croupier="local"))
        particulator = builder.build(products=(), attributes={
            "multiplicity": pop["multiplicity"][lo:hi],
            "volume": pop["water_mass"][lo:hi] / RHO_W, "cell id": pop["cell_id"][lo:hi] - c0})
        attrs, dyn = particulator.attributes, particulator.dynamics["Collision"]
        # the host model's state: two pinned buffer sets, a step reads one and writes the other
        host = [(torch.from_numpy(pop["multiplicity"][lo:hi].copy()).pin_memory(),
                 torch.from_numpy(pop["water_mass"][lo:hi].copy()).reshape(1, -1).pin_memory())
                for _ in range(2)]
        parts.append({"particulator": particulator, "attrs": attrs, "dyn": dyn, "host": host,
                      "mult": attrs["multiplicity"],
                      "ext": attrs.get_extensive_attribute_storage(),
                      "rate": torch.empty(c1 - c0, dtype=torch.int64).pin_memory(),
                      "stream": torch.cuda.Stream() if stripes > 1 else torch.cuda.current_stream()})
    main = torch.cuda.current_stream()

    def step():
        for p in parts:  # every stripe's upload is queued before anything waits
            p["stream"].wait_stream(main)
            with torch.cuda.stream(p["stream"]):
                p["mult"].upload(p["host"][0][0])
                p["ext"].upload(p["host"][0][1])
        for p in parts:
            with torch.cuda.stream(p["stream"]):
                for key in ("multiplicity", *p["attrs"].get_extensive_attribute_keys()):
                    p["attrs"].mark_updated(key)
                p["particulator"].run(1)
                p["mult"].download(p["host"][1][0], non_blocking=True)
                p["ext"].download(p["host"][1][1], non_blocking=True)
                p["dyn"].coalescence_rate.download(p["rate"], non_blocking=True)
            p["host"].reverse()  # the host model carries the state on
        for p in parts:  # the step's results are in host memory when step() returns
            p["stream"].synchronize()
            main.wait_stream(p["stream"])

    for _ in range(3):
        step()
    beg, end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    beg.record()
    for _ in range(args.e2e_steps):
        step()
    end.record()
    barrier()
    per_rank = gather(beg.elapsed_time(end))
    ms = max(per_rank)
    assert all(p["dyn"].n_fused_steps >= args.e2e_steps and int(p["rate"].sum()) > 0 for p in parts)
    world = len(per_rank)
    del parts
    torch.cuda.empty_cache()
    return {"value": world * n_sd * args.e2e_steps / (ms * 1e-3), "unit": "SD-steps/s",
            "h2d_bytes_per_step": n_sd * 16, "d2h_bytes_per_step": n_sd * 16 + n_cell * 8,
            "steps": args.e2e_steps, "ms_per_step": ms / args.e2e_steps,
            "per_rank_ms": per_rank, "stripes": stripes,
            "what": "plug-in: Builder(backend=B200()) + FusedCoalescence"
                    + (f", {stripes} Particulators over {stripes} stripes of cells, one CUDA "
                       "stream each (upload of one stripe overlaps the read-back of another)"
                       if stripes > 1 else "")
                    + "; per step Storage.upload of multiplicity and water mass from pinned host "
                      "memory, Particulator.run(1), Storage.download of multiplicity, water mass "
                      "and the coalescence counters into pinned host memory; PCIe-bound"}


def e2e_striped(args, pop, rank, barrier, gather):
    """the same metric through the repo's own host-facing API
    (pysdm_b200.host_step.HostCollisionStep): every step uploads the super-droplet state
    (multiplicity, water mass; the cell ids once -- no droplet changes cell in this workload,
    as in the plug-in leg) from pinned host memory, runs sort-by-cell + the
    collision step, and reads multiplicity, water mass and the coalescence counters back to
    pinned host memory; cells are cut into stripes so that upload, compute and download of
    different stripes overlap"""
    import torch  # pylint: disable=import-outside-toplevel

    from pysdm_b200.collision import Golovin  # pylint: disable=import-outside-toplevel
    from pysdm_b200.host_step import HostCollisionStep  # pylint: disable=import-outside-toplevel

    host = HostCollisionStep(
        multiplicity=pop["multiplicity"], attributes=pop["water_mass"].reshape(1, -1),
        cell_id=pop["cell_id"], n_cell=pop["n_cell"], n_stripes=args.e2e_stripes,
        seed=1000 * (rank + 1), kernel=Golovin(GOLOVIN_B), dt=1.0, dv=1.0, fused=True)
    host.step()  # (cell ids go up once: this workload moves no droplet between cells)
    for _ in range(3):
        host.step(cell_id_changed=False)
    launches0 = host.n_launches
    beg, end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    beg.record()
    for _ in range(args.e2e_steps):
        host.step(cell_id_changed=False)
    end.record()
    barrier()
    per_rank = gather(beg.elapsed_time(end))
    ms = max(per_rank)
    n_sd = host.n_sd
    out = {"value": len(per_rank) * n_sd * args.e2e_steps / (ms * 1e-3), "unit": "SD-steps/s",
           "h2d_bytes_per_step": host.h2d_bytes, "d2h_bytes_per_step": host.d2h_bytes,
           "steps": args.e2e_steps, "ms_per_step": ms / args.e2e_steps,
           "gpu_launches": host.n_launches - launches0, "stripes": len(host.stripes),
           "per_rank_ms": per_rank,
           "what": "HostCollisionStep: pinned host state (multiplicity, water mass; cell ids "
                   "unchanged, not re-sent) -> H2D -> sort by cell + fused collision step -> "
                   "D2H, cells in stripes over 3 streams"}
    del host
    torch.cuda.empty_cache()
    return out


def transport_run(args, rank, world, barrier, gather):
    """BASELINE configs[3] in kind, sharded over the ranks: see profiles/tools/transport_loop.py"""
    sys.path.insert(0, os.path.join(ROOT, "profiles", "tools"))
    import transport_loop  # pylint: disable=import-outside-toplevel,import-error

    try:
        return transport_loop.bench_entry(rank, world, args.transport_steps, barrier, gather)
    except Exception as exc:  # pylint: disable=broad-except
        return {"unavailable": f"{type(exc).__name__}: {exc}"[:300]}


def main():
    args = parse()
    if args.impl == "reference":
        reference_arm(args)
    else:
        gpu_arm(args)


if __name__ == "__main__":
    main()
