"""bench.py -- SD-steps/s of the SDM collision step (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]
    python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N ...

Workload (BASELINE.json configs[4], the one the metric is quoted
on): multi-cell Golovin coalescence, 2^24 super-droplets per GPU, 128 per cell (2^17 cells
per GPU), non-adaptive, "local" croupier, A = 1 extensive attribute, dt = 1 s, dv = 1 m^3.
A "step" is one model time step of the Collision dynamic on the whole population of every
rank.  Cells are sharded over ranks (weak scaling); the collision step needs no
communication -- the only collective is the timing barrier/max.

One JSON line on stdout (rank 0).  `--impl reference` times the CPU restatement of the
reference's Numba path (oracle/, kind "port": the reference is pure Python + Numba and
cannot travel to the GPU box) on the host cores instead.
"""

import argparse
import json
import os
import subprocess
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
                    help="bounded CPU sample: super-droplets per CPU step (log2)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-numba", action="store_true",
                    help="skip timing the reference's own Numba backend next to the port")
    ap.add_argument("--e2e-steps", type=int, default=10)
    ap.add_argument("--e2e-stripes", type=int, default=4)
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


# ---- CPU arm: the oracle port on the host cores -------------------------------------------
def cpu_run(args, steps, warmup):
    """times OracleCollision (C restatement of the Numba path, OpenMP where the reference
    uses prange) on a bounded sample: 2^cpu_log2_n_sd SDs of the same workload per step.
    The OpenMP thread count is swept (1, 2, 4, ... all cores) on one step each and the best
    is used: on shared/virtualised hosts "all threads" can be far slower than a few."""
    from oracle import sdm_oracle as orc  # pylint: disable=import-outside-toplevel
    from oracle.sdm_step import Golovin, OracleCollision  # pylint: disable=import-outside-toplevel

    n_sd = 1 << min(args.cpu_log2_n_sd, args.log2_n_sd)
    pop = population(n_sd, args.sd_per_cell, 0)
    sim = OracleCollision(kernel=Golovin(GOLOVIN_B), dt=1.0, dv=1.0, seed=44, **pop)
    ncpu = os.cpu_count() or 1
    counts = sorted({min(1 << k, ncpu) for k in range(0, 12)} | {ncpu})
    sweep = {}
    for _ in range(max(warmup, 1)):
        sim()
    for c in counts:
        orc.set_num_threads(c)
        sim()
        t0 = time.perf_counter()
        sim()
        sweep[c] = time.perf_counter() - t0
        if sweep[c] > 4 * min(sweep.values()):
            break  # oversubscribed: more threads only get slower
    cores = min(sweep, key=sweep.get)
    orc.set_num_threads(cores)
    t0 = time.perf_counter()
    for _ in range(steps):
        sim()
    dt = time.perf_counter() - t0
    return {
        "value": n_sd * steps / dt, "unit": "SD-steps/s", "cores": cores, "kind": "port",
        "sample": f"{steps} steps of 2^{int(np.log2(n_sd))} SDs ({args.sd_per_cell}/cell) of the "
                  f"same workload; C port of the Numba path (oracle/), numpy PCG64 RNG included; "
                  f"best of OpenMP thread sweep {sorted(sweep)} = {cores} threads on "
                  f"{ncpu} host cores",
        "ms_per_step": dt / steps * 1e3,
    }


def numba_run(args, steps=2):
    """the reference's OWN Numba CPU backend (unmodified PySDM from baseline/_ref, its stock
    Builder / Particulator / Coalescence path) on the same bounded sample, when it is importable
    on this box (numba + the install under baseline/_ref); reported next to the port, never used
    as the denominator of anything.  Returns a dict, or {"unavailable": why}."""
    ref = next((p for p in (os.path.join(ROOT, "baseline", "_ref"), "/root/reference")
                if os.path.isdir(os.path.join(p, "PySDM"))), None)
    if ref is None:
        return {"unavailable": "no PySDM install under baseline/_ref"}
    try:
        os.environ.setdefault("CI", "1")  # PySDM's own switch: do not import ThrustRTC
        sys.path.insert(0, os.path.join(ROOT, "oracle", "refshims"))  # pint / chempy stand-ins
        sys.path.insert(1, ref)
        np.NaN = np.nan  # the reference predates NumPy 2
        import numba  # pylint: disable=import-outside-toplevel
        from PySDM import Builder, Formulae  # pylint: disable=import-outside-toplevel
        from PySDM.backends import CPU  # pylint: disable=import-outside-toplevel
        from PySDM.dynamics import Coalescence  # pylint: disable=import-outside-toplevel
        from PySDM.dynamics.collisions.collision_kernels import Golovin  # pylint: disable=import-outside-toplevel
        from PySDM.environments import Box  # pylint: disable=import-outside-toplevel
        from PySDM.impl.mesh import Mesh  # pylint: disable=import-outside-toplevel

        n_sd = 1 << min(args.cpu_log2_n_sd, args.log2_n_sd)
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
                "sample": f"{steps} steps of 2^{int(np.log2(n_sd))} SDs ({args.sd_per_cell}/cell), "
                          f"PySDM Numba backend (baseline/_ref, stock Builder/Coalescence path), "
                          f"best of {sorted({1, ncpu})} numba threads"}
    except Exception as exc:  # pylint: disable=broad-except
        return {"unavailable": f"{type(exc).__name__}: {exc}"[:200]}


def reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    steps, warmup = max(1, min(args.steps, 5)), max(1, min(args.warmup, 2))
    r = cpu_run(args, steps, warmup)
    line = {
        "impl": "reference", "metric": "SD-steps/s, SDM collision step", "value": r["value"],
        "unit": "SD-steps/s", "n_gpus": args.gpus, "steps": steps, "warmup": warmup,
        "ms_per_step": r["ms_per_step"], "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": workload_name(args), "timed": r["sample"]},
        "cpu_baseline": {k: r[k] for k in ("value", "unit", "cores", "kind", "sample")},
        "e2e": {"value": r["value"], "unit": "SD-steps/s", "h2d_bytes_per_step": 0,
                "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    if not args.no_numba:
        line["cpu_baseline"]["numba_backend"] = numba_run(args)
    print(json.dumps(line))


# ---- GPU arm ---------------------------------------------------------------------------------
class ClockSampler(threading.Thread):
    """samples SM clocks / throttle reasons with nvidia-smi while the timed region runs"""

    QUERY = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.stop_flag = index, [], threading.Event()

    def run(self):
        if self._run_nvml():
            return
        while not self.stop_flag.is_set():
            try:
                out = subprocess.run(
                    ["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.QUERY}",
                     "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5)
                self.samples.append([x.strip() for x in out.stdout.strip().split(",")])
            except Exception:  # pylint: disable=broad-except
                pass
            self.stop_flag.wait(0.1)

    def _run_nvml(self):
        """NVML directly (a query takes ~0.1 ms, so a 25 ms timed region still yields a dozen
        samples; an nvidia-smi process takes longer than that to start); False = unavailable"""
        try:
            import pynvml  # pylint: disable=import-outside-toplevel

            pynvml.nvmlInit()
            visible = os.environ.get("CUDA_VISIBLE_DEVICES")
            index = int(visible.split(",")[self.index]) if visible else self.index
            handle = pynvml.nvmlDeviceGetHandleByIndex(index)
            max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(handle, pynvml.NVML_CLOCK_SM)
            bits = (("hw_slowdown", pynvml.nvmlClocksEventReasonHwSlowdown),
                    ("hw_thermal_slowdown", pynvml.nvmlClocksEventReasonHwThermalSlowdown),
                    ("sw_thermal_slowdown", pynvml.nvmlClocksEventReasonSwThermalSlowdown),
                    ("sw_power_cap", pynvml.nvmlClocksEventReasonSwPowerCap))
        except Exception:  # pylint: disable=broad-except
            return False
        while not self.stop_flag.is_set():
            try:
                mhz = pynvml.nvmlDeviceGetClockInfo(handle, pynvml.NVML_CLOCK_SM)
                mask = pynvml.nvmlDeviceGetCurrentClocksEventReasons(handle)
                self.samples.append([str(mhz), str(max_mhz)] +
                                    ["Active" if mask & bit else "Not Active" for _, bit in bits])
            except Exception:  # pylint: disable=broad-except
                pass
            self.stop_flag.wait(0.002)
        return True

    def summary(self):
        self.stop_flag.set()
        self.join(timeout=6)
        good = [s for s in self.samples if len(s) == 6 and s[0].isdigit()]
        if not good:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unsampled"]}
        names = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")
        reasons = [n for i, n in enumerate(names) if any(s[2 + i] == "Active" for s in good)]
        return {"sm_mhz": float(np.median([int(s[0]) for s in good])),
                "sm_max_mhz": float(good[0][1]), "reasons": reasons, "samples": len(good)}


def hbm_peak():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path, encoding="utf-8") as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def ncu_traffic():
    """dram bytes per launch of the fused kernel from the committed ncu capture, if any"""
    path = os.path.join(ROOT, "profiles", "fused_step_traffic.json")
    if os.path.exists(path):
        with open(path, encoding="utf-8") as f:
            return json.load(f).get("dram_bytes_per_launch")
    return None


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

    n_sd = 1 << args.log2_n_sd
    pop = population(n_sd, args.sd_per_cell, rank)
    sim = DeviceCollision(kernel=Golovin(GOLOVIN_B), dt=1.0, dv=1.0, seed=44 + rank, fused=True,
                          **pop)
    for _ in range(max(args.warmup, 3)):
        sim()
    assert sim.n_fused_steps > 0

    # ---- timed region: K steps, device-resident state, CUDA events on the launch stream ----
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
          for _ in range(args.steps)]
    t_beg, t_end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    sampler = ClockSampler(local)
    launches0 = sim.n_launches
    barrier()
    sampler.start()
    t_beg.record()
    for k in range(args.steps):
        ev[k][0].record()
        sim.step_async(mark=ev[k][1])  # the launch; `healthy` of step k-1 is read behind it
    sim.drain()              # `healthy` of the last step (the reference reads it every step)
    t_end.record()
    barrier()
    clocks = sampler.summary()
    elapsed_ms = t_beg.elapsed_time(t_end)
    kernel_ms = float(np.mean([a.elapsed_time(b) for a, b in ev]))
    launches = sim.n_launches - launches0
    if world > 1:
        t = torch.tensor([elapsed_ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        elapsed_ms = float(t.item())
    value = world * n_sd * args.steps / (elapsed_ms * 1e-3)

    # ---- e2e: host buffers in, host buffers out, every step (see DESIGN.md "measurement") ----
    e2e = None
    if not args.no_e2e:
        del sim
        torch.cuda.empty_cache()
        e2e = e2e_run(args, pop, rank, world, barrier)
    peak, peak_src = hbm_peak()
    achieved = BYTES_PER_SD * n_sd / (kernel_ms * 1e-3) / 1e9
    line = {
        "metric": "SD-steps/s, SDM collision step", "value": value, "unit": "SD-steps/s",
        "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
        "ms_per_step": elapsed_ms / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": workload_name(args), "n_sd_per_gpu": n_sd,
                   "n_cell_per_gpu": pop["n_cell"], "rng": "in-kernel Philox4x32-10",
                   "l2": "working set ~0.8 GB per GPU >> 126 MB L2, no flush needed",
                   "sharding": "cells sharded over ranks, no data-path collective"},
        "clocks": clocks, "gpu_launches": launches,
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                     "frac": achieved / peak, "traffic": ncu_traffic(),
                     "kernel": "k_fused_step<Golovin, Philox>", "kernel_ms": kernel_ms,
                     "algorithmic_bytes_per_launch": BYTES_PER_SD * n_sd, "peak_source": peak_src,
                     "frac_of_nominal_8TBs": achieved / 8000.0},
    }
    if e2e is not None:
        line["e2e"] = e2e
    if rank == 0:
        if not args.no_cpu_baseline and world == 1:
            r = cpu_run(args, 3, 1)
            line["cpu_baseline"] = {k: r[k] for k in ("value", "unit", "cores", "kind", "sample")}
            if not args.no_numba:
                line["cpu_baseline"]["numba_backend"] = numba_run(args)
        sys.stdout.flush()
        os.dup2(stdout_fd, 1)
        print(json.dumps(line), flush=True)
        os.dup2(2, 1)
    if world > 1:
        dist.destroy_process_group()


def e2e_run(args, pop, rank, world, barrier):
    """the same metric through the host-facing API (pysdm_b200.host_step.HostCollisionStep):
    every step uploads the super-droplet state (multiplicity, water mass, cell id) from pinned
    host memory, runs sort-by-cell + the collision step, and reads multiplicity, water mass
    and the coalescence counters back to pinned host memory; cells are cut into stripes so
    that upload, compute and download of different stripes overlap"""
    import torch  # pylint: disable=import-outside-toplevel
    import torch.distributed as dist  # pylint: disable=import-outside-toplevel

    from pysdm_b200.collision import Golovin  # pylint: disable=import-outside-toplevel
    from pysdm_b200.host_step import HostCollisionStep  # pylint: disable=import-outside-toplevel

    host = HostCollisionStep(
        multiplicity=pop["multiplicity"], attributes=pop["water_mass"].reshape(1, -1),
        cell_id=pop["cell_id"], n_cell=pop["n_cell"], n_stripes=args.e2e_stripes,
        seed=1000 * (rank + 1), kernel=Golovin(GOLOVIN_B), dt=1.0, dv=1.0, fused=True)
    for _ in range(3):
        host.step()
    launches0 = host.n_launches
    beg, end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    beg.record()
    for _ in range(args.e2e_steps):
        host.step()
    end.record()
    barrier()
    ms = beg.elapsed_time(end)
    if world > 1:
        t = torch.tensor([ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    n_sd = host.n_sd
    return {"value": world * n_sd * args.e2e_steps / (ms * 1e-3), "unit": "SD-steps/s",
            "h2d_bytes_per_step": host.h2d_bytes, "d2h_bytes_per_step": host.d2h_bytes,
            "steps": args.e2e_steps, "ms_per_step": ms / args.e2e_steps,
            "gpu_launches": host.n_launches - launches0, "stripes": len(host.stripes),
            "what": "HostCollisionStep: pinned host state -> H2D -> sort by cell + fused "
                    "collision step -> D2H, cells in stripes over 3 streams"}


def main():
    args = parse()
    if args.impl == "reference":
        reference_arm(args)
    else:
        gpu_arm(args)


if __name__ == "__main__":
    main()
