"""GPU time line of the adaptive sub-step loop (CUPTI through torch.profiler): start, gap to the
previous kernel and duration of every kernel / copy of three time steps of `DeviceCollision` on
the configs[3] population (2^22 SDs in 32768 cells, Geometric adaptive, optimized_random, A = 3),
plus ms per step and per sub-step without the profiler.

    python profiles/tools/adaptive_timeline.py [speculation depth, default 2] > profiles/<name>.txt
"""
import os
import sys

import numpy as np
import torch
from torch.profiler import ProfilerActivity, profile

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from pysdm_b200.collision import DeviceCollision, Geometric  # noqa: E402


def make(n_cell=32768, sd_per_cell=128, seed=1):
    n0 = n_cell * sd_per_cell
    rng = np.random.default_rng(seed)
    dv, dt = (1500 / 512) ** 2, 5.0
    volume = 4 / 3 * np.pi * (10e-6 * np.exp(0.47 * rng.standard_normal(n0))) ** 3
    return DeviceCollision(
        multiplicity=np.full(n0, int(100e6 * dv / sd_per_cell), dtype=np.int64),
        water_mass=1000.0 * volume, cell_id=(np.arange(n0) // sd_per_cell).astype(np.int64),
        n_cell=n_cell, dt=dt, dv=dv, seed=7, fused=True, kernel=Geometric(1.0), adaptive=True,
        optimized_random=True, dt_coal_range=(0.1, dt),
        extra_attributes=[volume * 1e-3, volume * 0.6e-3])


def main():
    depth = int(sys.argv[1]) if len(sys.argv) > 1 else 2
    sim = make()
    sim.substep_speculation = depth
    for _ in range(4):
        sim()
    torch.cuda.synchronize()
    n0, steps = sim.n_fused_steps, 20
    beg, end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    beg.record()
    for _ in range(steps):
        sim()
    end.record()
    torch.cuda.synchronize()
    sub = (sim.n_fused_steps - n0) / steps
    ms = beg.elapsed_time(end) / steps
    print(f"speculation depth {depth}: {ms:.3f} ms per time step, {sub:.2f} sub-steps per step, "
          f"{ms / sub * 1e3:.1f} us per sub-step (CUDA events, no profiler)")
    with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
        for _ in range(3):
            sim()
        torch.cuda.synchronize()
    events = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
    events.sort(key=lambda e: e.time_range.start)
    t0 = previous = events[0].time_range.start
    print("     start        gap   duration  kernel (three time steps under the profiler)")
    for e in events:
        r = e.time_range
        print(f"{r.start - t0:10.1f} {r.start - previous:+9.1f} {r.end - r.start:9.1f}  {e.name[:70]}")
        previous = r.end


if __name__ == "__main__":
    main()
