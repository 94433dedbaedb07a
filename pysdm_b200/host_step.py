"""Collision step over HOST-resident super-droplet state (the end-to-end entry point).

A host model that keeps its particle arrays in (pinned) host memory -- e.g. one that does
advection / condensation on the CPU, or stages them for I/O -- calls `HostCollisionStep.step()`:
multiplicity, extensive attributes and cell ids go to the GPU, the collision step of
`pysdm_b200.collision.DeviceCollision` runs (sort by cell + fused step, or the method-level
kernels when the fused step does not apply), and multiplicity, attributes and the per-cell
coalescence counters come back.

The step is PCIe-bound (40 B per super-droplet over the link, against 73 B per super-droplet
of HBM traffic at ~100x the bandwidth), so the only thing that matters is keeping both
directions of the link busy: cells are cut into `n_stripes` contiguous ranges ("stripes";
collisions never cross a cell), every stripe is an independent `DeviceCollision`,
and uploads (one stream), compute (one stream) and downloads (one stream) of different stripes
overlap.  Results are identical to running the stripes one after another.

Reference counterpart: `Particulator.run` -> `Collision.__call__` with numpy-backed Storage
(PySDM/particulator.py:46-56, PySDM/dynamics/collisions/collision.py:174-194), where the state
is host-resident by construction.
"""

import numpy as np
import torch

from .collision import DeviceCollision
from .sharding import cell_range


class HostCollisionStep:  # pylint: disable=too-many-instance-attributes
    """`multiplicity` i64[n], `attributes` f64[A, n] (row 0 = water mass), `cell_id` i64[n]:
    host arrays (numpy or torch CPU tensors; copied once into pinned memory, exposed as
    `.multiplicity`, `.attributes`, `.cell_id` for the caller to update in place between
    steps).  With n_stripes > 1 the super-droplets must be ordered by cell id."""

    def __init__(self, *, multiplicity, attributes, cell_id, n_cell, n_stripes=8, seed=0,
                 device="cuda", **collision_kw):
        if not torch.cuda.is_available():
            raise RuntimeError("pysdm_b200 needs a CUDA device (no CPU fallback)")

        def pinned(array, dtype):
            t = torch.as_tensor(np.ascontiguousarray(array) if isinstance(array, np.ndarray)
                                else array).to(dtype).contiguous()
            return t.clone().pin_memory()

        self.multiplicity = pinned(multiplicity, torch.int64)
        self.attributes = pinned(np.atleast_2d(attributes) if isinstance(attributes, np.ndarray)
                                 else attributes, torch.float64)
        self.cell_id = pinned(cell_id, torch.int64)
        self.n_sd, self.n_cell = self.multiplicity.numel(), int(n_cell)
        self.out_multiplicity = torch.empty_like(self.multiplicity).pin_memory()
        self.out_attributes = torch.empty_like(self.attributes).pin_memory()
        self.out_coalescence_rate = torch.empty(self.n_cell, dtype=torch.int64).pin_memory()
        n_stripes = max(1, min(int(n_stripes), self.n_cell))
        cid = self.cell_id.numpy()
        if n_stripes > 1 and self.n_sd > 1 and bool((np.diff(cid) < 0).any()):
            raise ValueError("n_stripes > 1 needs the super-droplets ordered by cell id")
        self.device = torch.device(device)
        self.s_in, self.s_compute, self.s_out = (torch.cuda.Stream(self.device) for _ in range(3))
        self.stripes = []
        for s in range(n_stripes):
            cb, ce = cell_range(self.n_cell, s, n_stripes)
            sb, se = (int(x) for x in np.searchsorted(cid, (cb, ce))) if n_stripes > 1 else \
                (0, self.n_sd)
            if se == sb:
                continue
            with torch.cuda.stream(self.s_compute):
                sim = DeviceCollision(
                    multiplicity=self.multiplicity[sb:se], water_mass=self.attributes[0, sb:se],
                    extra_attributes=[self.attributes[r, sb:se]
                                      for r in range(1, self.attributes.shape[0])],
                    cell_id=self.cell_id[sb:se] - cb, n_cell=ce - cb, seed=seed + s,
                    device=device, **collision_kw)
            self.stripes.append(_Stripe(sim, sb, se, cb, ce))
        torch.cuda.synchronize(self.device)
        self.h2d_bytes = self.n_sd * 8 * (2 + self.attributes.shape[0])
        self.d2h_bytes = self.n_sd * 8 * (1 + self.attributes.shape[0]) + self.n_cell * 8

    def step(self, cell_id_changed=True):
        """one model time step of the collision dynamic on the host arrays; returns when the
        outputs are in `out_multiplicity`, `out_attributes`, `out_coalescence_rate`.
        `cell_id_changed=False`: the host model did not move any super-droplet to another cell
        since the previous step (no transport between two collision steps) and did not revive
        removed ones: the cell ids on the device are kept -- a third of the upload -- and so is
        the permutation with its sort by cell, as `Particulator.run` keeps them from step to
        step (no identity permutation, no re-sort, no scan for zero multiplicities)"""
        self.h2d_bytes = self.n_sd * 8 * ((2 if cell_id_changed else 1) + self.attributes.shape[0])
        # every upload is queued up front: the copy engine works through them while the host
        # drives sort + step of the stripes that have arrived
        for st in self.stripes:
            sim, sl = st.sim, slice(st.sb, st.se)
            with torch.cuda.stream(self.s_in):
                sim.multiplicity.copy_(self.multiplicity[sl], non_blocking=True)
                sim.attributes.copy_(self.attributes[:, sl], non_blocking=True)
                if cell_id_changed:
                    sim.cell_id.copy_(self.cell_id[sl], non_blocking=True)
                st.arrived.record()
        for st in self.stripes:
            sim, sl = st.sim, slice(st.sb, st.se)
            with torch.cuda.stream(self.s_compute):
                self.s_compute.wait_event(st.arrived)
                if cell_id_changed:
                    if st.cb:
                        sim.cell_id.sub_(st.cb)  # stripe-local cell numbers
                    sim.reset_permutation()
                sim()
                st.computed.record()
            with torch.cuda.stream(self.s_out):
                self.s_out.wait_event(st.computed)
                self.out_multiplicity[sl].copy_(sim.multiplicity, non_blocking=True)
                self.out_attributes[:, sl].copy_(sim.attributes, non_blocking=True)
                self.out_coalescence_rate[st.cb:st.ce].copy_(sim.coalescence_rate,
                                                             non_blocking=True)
        self.s_out.synchronize()

    @property
    def n_launches(self):
        return sum(st.sim.n_launches for st in self.stripes)


class _Stripe:  # pylint: disable=too-few-public-methods
    def __init__(self, sim, sb, se, cb, ce):
        self.sim, self.sb, self.se, self.cb, self.ce = sim, sb, se, cb, ce
        self.arrived, self.computed = torch.cuda.Event(), torch.cuda.Event()
