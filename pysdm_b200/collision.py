"""Device-resident SDM collision step: state in HBM, one object = the collision part of a
`Particulator` (the `Coalescence` dynamic + the `ParticleAttributes` bookkeeping it drives).

It follows the reference's control flow call by call
  Collision.__call__/step/compute_gamma     PySDM/dynamics/collisions/collision.py:174-290
  ParticleAttributes (sort/permutation/sanitize/working length)   PySDM/impl/particle_attributes.py:42-111
  RandomGeneratorOptimizer                  PySDM/dynamics/impl/random_generator_optimizer.py:19-48
  Particulator.collision_coalescence_breakup   PySDM/particulator.py:157-213
so it can be used stand-alone (bench.py, tests on a box without PySDM) with results
identical to `Builder(...).build(...)` + `particulator.run()` under the reference backend.

Two execution modes over the same state:
  * method level -- one CUDA kernel (group) per reference backend method (pysdm_b200.methods);
  * fused        -- one launch per (sub-)step (csrc/fused*.cu) for Coalescence with the "local"
                    croupier and a Golovin / ConstantK / Linear kernel, adaptive time step
                    included, when every cell fits a CTA batch.  Selected automatically;
                    bit-identical to the first mode.
There is no CPU path.
"""

import math

import numpy as np
import torch

from . import methods as m
from .random import HostReplayRandom, Random


class Golovin:
    """collision_kernels/golovin.py:14-16"""
    fused = "Golovin"

    def __init__(self, b):
        self.a, self.b = 0.0, b

    def __call__(self, sim, output, flag):
        m.pair_op("sum", output, sim.volume(), flag, sim.idx, sim.length)
        m.elementwise("mul", output, output, None, self.b)


class ConstantK:
    """collision_kernels/constantK.py"""
    fused = "ConstantK"

    def __init__(self, a):
        self.a, self.b = a, 0.0

    def __call__(self, sim, output, flag):
        output.fill_(self.a)


class Linear:
    """collision_kernels/linear.py"""
    fused = "Linear"

    def __init__(self, a, b):
        self.a, self.b = a, b

    def __call__(self, sim, output, flag):
        m.pair_op("sum", output, sim.volume(), flag, sim.idx, sim.length)
        m.elementwise("mul", output, output, None, self.b)
        m.elementwise("add", output, output, None, self.a)


class Geometric:
    """collision_kernels/geometric.py:15-22 with the Gunn-Kinzer fall velocity (PySDM's default
    `terminal_velocity`): `sum_pair(radius)`, `**= 2`, `*= PI * collection_efficiency`,
    `distance_pair(relative fall velocity)`, `*=`.  Radius and fall velocity are per-SD arrays
    derived from the water mass (attributes/physics/{volume,radius,terminal_velocity}.py); the
    simulation object owns them (`DeviceCollision.radius`, `.fall_velocity`) -- computed in one
    pass up front, then kept current by the fused step for the super-droplets a coalescence
    changed, or recomputed in full before every step of the method-level path."""
    fused = "Geometric"
    needs_dependents = True
    PI = 3.141592653589793
    PI_4_3 = PI * 4 / 3  # PySDM/physics/constants_defaults.py

    def __init__(self, collection_efficiency=1.0):
        self.a, self.b = self.PI * collection_efficiency, 0.0

    def __call__(self, sim, output, flag):
        m.pair_op("sum", output, sim.radius, flag, sim.idx, sim.length)
        m.elementwise("pow", output, output, None, 2.0)
        m.elementwise("mul", output, output, None, self.a)
        m.pair_op("distance", sim.pair_tmp, sim.fall_velocity, flag, sim.idx, sim.length)
        m.elementwise("mul", output, output, sim.pair_tmp)


def gunn_kinzer_table(device):
    """the reference's interpolation table (pysdm_b200/data, written by oracle/gen_gk_table.py
    from the reference object) on the device"""
    import os  # pylint: disable=import-outside-toplevel

    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "data", "gunn_kinzer_1949.npz")
    with np.load(path) as table:
        return {"table_b": torch.from_numpy(table["a"]).to(device),
                "table_c": torch.from_numpy(table["b"]).to(device),
                "table_factor": float(table["factor"]),
                "radius_max": float(table["maximum_radius"])}


class DeviceCollision:  # pylint: disable=too-many-instance-attributes
    """`Coalescence(collision_kernel=kernel, ...)` on a Box or a mesh of `n_cell` cells"""

    def __init__(self, *, multiplicity, water_mass, cell_id=None, n_cell=1, dt, dv, kernel, seed,
                 adaptive=False, substeps=1, dt_coal_range=(0.1, 100.0), croupier="local",
                 optimized_random=False, rho_w=1000.0, extra_attributes=None, rnd=None,
                 host_rng=False, fused="auto", device="cuda"):
        assert substeps == 1 or adaptive is False
        if not torch.cuda.is_available():
            raise RuntimeError("pysdm_b200 needs a CUDA device (no CPU fallback)")
        dev = torch.device(device)
        self.device = dev
        self.n_sd = len(multiplicity)
        self.n_cell = int(n_cell)
        self.dt, self.dv, self.kernel, self.rho_w = dt, dv, kernel, rho_w
        self.adaptive, self.substeps, self.croupier = adaptive, substeps, croupier
        self.dt_coal_range = (dt_coal_range[0], min(dt_coal_range[1], dt))

        def up(array, dtype):
            if isinstance(array, torch.Tensor):
                return array.to(device=dev, dtype=dtype).contiguous().clone()
            return torch.from_numpy(np.ascontiguousarray(array)).to(device=dev, dtype=dtype)

        def full(n, value, dtype):
            return torch.full((n,), value, dtype=dtype, device=dev)

        i64, f64 = torch.int64, torch.float64
        self.multiplicity = up(multiplicity, i64)
        rows = [water_mass] + list(extra_attributes or [])
        self.attributes = torch.stack([up(r, f64) for r in rows]).contiguous()
        self.cell_id = full(self.n_sd, 0, i64) if cell_id is None else up(cell_id, i64)
        # ParticleAttributes.__init__ (impl/particle_attributes.py:24-40)
        self.idx = torch.arange(self.n_sd, dtype=i64, device=dev)
        self.tmp_idx = full(self.n_sd, -1, i64)
        self.length = self.n_sd
        self.valid_n_sd = self.n_sd
        self.cell_idx = torch.arange(self.n_cell, dtype=i64, device=dev)
        self.cell_start = full(self.n_cell + 1, 0, i64)
        self.sorted = False
        self._cell_idx_identity, self._sort_identity = True, False
        self.sorted_length = 0
        self.max_cell = None
        # words the host reads back after a step: [0] healthy, [1:3] the fused adaptive
        # sub-step's adaptive_sdm_end result (pysdm_b200.h `adaptive_end`)
        # healthy, end, end, status, end, end, end, end, gate
        self._words = full(9, 0, i64)
        self.healthy = self._words[0:1]
        self.healthy.fill_(1)
        self.adaptive_end = self._words[1:3]
        self.rank_to_cell = torch.arange(self.n_cell, dtype=i64, device=dev)
        self.rank_info = full(self.n_cell, 0, i64)
        # Collision.register (collision.py:101-172)
        n_pairs = self.n_sd // 2
        nan = float("nan")
        self.kernel_temp = full(n_pairs, nan, f64)
        self.gamma = full(n_pairs, nan, f64)
        self.flag = torch.ones(self.n_sd, dtype=torch.bool, device=dev)
        self.norm_factor = full(self.n_cell, nan, f64)
        self.dt_left = full(self.n_cell, nan, f64)
        self.stats_n_substep = full(self.n_cell, 0 if adaptive else substeps, i64)
        self.stats_dt_min = full(self.n_cell, nan, f64)
        self.collision_rate = full(self.n_cell, 0, i64)
        self.collision_rate_deficit = full(self.n_cell, 0, i64)
        self.coalescence_rate = full(self.n_cell, 0, i64)
        self.status = self._words[3:4]  # fused step: a batch of cells did not fit (wrong hint)
        self.nf_slot = full(n_pairs, nan, f64)  # fused step: normalize factor per pair slot
        self.nf_valid = False  # False after anything that changes cell_start/cell_idx/cell_id
        self.norm_factor_valid = False  # norm_factor[] matches cell_start (adaptive presort)
        # RandomGeneratorOptimizer.register
        self.optimized_random = optimized_random
        shift = math.ceil(dt / dt_coal_range[0]) if optimized_random else 0
        self.pairs_rand = full(self.n_sd + shift, nan, f64)
        self.rand = full(n_pairs, nan, f64)
        if rnd is None:
            rnd = (HostReplayRandom if host_rng else Random)(self.n_sd + shift, seed)
        self.rnd = rnd
        self.substep = 0
        self._volume = full(self.n_sd, nan, f64)
        self.n_launches = 0  # CUDA launches issued through the C-ABI (entry-point calls)
        # gravitational kernels: radius and fall velocity per SD (see `Geometric`)
        self.dependents = None
        if getattr(kernel, "needs_dependents", False):
            self.radius, self.fall_velocity = full(self.n_sd, nan, f64), full(self.n_sd, nan, f64)
            self.pair_tmp = full(n_pairs, nan, f64)
            self.dependents = {"volume": self._volume, "radius": self.radius,
                               "velocity": self.fall_velocity,
                               "radius_scale": 1 / kernel.PI_4_3, "radius_exponent": 1 / 3,
                               "status": full(1, 0, i64), **gunn_kinzer_table(dev)}
            self.refresh_dependents()
        self.dt_min_warning = False
        self.fused = fused
        self.n_fused_steps = 0
        self._epoch = 0
        self._launch_cache = {}  # argument block of the fused launch (every array is allocated once)
        self._healthy_host = torch.ones(1, dtype=i64).pin_memory()
        self._words_host = torch.zeros(9, dtype=i64).pin_memory()
        # speculative stepping (step_async): gate word of the fused kernel, a ring of pinned
        # slots for the `healthy` flags in flight, the steps whose flags have not been read yet
        self.gate = self._words[8:9]
        self.speculation_depth = 8
        self.substep_speculation = 2  # adaptive loop: sub-steps launched ahead of the read-back
        self._spec_host = self._cell_idx_ring = None
        self._healthy_slots = torch.ones(self.speculation_depth + 1, dtype=i64).pin_memory()
        self._in_flight = []  # (event, slot), oldest first
        self._slot = 0
        # Builder.build: zero multiplicities are removed up front
        if bool((self.multiplicity == 0).any().item()):
            self.healthy.zero_()
            self.sanitize()

    # ---- attributes -------------------------------------------------------------------
    def volume(self):
        m.volume_of_water_mass(self._volume, self.attributes[0], self.rho_w)
        return self._volume

    def refresh_dependents(self):
        """volume, radius and fall velocity of every super-droplet from the water mass (what the
        orchestration recomputes when the water mass was marked updated), one pass"""
        d = self.dependents
        m.dependent_attributes(self.attributes[0], self.rho_w, volume=d["volume"],
                               radius=d["radius"], velocity=d["velocity"], status=d["status"],
                               **{k: d[k] for k in ("radius_scale", "radius_exponent",
                                                    "table_factor", "radius_max", "table_b",
                                                    "table_c")})
        self.n_launches += 1

    def sort_by_cell_id(self):
        m.counting_sort_by_cell_id(self.tmp_idx, self.idx, self.cell_id, self.cell_idx,
                                   self.length, self.cell_start)
        self.idx, self.tmp_idx = self.tmp_idx, self.idx
        m.invert_index(self.rank_to_cell, self.cell_idx)  # the cell each rank of cell_start holds
        self.sorted = True
        self._sort_identity = self._cell_idx_identity
        self.sorted_length = self.length  # = cell_start[-1], kept on the host
        self.max_cell = None
        self.nf_valid = self.norm_factor_valid = False

    def reset_permutation(self):
        """after the caller has overwritten multiplicity / attributes / cell_id in place:
        identity permutation over all super-droplets, cell order to be rebuilt
        (ParticleAttributes.__init__ + recalculate_cell_id, impl/particle_attributes.py:24-40,98-105)"""
        self.drain()
        m.identity_index(self.idx)
        self.length = self.valid_n_sd = self.n_sd
        self.sorted = False
        self.n_launches += 1
        if self.dependents is not None:
            self.refresh_dependents()
        if bool((self.multiplicity == 0).any().item()):  # as Builder.build does up front
            self.healthy.zero_()
            self.sanitize()

    def get_cell_start(self):
        if not self.sorted:
            self.sort_by_cell_id()
        return self.cell_start

    def is_healthy(self):
        self._words_host.copy_(self._words, non_blocking=True)
        torch.cuda.current_stream().synchronize()
        self._check_status(self._words_host)
        return bool(self._words_host[0])

    @staticmethod
    def _check_status(words):
        if words[3]:
            raise RuntimeError("fused collision step: a batch of cells exceeded the size the "
                               "last sort by cell reported")

    def sanitize(self):
        if not self.is_healthy():
            self._remove_zero_multiplicities()

    def _remove_zero_multiplicities(self):
        self.length = self.valid_n_sd
        self.length = m.remove_zero_n_or_flagged(self.multiplicity, self.idx, self.length)
        self.valid_n_sd = self.length
        self.healthy.fill_(1)
        self.gate.zero_()
        self.sorted = False

    def permutation(self, u01, local):
        if local:
            cell_start = self.get_cell_start()
            # the reference walks every cell of cell_start, also beyond a cut working length
            if self.max_cell is None:
                self.max_cell = m.max_cell_size(self.cell_start)
            m.shuffle_local(self.idx, u01, cell_start, self.sorted_length, self.max_cell)
        else:
            m.shuffle_global(self.idx, self.length, u01)
            self.sorted = False

    # ---- random numbers -----------------------------------------------------------------
    def get_random_arrays(self):
        if self.optimized_random:
            shift = self.substep
            if self.substep == 0:
                self.rnd(_View(self.pairs_rand))
                self.rnd(_View(self.rand))
        else:
            shift = 0
            self.rnd(_View(self.pairs_rand))
            self.rnd(_View(self.rand))
        self.substep += 1
        return self.pairs_rand[shift:self.n_sd + shift], self.rand

    # ---- the dynamic ---------------------------------------------------------------------
    def __call__(self):
        if not self.adaptive:
            for _ in range(self.substeps):
                self.step()
        else:
            self.dt_left.fill_(self.dt)
            if (self.substep_speculation > 0 and self.fused is not False
                    and (self.optimized_random or isinstance(self.rnd, Random))):
                self._adaptive_loop_speculative()
            while self.length != 0:
                prepass_done = self._open_substep(self.cell_idx, self.adaptive_end, None)
                self._cell_idx_identity = False
                if self._fused_ok():
                    # the launch also leaves adaptive_sdm_end's result next to `healthy`: one
                    # read-back per sub-step (valid unless SDs were removed: the reference then
                    # re-sorts by cell before adaptive_sdm_end looks at cell_start)
                    self.step_fused(prepass_done=prepass_done)
                    self._words_host.copy_(self._words, non_blocking=True)
                    torch.cuda.current_stream().synchronize()
                    self._check_status(self._words_host)
                    if self._words_host[0]:
                        end = int(self._words_host[1])
                    else:
                        self._remove_zero_multiplicities()
                        end = m.adaptive_sdm_end(self.dt_left, self.get_cell_start())
                else:
                    self.step()
                    end = m.adaptive_sdm_end(self.dt_left, self.get_cell_start())
                assert end <= self.length
                self.length = end
            self.length = self.valid_n_sd
            m.identity_index(self.cell_idx)
            self._cell_idx_identity = True
            if self.sorted and self._sort_identity and self.sorted_length == self.length:
                # idx is still in the order of a sort under the identity cell order: the stable
                # re-sort of reset_cell_idx (impl/particle_attributes.py:86-88) would change nothing
                self.nf_valid = False
            else:
                self.sort_by_cell_id()
        self.substep = 0

    def _adaptive_loop_speculative(self):
        """the adaptive sub-step loop (collision.py:178-190) without a host wait per sub-step:
        sub-step k + 1 is launched before `healthy` and adaptive_sdm_end's result of sub-step k
        have been read.  The launch takes its working length from the device word its
        predecessor wrote (pysdm_b200.h `length_dev` / `adaptive_end`) and does nothing when
        that is 0 (the loop has ended) or when a predecessor closed the gate (it left zero
        multiplicities: the host removes them, takes adaptive_sdm_end from the re-sorted
        population as the reference does, and restarts the chain).  The host bookkeeping of
        launches that did nothing is rolled back, so the results are those of the
        step-by-step loop (tests/test_gpu_step.py).  Leaves self.length == 0 when it ran the
        loop to its end; returns early (length != 0) if the fused step stops applying."""
        depth = self.substep_speculation
        ring = depth + 2
        if self._spec_host is None or self._spec_host.shape[0] != ring:
            self._spec_host = torch.zeros((ring, 9), dtype=torch.int64).pin_memory()
            self._cell_idx_ring = [torch.empty_like(self.cell_idx) for _ in range(ring)]
            self._side_stream = torch.cuda.Stream()
        # adaptive_sdm_end's result of sub-step k lives in its own word of a ring as long as the
        # ring of read-backs: the word is rewritten `ring` sub-steps later, by when the host has
        # read it -- so the read-back can leave the main stream (a side stream copies the words
        # once the launch is done) and no copy sits between two sub-steps
        assert ring <= 6
        slots = (1, 2, 4, 5, 6, 7)[:ring]
        end_word = [self._words[i:i + 1] for i in slots]
        main, side = torch.cuda.current_stream(), self._side_stream
        k = chain_start = 0
        pending = []  # (sub-step number, event behind its read-back), oldest first
        while self.length != 0:
            cell_idx = self._cell_idx_ring[k % ring]
            # (skipped on the device once the predecessor's adaptive_sdm_end came out 0)
            prepass_done = self._open_substep(
                cell_idx, end_word[k % ring],
                None if k == chain_start else end_word[(k - 1) % ring])
            self.cell_idx, self._cell_idx_identity = cell_idx, False
            if not self._fused_ok() or self.max_cell > m.fused_max_cell_size():
                break  # (the streaming step for large cells takes its length from the host)
            self.step_fused(gated=True, adaptive_end=end_word[k % ring],
                            length_dev=None if k == chain_start else end_word[(k - 1) % ring],
                            prepass_done=prepass_done)
            launched = torch.cuda.Event()
            launched.record(main)
            event = torch.cuda.Event()
            with torch.cuda.stream(side):
                side.wait_event(launched)
                self._spec_host[k % ring].copy_(self._words, non_blocking=True)
                event.record(side)
            pending.append((k, event, self._epoch))
            k += 1
            while pending and (len(pending) > depth or pending[0][1].query()):
                j, event, epoch = pending.pop(0)
                event.synchronize()
                words = self._spec_host[j % ring]
                self._check_status(words)
                # (the copy may have been taken while later sub-steps were running: the gate
                # holds the epoch of the launch that closed it, which tells whose zeros these are)
                if words[8] != 0 and words[8] <= epoch:
                    # sub-step j left zero multiplicities: the later ones were gated
                    self._rollback(k - 1 - j, self._cell_idx_ring[j % ring])
                    pending, k = [], j + 1
                    chain_start = k
                    self._remove_zero_multiplicities()
                    self.length = m.adaptive_sdm_end(self.dt_left, self.get_cell_start())
                    break
                self.length = int(words[slots[j % ring]])
                if self.length == 0:  # the loop ended with sub-step j: the later ones did nothing
                    self._rollback(k - 1 - j, self._cell_idx_ring[j % ring])
                    pending = []
                    break

    def _open_substep(self, cell_idx, adaptive_end, skip_if_zero):
        """`cell_idx.sort_by_key(dt_left)` (collision.py:181).  When the population is sorted by
        cell and the batched fused step is going to take the sub-step, the same launch also does
        what is left of that step's pre-pass (pysdm_b200_adaptive_presort): returns True then"""
        if (self.sorted and self.fused is not False and self.croupier == "local"
                and getattr(self.kernel, "fused", None) is not None and self.max_cell is not None
                and self.max_cell <= m.fused_max_cell_size() and self.n_sd < 2 ** 31):
            m.adaptive_presort(cell_idx, self.dt_left, cell_start=self.cell_start,
                               norm_factor=None if self.norm_factor_valid else self.norm_factor,
                               timestep=self.dt, dv=self.dv, adaptive_end=adaptive_end,
                               skip_if_zero=skip_if_zero)
            self.norm_factor_valid = True  # (until the next sort by cell)
            return True
        m.sort_by_key(cell_idx, self.dt_left, skip_if_zero=skip_if_zero)
        return False

    def _rollback(self, n_steps, cell_idx):
        """host bookkeeping of `n_steps` launches that did nothing"""
        if n_steps % 2:
            self.idx, self.tmp_idx = self.tmp_idx, self.idx
        if isinstance(self.rnd, Random) and not self.optimized_random:
            self.rnd.call -= 2 * n_steps
        else:
            self.substep -= n_steps
        self.n_fused_steps -= n_steps
        self.cell_idx = cell_idx

    def run(self, steps):
        if self.adaptive or self.substeps != 1:
            for _ in range(steps):
                self()
            return
        for _ in range(steps):
            self.step_async()
            self.substep = 0
        self.drain()

    def _fused_ok(self):
        if self.fused is False or self.croupier != "local":
            return False
        if getattr(self.kernel, "fused", None) is None:
            return False
        self.get_cell_start()
        if self.max_cell is None:
            self.max_cell = m.max_cell_size(self.cell_start)
        # (cells beyond m.fused_max_cell_size() -- boxes -- take the shuffle replay + streaming
        # kernel inside the same entry point, csrc/stream.cu)
        return self.n_sd < 2 ** 31

    def step(self):
        if self._fused_ok():
            self.step_fused()
        else:
            self.step_methods()
        self.sanitize()

    # ---- speculative stepping: no host wait between fused launches ---------------------------
    def step_async(self, mark=None):
        """`step()` without the GPU idling while the host reads `healthy` (the reference reads
        it after every step, impl/particle_attributes.py:67-73): fused launches are queued up to
        `speculation_depth` steps ahead of the oldest flag that has not been read.  A launch
        that leaves zero multiplicities closes the gate (pysdm_b200.h `gate`), so the launches
        queued behind it find it closed and do nothing: their host bookkeeping is undone, the
        dead super-droplets are removed and the steps are issued again -- same results as
        step(), step by step.  Needs the device Philox generator (its stream position can be
        rewound); otherwise, and whenever the fused step does not apply, this is step().
        Call drain() before reading state.  `mark`: a CUDA event recorded right behind the
        fused launch (timing)."""
        if not isinstance(self.rnd, Random) or self.optimized_random or not self._fused_ok():
            self.drain()
            self.step()
            if mark is not None:
                mark.record()
            return
        self.step_fused(gated=True)
        if mark is not None:
            mark.record()
        slot = self._slot
        self._slot = (slot + 1) % self._healthy_slots.numel()
        self._healthy_slots[slot:slot + 1].copy_(self.healthy, non_blocking=True)
        event = torch.cuda.Event()
        event.record()
        self._in_flight.append((event, slot))
        # read the flags that have arrived; wait only when the ring is full
        while self._in_flight and (len(self._in_flight) > self.speculation_depth
                                   or self._in_flight[0][0].query()):
            if not self._resolved_healthy(self._in_flight.pop(0)):
                redo = len(self._in_flight)  # the launches queued behind it did nothing
                self._undo(redo)
                for _ in range(redo):
                    self.step_async()
                return

    def _undo(self, n_steps):
        """the last `n_steps` launches found the gate closed: undo their host bookkeeping, then
        remove the dead super-droplets (what step() would have done after the unhealthy step)"""
        if n_steps % 2:
            self.idx, self.tmp_idx = self.tmp_idx, self.idx
        self.rnd.call -= 2 * n_steps
        self.n_fused_steps -= n_steps
        self._in_flight = []
        self._remove_zero_multiplicities()

    def _resolved_healthy(self, pending):
        event, slot = pending
        event.synchronize()
        return bool(self._healthy_slots[slot])

    def drain(self):
        """waits for the steps in flight and sanitizes after them, as step() would have"""
        if self._in_flight:  # (the error flag of the speculated launches, once per drain)
            self._in_flight[-1][0].synchronize()
            self._check_status(self._words.cpu())
        while self._in_flight:
            if not self._resolved_healthy(self._in_flight.pop(0)):
                redo = len(self._in_flight)
                self._undo(redo)
                for _ in range(redo):
                    self.step_async()

    def step_fused(self, gated=False, adaptive_end=None, length_dev=None, prepass_done=False):
        """one launch: csrc/fused.cu"""
        kw = {}
        if self.adaptive:
            kw["adaptive"] = {
                "prepass_done": prepass_done,
                "dt_left": self.dt_left, "stats_n_substep": self.stats_n_substep,
                "stats_dt_min": self.stats_dt_min, "dt_min": self.dt_coal_range[0],
                "dt_max": self.dt_coal_range[1], "rank_to_cell": self.rank_to_cell,
                "rank_info": self.rank_info, "length_dev": length_dev,
                "adaptive_end": self.adaptive_end if adaptive_end is None else adaptive_end}
        if self.dependents is not None:
            kw.update(radius=self.radius, fall_velocity=self.fall_velocity,
                      dependents=self.dependents)
        if isinstance(self.rnd, Random) and not self.optimized_random:
            # uniforms generated in-kernel from the same stream
            kw.update(seed=self.rnd.seed, call_u01=self.rnd.call, call_rand=self.rnd.call + 1)
            self.rnd.call += 2
        else:
            pairs_rand, rand = self.get_random_arrays()
            kw.update(u01=pairs_rand, rand=rand)
        err = m.fused_collision_step(
            idx_in=self.idx, idx_out=self.tmp_idx, cell_start=self.cell_start,
            cell_idx=self.cell_idx, cell_id_=self.cell_id, multiplicity=self.multiplicity,
            attributes=self.attributes, is_first_in_pair=self.flag, gamma=self.gamma,
            norm_factor=self.norm_factor, collision_rate=self.collision_rate,
            collision_rate_deficit=self.collision_rate_deficit,
            coalescence_rate=self.coalescence_rate, healthy=self.healthy, status=self.status,
            length=self.length, sorted_length=self.sorted_length, max_cell=self.max_cell,
            kernel=self.kernel.fused,
            kernel_a=self.kernel.a, kernel_b=self.kernel.b, timestep=self.dt, dv=self.dv,
            rho_w=self.rho_w, substeps=1 if self.adaptive else self.substeps,
            nf_slot=self.nf_slot, refresh_nf=not self.nf_valid, gate=self.gate if gated else None,
            epoch=self._epoch + 1, cache=self._launch_cache, **kw)
        m.check(err, "fused_collision_step")
        self.n_launches += 2 if self.adaptive else (1 if self.nf_valid else 3)
        self.nf_valid = not self.adaptive  # (cell_idx is re-sorted before every adaptive sub-step)
        # positions beyond the sorted range keep their content (flagged SDs)
        if self.sorted_length < self.n_sd:
            self.tmp_idx[self.sorted_length:] = self.idx[self.sorted_length:]
        self.idx, self.tmp_idx = self.tmp_idx, self.idx
        self.n_fused_steps += 1
        self._epoch += 1  # never reused: a closed gate carries the epoch of the launch that closed it

    def step_methods(self):
        """the reference's sequence of backend calls, one kernel (group) each"""
        if self.dependents is not None:  # the orchestration's recalculation of derived attributes
            self.refresh_dependents()
        pairs_rand, rand = self.get_random_arrays()
        self.permutation(pairs_rand, self.croupier == "local")
        cell_start = self.get_cell_start()
        m.find_pairs(cell_start, self.flag, self.cell_id, self.cell_idx, self.idx, self.length)
        m.sort_within_pair_by_attr(self.idx, self.length, self.flag, self.multiplicity)
        prob = self.gamma
        self.kernel(self, self.kernel_temp, self.flag)
        m.pair_op("max", prob, self.multiplicity, self.flag, self.idx, self.length)
        m.elementwise("mul", prob, prob, self.kernel_temp)
        m.normalize(prob, self.cell_id, self.cell_idx, self.get_cell_start(), self.norm_factor,
                    self.dt, self.dv)
        if self.adaptive:
            m.scale_prob_for_adaptive_sdm_gamma(prob, self.idx, self.length, self.multiplicity,
                                                self.cell_id, self.dt_left, self.dt,
                                                self.dt_coal_range, self.flag,
                                                self.stats_n_substep, self.stats_dt_min)
        else:
            m.elementwise("div", prob, prob, None, self.substeps)
        m.compute_gamma(prob, rand, self.idx, self.length, self.multiplicity, self.cell_id,
                        self.collision_rate_deficit, self.collision_rate, self.flag, self.gamma)
        m.collision_coalescence(self.multiplicity, self.idx, self.length, self.attributes,
                                self.gamma, self.healthy, self.cell_id, self.coalescence_rate,
                                self.flag)
        self.n_launches += 14

    # ---- host views for tests -------------------------------------------------------------
    def host(self, name):
        return getattr(self, name).cpu().numpy()


class _View:  # pylint: disable=too-few-public-methods
    """the minimal Storage face (`.data`, `.shape`) the Random classes fill"""

    def __init__(self, tensor):
        self.data = tensor
        self.shape = tuple(tensor.shape)
