"""`FusedCoalescence` / `FusedCollision`: PySDM's `Coalescence` / `Collision` dynamics with
`step()` mapped onto the one-launch
fused kernel (csrc/fused*.cu) -- for users of the unchanged PySDM orchestration
(`Builder`, `Particulator`, attributes, products) who want the collision step at the speed
`bench.py` measures instead of ~15 backend calls per step:

    from pysdm_b200 import B200, FusedCoalescence
    builder = Builder(n_sd, backend=B200(), environment=...)
    builder.add_dynamic(FusedCoalescence(collision_kernel=Golovin(b=1.5e3)))

It derives from the reference's own class (PySDM/dynamics/collisions/collision.py:46-290,
`Coalescence` :293-330), registers under the same name ("Collision"), allocates the same
storages, and overrides `step()` only: when the configuration is one the fused kernel
covers -- B200 backend, "local" croupier, Golovin / ConstantK / Linear / Geometric /
Hydrodynamic kernel, no breakup, every cell within the fused kernel's size limit; fixed
sub-steps or the adaptive time step, with or without `optimized_random` -- one launch replaces
the sequence `Collision.step` issues (:196-234) and writes every storage that sequence writes
(idx, pair flags, gamma, multiplicities, attributes, norm factors, rate counters, `healthy`,
and for the adaptive time step `dt_left`, `stats_n_substep`, `stats_dt_min`), bit for bit
(tests/test_gpu_pysdm_backend.py).  The adaptive loop itself (`Collision.__call__`, :174-194:
sort of `cell_idx` by `dt_left`, `adaptive_sdm_end`, the working-length cut) stays the
reference's.  Cells beyond the fused kernel's shared-memory batches (boxes) take the shuffle
replay + streaming kernel inside the same entry point (csrc/stream.cu).  `FusedCollision`
(PySDM's `Collision`, i.e. with breakup) runs the step up to compute_gamma the same way and
the breakup half -- coalescence efficiency (ConstEc / Straub2010Ec / Berry1967), ConstEb,
fragment mass (AlwaysN / Straub2010Nf), collision_coalescence_breakup -- as one more launch
(csrc/breakup_fused.cu).  Anything else falls through to the reference's `step()` and the
method-level kernels.
"""

import warnings

import torch

from . import methods as m
from .random import Random
from .recorder import RECORDER


def _geometric(k):
    """collision_kernels/geometric.py:18: the factor `output *= const.PI * collection_efficiency`"""
    from PySDM.physics import constants as const  # pylint: disable=import-outside-toplevel

    return (const.PI * k.collection_efficiency, 0.0)


def _parameterized(k):  # collision_kernels/impl/parameterized.py:23: `output *= const.PI`
    from PySDM.physics import constants as const  # pylint: disable=import-outside-toplevel

    return (const.PI, 0.0)


_KERNELS = {"Golovin": lambda k: (0.0, k.b), "ConstantK": lambda k: (k.a, 0.0),
            "Linear": lambda k: (k.a, k.b), "Geometric": _geometric,
            "Hydrodynamic": _parameterized}


_CLASSES = {}


def make_fused_coalescence(which="FusedCoalescence"):
    """builds the classes lazily: importing pysdm_b200 must not need PySDM"""
    if not _CLASSES:
        _CLASSES.update(_make_classes())
    return _CLASSES[which]


def _make_classes():
    from PySDM.dynamics import Coalescence, Collision  # pylint: disable=import-outside-toplevel
    from PySDM.dynamics.collisions.coalescence_efficiencies._parameterized import (  # pylint: disable=import-outside-toplevel
        Parameterized as ParameterizedEc)
    from PySDM.dynamics.collisions.collision_kernels.impl.parameterized import Parameterized  # pylint: disable=import-outside-toplevel
    from PySDM.physics import constants as const  # pylint: disable=import-outside-toplevel

    from .backend import B200  # pylint: disable=import-outside-toplevel

    class FusedStep:  # mixed into the reference's dynamics below
        def __init__(self, *args, **kwargs):
            super().__init__(*args, **kwargs)
            self.n_fused_steps = 0
            self._nf_slot = self._status = self._nf_facts = self._rank_info = None
            self._dependents = None  # derived attributes the fused step keeps current itself
            self._last_facts = None  # caretaker of the sort the last fused (sub-)step ran on
            self._presort_nf = None  # the sort by cell norm_factor_temp was computed for (presort)
            self._launch_cache = {}

        def register(self, builder):
            super().register(builder)
            dev = self.gamma.data.device
            self._nf_slot = torch.full((self.particulator.n_sd // 2,), float("nan"),
                                       dtype=torch.float64, device=dev)
            self._rank_info = torch.zeros(self.particulator.mesh.n_cell, dtype=torch.int64,
                                          device=dev)
            # what the host reads back after a fused (sub-)step, in one copy: [0] `healthy`,
            # [1] adaptive_sdm_end's result, [2] a derived radius left the Gunn-Kinzer table,
            # [3] `status` of the fused step (a batch of cells did not fit: wrong size hint)
            self._words = torch.zeros(4, dtype=torch.int64, device=dev)
            self._words_host = torch.zeros(4, dtype=torch.int64).pin_memory()
            self._status = self._words[3:4]
            self._dependents = self._dependent_chain(builder)

        def _dependent_chain(self, builder):
            """the attributes PySDM derives from the water mass for the gravitational kernels --
            "volume" (attributes/physics/volume.py), "radius" (radius.py:15-17), the fall
            velocity (terminal_velocity.py with the Gunn-Kinzer `interpolation`,
            dynamics/terminal_velocity/gunn_and_kinzer.py:128-133) -- when they are exactly these
            stock classes: the fused step then refreshes them for the super-droplets a
            coalescence changed (pysdm_b200.h `dep_*`), instead of the orchestration
            recomputing all of them after every (sub-)step in five launches and a host sync.
            None = leave it to the orchestration."""
            name = type(self.collision_kernel).__name__
            if name != "Geometric" and not isinstance(self.collision_kernel, Parameterized):
                return None
            formulae = builder.particulator.backend.formulae
            if getattr(formulae.particle_shape_and_density, "__name__", "") != "LiquidSpheres":
                return None
            mass, volume, radius, velocity = (builder.get_attribute(k) for k in (
                "water mass", "volume", "radius", "relative fall velocity"))
            approximation = getattr(velocity, "approximation", None)
            if ([type(x).__name__ for x in (mass, volume, radius, velocity, approximation)]
                    != ["ViewWaterMass", "Volume", "Radius", "TerminalVelocity", "GunnKinzer1949"]):
                return None
            return {"chain": (mass, volume, radius, velocity), "approximation": approximation,
                    "radius_scale": 1 / formulae.constants.PI_4_3, "radius_exponent": 1 / 3,
                    "status": self._words[2:3]}

        def __call__(self):
            fused_before = self.n_fused_steps
            if self.enable and self.adaptive and self._kernel_name() is not None:
                self._adaptive_loop()  # (falls back per sub-step when a cell is too large)
            else:
                super().__call__()
            if self.adaptive and self.n_fused_steps > fused_before:
                # the reference looks at this after every sub-step (collision.py:276-277); the
                # fused sub-steps leave it to one look per time step (stats_dt_min only falls)
                if self.stats_dt_min.amin() == self.dt_coal_range[0]:
                    warnings.warn("adaptive time-step reached dt_min")

        def _adaptive_loop(self):
            """`Collision.__call__` for the adaptive time step (collision.py:178-194), call for
            call, with the fused sub-step: the launch also computes adaptive_sdm_end's result
            (pysdm_b200.h `adaptive_end`), so one read-back per sub-step brings `healthy` and the
            next working length instead of two synchronising calls"""
            part, attrs = self.particulator, self.particulator.attributes
            self.dt_left[:] = part.dt
            in_order = False
            while attrs.get_working_length() != 0:
                prepass_done = self._presort()
                if not prepass_done:
                    attrs.cell_idx.sort_by_key(self.dt_left)
                end = self.step(want_end=True, prepass_done=prepass_done)
                in_order = end is not None
                if end is None:  # not a fused sub-step, or super-droplets were removed
                    end = part.adaptive_sdm_end(self.dt_left)
                attrs.cut_working_length(end)
            attrs.reset_working_length()
            facts = self._last_facts
            if (in_order and facts is not None and facts.identity_sort
                    and facts.length == attrs.get_working_length()):
                # idx is still in the order of a sort under the identity cell order (the local
                # shuffles keep every cell's range, nothing was removed or re-sorted since): the
                # stable re-sort of reset_cell_idx (impl/particle_attributes.py:86-88) would
                # reproduce idx and cell_start as they are
                attrs.cell_idx.reset_index()
            else:
                attrs.reset_cell_idx()
            self.rnd_opt_coll.reset()
            if self.enable_breakup:
                self.rnd_opt_proc.reset()
                self.rnd_opt_frag.reset()

        def _presort(self):
            """opens a sub-step with `cell_idx.sort_by_key(dt_left)` and the fused sub-step's
            per-rank pre-pass in one launch, when the population is sorted by cell (nothing was
            removed since) and the batched fused kernel is going to take the sub-step"""
            part, attrs = self.particulator, self.particulator.attributes
            if (not attrs._ParticleAttributes__sorted  # pylint: disable=protected-access
                    or self._kernel_name() is None):
                return False
            keys = attrs._ParticleAttributes__extensive_keys  # pylint: disable=protected-access
            facts = part.backend.sort_facts(attrs.cell_start.data)
            if (facts is None or facts.max_cell is None or part.n_sd >= 2 ** 31
                    or facts.max_cell > m.fused_max_cell_size()
                    or ("signed water mass" not in keys and "water mass" not in keys)):
                return False
            sort_id = (facts, facts.n_sorts)
            part.backend.adaptive_presort(
                attrs.cell_idx, self.dt_left, cell_start=attrs.cell_start,
                norm_factor=None if self._presort_nf == sort_id else self.norm_factor_temp,
                timestep=part.dt, dv=part.mesh.dv, adaptive_end=self._words[1:2])
            self._presort_nf = sort_id
            return True

        def _kernel_name(self):
            """name of the fused kernel's instantiation for this configuration, or None when the
            fused step does not cover it (decided without touching the particle state)"""
            part = self.particulator
            name = type(self.collision_kernel).__name__
            if isinstance(self.collision_kernel, Parameterized):
                name = "Hydrodynamic"  # (also `Electric`: Berry's form with other parameters)
            if self.enable_breakup and self._breakup_plan() is None:
                return None
            if (not isinstance(part.backend, B200)
                    or self.croupier != "local" or name not in _KERNELS
                    or part.n_sd >= 2 ** 32
                    or getattr(part.backend.formulae.particle_shape_and_density, "__name__",
                               "LiquidSpheres") != "LiquidSpheres"):
                return None
            return name

        def _breakup_plan(self):
            """keyword arguments of methods.fused_breakup for this dynamic's efficiencies and
            fragmentation function, or None when the fused breakup kernel does not cover them"""
            ec, eb, nf = (self.compute_coalescence_efficiency, self.compute_breakup_efficiency,
                          self.compute_number_of_fragments)
            plan = {}
            if type(eb).__name__ != "ConstEb":
                return None
            plan["eb_const"] = eb.Eb
            if type(ec).__name__ == "ConstEc":
                plan.update(ec_kind=0, ec_const=ec.Ec)
            elif type(ec).__name__ == "Straub2010Ec":
                plan.update(ec_kind=1)
            elif isinstance(ec, ParameterizedEc):  # Berry1967
                plan.update(ec_kind=2, ec_params=ec.params, ec_unit=const.si.um)
            else:
                return None
            if type(nf).__name__ == "AlwaysN":
                plan.update(nf_kind=0, nf_n=nf.N)
            elif type(nf).__name__ == "Straub2010Nf":
                plan.update(nf_kind=1, vmin=nf.vmin, nfmax=nf.nfmax)
            else:
                return None
            return plan

        def _fused_plan(self):
            """(kernel name, a, b, caretaker of the current sort) or None"""
            part = self.particulator
            name = self._kernel_name()
            if name is None:
                return None
            cell_start = part.attributes.cell_start  # sorts by cell when needed
            facts = part.backend.sort_facts(cell_start.data)
            if facts is None or part.n_sd >= 2 ** 31:
                return None
            return (name, *_KERNELS[name](self.collision_kernel), facts)

        def _fused_breakup(self, idx, healthy, ext, length, mass_row, dependents):
            """collision.py:206-234 after compute_gamma, in one launch"""
            part, attrs = self.particulator, self.particulator.attributes
            plan = self._breakup_plan()
            kw = {}
            proc, frag = self.rnd_opt_proc, self.rnd_opt_frag
            if (isinstance(proc.rnd, Random) and not proc.optimized_random
                    and proc.rnd.seed == frag.rnd.seed):
                # device Philox: the two uniforms are generated inside the kernel
                kw.update(seed=proc.rnd.seed, call_proc=proc.rnd.call, call_frag=frag.rnd.call)
                for opt in (proc, frag):
                    opt.rnd.call += 1
                    opt.substep += 1
            else:
                kw.update(rand_proc=proc.get_random_arrays().data,
                          rand_frag=frag.get_random_arrays().data)
            if plan["ec_kind"] == 1 or plan["nf_kind"] == 1:
                kw.update(volume=attrs["volume"].data,
                          fall_velocity=attrs["relative fall velocity"].data)
            if plan["ec_kind"] == 2 or plan["nf_kind"] == 1:
                kw["radius"] = attrs["radius"].data
            constants = part.backend.formulae.constants
            RECORDER.flush_all()
            n_overflow = m.fused_breakup(
                multiplicity=attrs["multiplicity"].data, idx=idx.data, attributes=ext.data,
                gamma=self.gamma.data, is_first_in_pair=self.is_first_in_pair.indicator.data,
                cell_id_=attrs["cell id"].data, healthy=healthy.data,
                coalescence_rate=self.coalescence_rate.data, breakup_rate=self.breakup_rate.data,
                breakup_rate_deficit=self.breakup_rate_deficit.data, length=length,
                mass_row=mass_row, max_multiplicity=self.max_multiplicity,
                handle_all_breakups=bool(part.backend.formulae.handle_all_breakups),
                constants={"rho_w": constants.rho_w, "PI": constants.PI,
                           "sgm_w": constants.sgm_w, "uJ": const.si.uJ},
                dependents=dependents, count_overflows=bool(self.warn_overflows), **plan, **kw)
            if self.warn_overflows and n_overflow:
                import sys  # pylint: disable=import-outside-toplevel

                print(f"overflow ({n_overflow} pairs)\n\tfile: {__file__}"
                      "\n\tcontext:  collision_coalescence_breakup", file=sys.stderr)

        def step(self, want_end=False, prepass_done=False):  # pylint: disable=arguments-differ
            """one (sub-)step.  `want_end` (adaptive loop): returns adaptive_sdm_end's result
            when the fused launch computed it and it is still valid, else None"""
            plan = self._fused_plan()
            if plan is None:
                super().step()
                return None
            name, kernel_a, kernel_b, facts = plan
            part, attrs = self.particulator, self.particulator.attributes
            # the reference reaches these through the same name-mangled handles
            # (PySDM/particulator.py:174-176)
            idx = attrs._ParticleAttributes__idx  # pylint: disable=protected-access
            healthy = attrs._ParticleAttributes__healthy_memory  # pylint: disable=protected-access
            keys = attrs._ParticleAttributes__extensive_keys  # pylint: disable=protected-access
            mass_row = keys.get("signed water mass", keys.get("water mass"))
            if mass_row is None:
                super().step()
                return None
            rnd = self.rnd_opt_coll.rnd
            if isinstance(rnd, Random) and not self.rnd_opt_coll.optimized_random:
                # device Philox: generated inside the kernel
                kw = {"seed": rnd.seed, "call_u01": rnd.call, "call_rand": rnd.call + 1}
                rnd.call += 2
                self.rnd_opt_coll.substep += 1
            else:  # injected uniforms (parity runs with the Numba backend's stream), or the
                # arrays `optimized_random` fills once per time step and reads through shifted views
                pairs_rand, rand = self.rnd_opt_coll.get_random_arrays()
                kw = {"u01": pairs_rand.data, "rand": rand.data}
            if name == "Hydrodynamic":  # Berry's parameterization, parameterized.py:14-20
                from PySDM.physics import constants as const  # pylint: disable=import-outside-toplevel

                kw["berry_params"] = self.collision_kernel.params
                kw["berry_unit"] = const.si.um
            if name in ("Geometric", "Hydrodynamic"):
                # dependent attributes, recomputed by the orchestration when the water mass was
                # marked updated (attributes/physics/radius.py:15-17, relative_fall_velocity.py,
                # terminal_velocity.py) -- the operands geometric.py:15-22 gathers
                kw["radius"] = attrs["radius"].data
                kw["fall_velocity"] = attrs["relative fall velocity"].data
                if self._dependents is not None:
                    dep, approx = self._dependents, self._dependents["approximation"]
                    kw["dependents"] = {
                        "volume": attrs["volume"].data, "radius": kw["radius"],
                        "velocity": kw["fall_velocity"], "table_b": approx.a.data,
                        "table_c": approx.b.data, "table_factor": approx.factor,
                        "radius_max": approx.maximum_radius, "status": dep["status"],
                        "radius_scale": dep["radius_scale"],
                        "radius_exponent": dep["radius_exponent"]}
            substeps = 1
            if self.adaptive:
                kw["adaptive"] = {
                    "dt_left": self.dt_left.data, "stats_n_substep": self.stats_n_substep.data,
                    "stats_dt_min": self.stats_dt_min.data, "dt_min": self.dt_coal_range[0],
                    "dt_max": self.dt_coal_range[1], "rank_to_cell": facts.rank_to_cell,
                    "rank_info": self._rank_info, "prepass_done": prepass_done,
                    "adaptive_end": self._words[1:3] if want_end else None}
            else:
                substeps = self._Collision__substeps  # pylint: disable=no-member
            length = len(idx)
            ext = attrs.get_extensive_attribute_storage()
            sort_id = (facts, facts.n_sorts)
            RECORDER.flush_all()  # (this class drives the kernels directly: nothing may be pending)
            # the argument block is kept between launches; this identifies everything in it that
            # is not rewritten per launch (methods.fused_collision_step, `cache`)
            static = [attrs.cell_start.data, attrs["cell id"].data, attrs["multiplicity"].data,
                      ext.data, self.is_first_in_pair.indicator.data, self.gamma.data,
                      self.norm_factor_temp.data, self.collision_rate.data,
                      self.collision_rate_deficit.data, self.coalescence_rate.data, healthy.data]
            for group in (kw, kw.get("adaptive", {}), kw.get("dependents", {})):
                static.extend(v for k, v in group.items()
                              if isinstance(v, torch.Tensor) and k not in ("u01", "rand",
                                                                            "adaptive_end"))
            cache_key = (name, kernel_a, kernel_b, part.dt, part.mesh.dv, substeps, mass_row,
                         bool(self.enable_breakup), *(t.data_ptr() for t in static))
            err = m.fused_collision_step(
                cache=self._launch_cache, cache_key=cache_key,
                idx_in=idx.data, idx_out=facts.tmp_idx.data, cell_start=attrs.cell_start.data,
                cell_idx=attrs.cell_idx.data, cell_id_=attrs["cell id"].data,
                multiplicity=attrs["multiplicity"].data, attributes=ext.data,
                is_first_in_pair=self.is_first_in_pair.indicator.data, gamma=self.gamma.data,
                norm_factor=self.norm_factor_temp.data, collision_rate=self.collision_rate.data,
                collision_rate_deficit=self.collision_rate_deficit.data,
                coalescence_rate=self.coalescence_rate.data, healthy=healthy.data,
                status=self._status, length=length, sorted_length=facts.length,
                max_cell=facts.max_cell, kernel=name,
                kernel_a=kernel_a, kernel_b=kernel_b, timestep=part.dt, dv=part.mesh.dv,
                rho_w=part.backend.rho_w, substeps=substeps, mass_row=mass_row,
                nf_slot=self._nf_slot, refresh_nf=self._nf_facts != sort_id,
                skip_coalescence=self.enable_breakup, **kw)
            m.check(err, "fused_collision_step")
            self._nf_facts = sort_id
            self._last_facts = facts
            if facts.length < idx.data.numel():  # flagged positions beyond the sorted range stay
                facts.tmp_idx.data[facts.length:] = idx.data[facts.length:]
            idx.data, facts.tmp_idx.data = facts.tmp_idx.data, idx.data
            self.is_first_in_pair.length = length
            self.n_fused_steps += 1
            if self.enable_breakup:
                self._fused_breakup(idx, healthy, ext, length, mass_row, kw.get("dependents"))
            # Particulator.collision_coalescence_breakup, PySDM/particulator.py:210-213
            # one read-back: `healthy` (what sanitize() would read), the next working length, and
            # the two error flags of the launch
            end = None
            self._words[0:1].copy_(healthy.data)
            self._words_host.copy_(self._words, non_blocking=True)
            torch.cuda.current_stream().synchronize()
            if self._words_host[3]:
                raise RuntimeError("fused collision step: a batch of cells exceeded the size the "
                                   "last sort by cell reported")
            if self._words_host[2]:  # gunn_and_kinzer.py:129-133
                raise ValueError("Radii can be interpolated up to "
                                 f"{self._dependents['approximation'].maximum_radius} m")
            if not self._words_host[0]:
                attrs.sanitize()  # removes the dead super-droplets; idx is re-sorted next
            elif want_end:
                end = int(self._words_host[1])
            attrs.mark_updated("multiplicity")
            for key in attrs.get_extensive_attribute_keys():
                attrs.mark_updated(key)
            if "dependents" in kw:
                # volume / radius / fall velocity are current already: stamp them as derived from
                # the water mass as it is now (DerivedAttribute.update, attributes/impl/
                # derived_attribute.py:18-26, compares these time stamps), in dependency order
                for attribute in self._dependents["chain"]:
                    for dependency in attribute.dependencies:
                        dependency.update()
                    attribute.timestamp = sum(d.timestamp for d in attribute.dependencies)
            return end

    class FusedCoalescence(FusedStep, Coalescence):
        pass

    class FusedCollision(FusedStep, Collision):
        pass

    return {"FusedCoalescence": FusedCoalescence, "FusedCollision": FusedCollision}
