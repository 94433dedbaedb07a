"""`B200`: the backend class behind PySDM's backend interface
(`Builder(n_sd, backend=B200(), environment=...)`), for the collision / coalescence path.

It exposes what `PySDM.backends.Numba` / `ThrustRTC` expose (PySDM/backends/numba.py:18-82,
thrust_rtc.py:32-74): `Storage`, `Random`, `formulae`, `formulae_flattened`,
`default_croupier`, and the backend methods called by the unchanged orchestration code
(list in SURVEY.md 8b; each method below cites the Numba method it replaces, paths
relative to PySDM/backends/impl_numba/methods/).  Arguments are the reference's Storage /
Index / IndexedStorage / PairIndicator / PairwiseStorage objects (built by the
`make_*` factories over `B200.Storage`); methods unwrap `.data` (CUDA tensors) and call the
C-ABI through pysdm_b200.methods.  Methods outside the collision path raise
NotImplementedError, as ThrustRTC does for chemistry/seeding/deposition.

No CPU fallback: constructing `B200` without the CUDA library or a CUDA device fails.
"""

import weakref

import numpy as np
import torch

from . import _lib
from . import methods as m
from .random import DeferredRandom, HostReplayRandom, Random
from .recorder import RECORDER, tensors_of
from .storage import Storage

try:  # drop-in: derive from the reference's own base class when PySDM is importable
    from PySDM.backends.impl_common.backend_methods import BackendMethods
except ImportError:  # stand-alone use (GPU box without PySDM)
    class BackendMethods:  # pylint: disable=too-few-public-methods
        def __init__(self):
            pass


def _unsupported(name):
    def method(self, *args, **kwargs):
        raise NotImplementedError(
            f"B200 backend implements the collision/coalescence/breakup path only: {name}")
    method.__name__ = name
    return method


class B200(BackendMethods):  # pylint: disable=too-many-public-methods
    Storage = Storage
    Random = Random
    HostReplayRandom = HostReplayRandom
    default_croupier = "local"

    def __init__(self, formulae=None, double_precision=True, device="cuda", host_rng=False,
                 defer=True):
        if not double_precision:
            raise NotImplementedError("B200 backend is fp64/int64 only")
        _lib.load()  # raises if libpysdm_b200.so has not been built
        if not torch.cuda.is_available():
            raise RuntimeError("B200 backend needs a CUDA device (there is no CPU fallback)")
        if formulae is None:
            from PySDM.formulae import Formulae  # pylint: disable=import-outside-toplevel

            formulae = Formulae()
        self.formulae = formulae
        self.formulae_flattened = formulae.flatten
        self.device = torch.device(device)
        if host_rng:  # parity runs: the Numba backend's PCG64 stream, uploaded
            self.Random = HostReplayRandom
        # cell caretakers made by this backend, by the device pointer of the cell_start they
        # last sorted into (weak references: a caretaker lives as long as its ParticleAttributes)
        self._caretakers = {}
        # device pointers of index tensors known to hold the identity permutation (written by
        # identity_index, not sorted since): lets a caretaker tell a sort under the identity
        # cell order, after which a repeated sort of the untouched idx is a no-op
        self._identity_ptrs = set()
        # defer=True: the backend calls of a `Collision.step()` are recorded and executed as one
        # fused launch when they form a step the fused kernel covers (pysdm_b200/recorder.py);
        # defer=False: every call runs its method-level kernel(s) at once
        self.defer = bool(defer)
        if self.defer and not host_rng:
            self.Random = DeferredRandom
        self._volume_of = {}  # data pointer of a "volume" tensor -> the mass tensor it came from
        self._versions = {}   # data pointer of an index tensor -> number of rewrites seen
        BackendMethods.__init__(self)

    # ---- recording (pysdm_b200/recorder.py) ---------------------------------------------------
    def _record(self, name, run, args, *holders, writes=(), start=False):
        """records the call when a step is being recorded (or `start`s one); else runs it.
        `holders`: every storage involved, `writes`: those it writes"""
        if self.defer and RECORDER.enabled and (start or RECORDER.queue):
            tensors, written = [], []
            for h in holders:
                tensors_of(h, tensors)
            for h in writes:
                tensors_of(h, written, with_idx=False)
            args["backend"] = self
            if RECORDER.defer(name, run, args, tensors, written):
                return
        if RECORDER.busy():
            tensors = []
            for h in holders:
                tensors_of(h, tensors)
            RECORDER.sync(tensors)
        run()

    def volume_source(self, volume):
        """the mass tensor `volume` was last computed from by volume_of_water_mass, or None"""
        entry = self._volume_of.get(volume.data_ptr())
        return entry() if entry is not None else None

    def version_of(self, tensor):
        return self._versions.get(tensor.data_ptr(), 0)

    @property
    def rho_w(self):
        return float(self.formulae.constants.rho_w)

    # ---- index methods (index_methods.py) ------------------------------------------------
    def identity_index(self, idx):  # :13-20
        m.identity_index(idx)
        self._identity_ptrs.add(idx.data_ptr())
        self._versions[idx.data_ptr()] = self._versions.get(idx.data_ptr(), 0) + 1

    @staticmethod
    def shuffle_global(idx, length, u01):  # :22-30
        m.shuffle_global(idx, int(length), u01)

    def sort_facts(self, cell_start):
        """the caretaker whose last sort by cell wrote the device tensor `cell_start` (it knows
        the sorted length, the largest cell and the rank -> cell map), or None"""
        ref = self._caretakers.get(cell_start.data_ptr())
        caretaker = ref() if ref is not None else None
        if caretaker is None or caretaker.cell_start_ref() is not cell_start:
            return None
        return caretaker

    def shuffle_local(self, idx, u01, cell_start):  # :32-44
        # what the last sort by cell learnt about `cell_start`: the working length and the largest
        # cell, so that no device sync is needed and populations of small cells take the
        # one-launch shuffle
        def run():
            facts = self.sort_facts(cell_start)
            if facts is None:
                m.shuffle_local(idx, u01, cell_start, None, None)
            else:
                m.shuffle_local(idx, u01, cell_start, facts.length, facts.max_cell)

        self._record("shuffle_local", run, {"idx": idx, "u01": u01, "cell_start": cell_start},
                     idx, u01, cell_start, writes=(idx,), start=True)

    def sort_by_key(self, idx, attr):  # :46-48
        m.sort_by_key(idx.data, attr.data)
        self._identity_ptrs.discard(idx.data.data_ptr())
        self._versions[idx.data.data_ptr()] = self._versions.get(idx.data.data_ptr(), 0) + 1

    def adaptive_presort(self, cell_idx, dt_left, *, cell_start, norm_factor, timestep, dv,
                         adaptive_end):
        """`cell_idx.sort_by_key(dt_left)` of the adaptive loop (collision.py:181) opening a fused
        sub-step, one launch (pysdm_b200_adaptive_presort); `norm_factor`: None when the
        per-cell factors of normalize() are valid for the current sort by cell"""
        RECORDER.flush_all()
        m.adaptive_presort(cell_idx.data, dt_left.data, cell_start=cell_start.data,
                           norm_factor=None if norm_factor is None else norm_factor.data,
                           timestep=timestep, dv=dv, adaptive_end=adaptive_end)
        self._identity_ptrs.discard(cell_idx.data.data_ptr())
        self._versions[cell_idx.data.data_ptr()] = self._versions.get(cell_idx.data.data_ptr(), 0) + 1

    @staticmethod
    def remove_zero_n_or_flagged(multiplicity, idx, length):  # collisions_methods.py:664-680
        return m.remove_zero_n_or_flagged(multiplicity, idx, int(length))

    @staticmethod
    def cell_id(cell_id, cell_origin, strides):  # collisions_methods.py:407-416
        m.cell_id(cell_id.data, cell_origin.data, strides.data)

    def make_cell_caretaker(self, idx_shape, idx_dtype, cell_start_len, scheme="default"):
        """collisions_methods.py:587-631: owns tmp_idx, swaps the .data of idx and tmp_idx"""
        registry, identity_ptrs = self._caretakers, self._identity_ptrs

        class CellCaretaker:  # pylint: disable=too-few-public-methods,too-many-instance-attributes
            def __init__(self):
                self.scheme = scheme
                self.tmp_idx = Storage.empty(idx_shape, idx_dtype)
                # facts of the last sort (for shuffle_local and the fused step)
                self.length = self.max_cell = None
                self.identity_sort = False  # the last sort ran under the identity cell order
                self.cell_start_ref = lambda: None
                self.n_sorts = 0
                self.rank_to_cell = torch.arange(cell_start_len - 1, dtype=torch.int64,
                                                 device=self.tmp_idx.data.device)

            def __call__(self, cell_id, cell_idx, cell_start, idx):
                if RECORDER.busy():
                    RECORDER.sync([cell_id.data, cell_idx.data, cell_start.data, idx.data,
                                   self.tmp_idx.data])
                m.counting_sort_by_cell_id(self.tmp_idx.data, idx.data, cell_id.data,
                                           cell_idx.data, len(idx), cell_start.data)
                idx.data, self.tmp_idx.data = self.tmp_idx.data, idx.data
                m.invert_index(self.rank_to_cell, cell_idx.data)
                self.length = int(len(idx))
                self.max_cell = m.max_cell_size(cell_start.data)
                self.identity_sort = cell_idx.data.data_ptr() in identity_ptrs
                self.cell_start_ref = weakref.ref(cell_start.data)
                self.n_sorts += 1
                for key in [k for k, v in registry.items() if v() is None]:
                    del registry[key]  # caretakers that are gone
                registry[cell_start.data.data_ptr()] = weakref.ref(self)

        assert cell_start_len >= 1
        return CellCaretaker()

    # ---- pair methods (pair_methods.py) ---------------------------------------------------
    def find_pairs(self, cell_start, is_first_in_pair, cell_id, cell_idx, idx):  # :34-55
        def run():
            m.find_pairs(cell_start.data, is_first_in_pair.indicator.data, cell_id.data,
                         cell_idx.data, idx.data, len(idx))

        self._record("find_pairs", run, {"cell_start": cell_start, "flags": is_first_in_pair,
                                         "cell_id": cell_id, "cell_idx": cell_idx, "idx": idx},
                     cell_start, is_first_in_pair, cell_id, cell_idx, idx,
                     writes=(is_first_in_pair,))

    def sort_within_pair_by_attr(self, idx, is_first_in_pair, attr):  # :126-140
        def run():
            m.sort_within_pair_by_attr(idx.data, len(idx), is_first_in_pair.indicator.data,
                                       attr.data)

        self._record("sort_within_pair_by_attr", run,
                     {"idx": idx, "flags": is_first_in_pair, "attr": attr},
                     idx, is_first_in_pair, attr, writes=(idx,))

    def _pair(self, op, data_out, data_in, is_first_in_pair, idx):
        def run():
            m.pair_op(op, data_out.data, data_in.data, is_first_in_pair.indicator.data, idx.data,
                      len(idx))

        self._record("pair_op", run, {"op": op, "out": data_out, "in": data_in,
                                      "flags": is_first_in_pair, "idx": idx},
                     data_out, data_in, is_first_in_pair, idx, writes=(data_out,))

    def distance_pair(self, data_out, data_in, is_first_in_pair, idx):  # :14-32
        self._pair("distance", data_out, data_in, is_first_in_pair, idx)

    def max_pair(self, data_out, data_in, is_first_in_pair, idx):  # :57-75
        self._pair("max", data_out, data_in, is_first_in_pair, idx)

    def min_pair(self, data_out, data_in, is_first_in_pair, idx):  # :77-95
        self._pair("min", data_out, data_in, is_first_in_pair, idx)

    def sum_pair(self, data_out, data_in, is_first_in_pair, idx):  # :142-160
        self._pair("sum", data_out, data_in, is_first_in_pair, idx)

    def multiply_pair(self, data_out, data_in, is_first_in_pair, idx):  # :162-180
        self._pair("multiply", data_out, data_in, is_first_in_pair, idx)

    @staticmethod
    def sort_pair(data_out, data_in, is_first_in_pair, idx):  # :97-124
        m.sort_pair(data_out.data, data_in.data, is_first_in_pair.indicator.data, idx.data,
                    len(idx))

    # ---- collision methods (collisions_methods.py) -----------------------------------------
    def normalize(self, prob, cell_id, cell_idx, cell_start, norm_factor, timestep, dv):  # :633-662
        def run():
            m.normalize(prob.data, cell_id.data, cell_idx.data, cell_start.data,
                        norm_factor.data, timestep, dv)

        self._record("normalize", run, {"prob": prob, "cell_id": cell_id, "cell_idx": cell_idx,
                                        "cell_start": cell_start, "norm_factor": norm_factor,
                                        "timestep": timestep, "dv": dv},
                     prob, cell_id, cell_idx, cell_start, norm_factor,
                     writes=(prob, norm_factor))

    @staticmethod
    def adaptive_sdm_end(dt_left, cell_start):  # :313-328
        return m.adaptive_sdm_end(dt_left.data, cell_start.data)

    @staticmethod
    def scale_prob_for_adaptive_sdm_gamma(*, prob, multiplicity, cell_id, dt_left, dt, dt_range,
                                          is_first_in_pair, stats_n_substep, stats_dt_min):
        # :330-405
        m.scale_prob_for_adaptive_sdm_gamma(
            prob.data, multiplicity.idx.data, len(multiplicity), multiplicity.data, cell_id.data,
            dt_left.data, dt, dt_range, is_first_in_pair.indicator.data, stats_n_substep.data,
            stats_dt_min.data)

    def compute_gamma(self, *, prob, rand, multiplicity, cell_id, collision_rate_deficit,
                      collision_rate, is_first_in_pair, out):  # :522-585
        def run():
            m.compute_gamma(prob.data, rand.data, multiplicity.idx.data, len(multiplicity),
                            multiplicity.data, cell_id.data, collision_rate_deficit.data,
                            collision_rate.data, is_first_in_pair.indicator.data, out.data)

        self._record("compute_gamma", run,
                     {"prob": prob, "rand": rand, "multiplicity": multiplicity, "cell_id": cell_id,
                      "collision_rate_deficit": collision_rate_deficit,
                      "collision_rate": collision_rate, "flags": is_first_in_pair, "out": out},
                     prob, rand, multiplicity, cell_id, collision_rate_deficit, collision_rate,
                     is_first_in_pair, out,
                     writes=(out, collision_rate_deficit, collision_rate))

    def collision_coalescence(self, *, multiplicity, idx, attributes, gamma, healthy, cell_id,
                              coalescence_rate, is_first_in_pair):  # :418-477
        def run():
            m.collision_coalescence(multiplicity.data, idx.data, len(idx), attributes.data,
                                    gamma.data, healthy.data, cell_id.data,
                                    coalescence_rate.data, is_first_in_pair.indicator.data)

        self._record("collision_coalescence", run,
                     {"multiplicity": multiplicity, "idx": idx, "attributes": attributes,
                      "gamma": gamma, "healthy": healthy, "cell_id": cell_id,
                      "coalescence_rate": coalescence_rate, "flags": is_first_in_pair},
                     multiplicity, idx, attributes, gamma, healthy, cell_id, coalescence_rate,
                     is_first_in_pair,
                     writes=(multiplicity, attributes, healthy, coalescence_rate))

    def collision_coalescence_breakup(self, *, multiplicity, idx, attributes, gamma, rand, Ec,
                                      Eb, fragment_mass, healthy, cell_id, coalescence_rate,
                                      breakup_rate, breakup_rate_deficit, is_first_in_pair,
                                      warn_overflows, particle_mass, max_multiplicity):
        # :479-520 (body :246-311; break_up :135-175 or break_up_while :178-243)
        n_overflow = m.collision_coalescence_breakup(
            multiplicity.data, idx.data, len(idx), attributes.data, gamma.data, rand.data,
            Ec.data, Eb.data, fragment_mass.data, healthy.data, cell_id.data,
            coalescence_rate.data, breakup_rate.data, breakup_rate_deficit.data,
            is_first_in_pair.indicator.data, max_multiplicity, particle_mass.data,
            handle_all_breakups=bool(self.formulae.handle_all_breakups),
            count_overflows=bool(warn_overflows))
        if warn_overflows and n_overflow:
            # impl_numba/warnings.py:12-21: the reference prints one message per pair from
            # inside the kernel; here one message carries the count
            import sys  # pylint: disable=import-outside-toplevel

            print(f"overflow ({n_overflow} pairs)\n\tfile: {__file__}"
                  "\n\tcontext:  collision_coalescence_breakup", file=sys.stderr)

    def linear_collection_efficiency(self, *, params, output, radii, is_first_in_pair, unit):
        # :743-782
        def run():
            m.linear_collection_efficiency(params, m.int_mask_of(params), output.data, radii.data,
                                           is_first_in_pair.indicator.data, radii.idx.data,
                                           len(is_first_in_pair), unit)

        self._record("lce", run, {"params": params, "output": output, "radii": radii,
                                  "flags": is_first_in_pair, "unit": unit},
                     output, radii, is_first_in_pair, writes=(output,))

    # ---- terminal velocities (terminal_velocity_methods.py) ---------------------------------
    @staticmethod
    def interpolation(*, output, radius, factor, b, c):  # :13-30
        m.interpolation(output.data, radius.data, factor, b.data, c.data)

    def terminal_velocity(self, *, values, radius):  # :32-44, values/radius are raw .data
        name = type(self.formulae.terminal_velocity).__name__
        if name != "RogersYau":
            raise NotImplementedError(f"terminal_velocity formula {name}")
        c = self.formulae.constants
        m.terminal_velocity_rogers_yau(
            values, radius,
            (c.ROGERS_YAU_TERM_VEL_SMALL_K, c.ROGERS_YAU_TERM_VEL_MEDIUM_K,
             c.ROGERS_YAU_TERM_VEL_LARGE_K, c.ROGERS_YAU_TERM_VEL_SMALL_R_LIMIT,
             c.ROGERS_YAU_TERM_VEL_MEDIUM_R_LIMIT))

    @staticmethod
    def power_series(*, values, radius, num_terms, prefactors, powers):  # :46-64
        m.power_series(values, radius, list(prefactors)[:num_terms], list(powers)[:num_terms])

    # ---- fragmentation functions (fragmentation_methods.py) ---------------------------------
    @staticmethod
    def slams_fragmentation(n_fragment, frag_volume, x_plus_y, probs, rand, vmin, nfmax):
        m.slams_fragmentation(n_fragment.data, frag_volume.data, x_plus_y.data, probs.data,
                              rand.data, vmin, nfmax)  # :106-133

    @staticmethod
    def exp_fragmentation(*, n_fragment, scale, frag_volume, x_plus_y, rand, vmin, nfmax,
                          tol=1e-5):  # :136-170
        m.exp_fragmentation(n_fragment.data, scale, frag_volume.data, x_plus_y.data, rand.data,
                            vmin, nfmax, tol)

    def feingold1988_fragmentation(self, *, n_fragment, scale, frag_volume, x_plus_y, rand,
                                   fragtol, vmin, nfmax):  # :172-198
        self._need_fragmentation_function("Feingold1988")
        m.feingold1988_fragmentation(n_fragment.data, scale, frag_volume.data, x_plus_y.data,
                                     rand.data, fragtol, vmin, nfmax)

    @staticmethod
    def gauss_fragmentation(*, n_fragment, mu, sigma, frag_volume, x_plus_y, rand, vmin, nfmax):
        m.gauss_fragmentation(n_fragment.data, mu, sigma, frag_volume.data, x_plus_y.data,
                              rand.data, vmin, nfmax)  # :200-216

    def straub_fragmentation(self, *, n_fragment, CW, gam, ds, frag_volume, v_max, x_plus_y,
                             rand, vmin, nfmax, Nr1, Nr2, Nr3, Nr4, Nrt, d34):  # :218-260
        self._need_fragmentation_function("Straub2010Nf")
        m.straub_fragmentation(n_fragment.data, CW.data, gam.data, ds.data, frag_volume.data,
                               v_max.data, x_plus_y.data, rand.data, vmin, nfmax, Nr1.data,
                               Nr2.data, Nr3.data, Nr4.data, Nrt.data, d34.data)

    def ll82_fragmentation(self, *, n_fragment, CKE, W, W2, St, ds, dl, dcoal, frag_volume,
                           x_plus_y, rand, vmin, nfmax, Rf, Rs, Rd, tol=1e-8):  # :262-303
        self._need_fragmentation_function("LowList1982Nf")
        m.ll82_fragmentation(n_fragment.data, CKE.data, W.data, W2.data, St.data, ds.data,
                             dl.data, dcoal.data, frag_volume.data, x_plus_y.data, rand.data,
                             vmin, nfmax, Rf.data, Rs.data, Rd.data, tol)

    @staticmethod
    def ll82_coalescence_check(*, Ec, dl):  # :305-319
        m.ll82_coalescence_check(Ec.data, dl.data)

    def _need_fragmentation_function(self, name):
        """the reference reaches these formulae through Formulae(fragmentation_function=name)
        and fails at JIT time otherwise; fail the same way, early"""
        frag = self.formulae.fragmentation_function  # SimpleNamespace named by _pick, formulae.py:371
        have = getattr(frag, "__name__", type(frag).__name__)
        if have != name:
            raise AttributeError(
                f"Formulae(fragmentation_function='{name}') required (got '{have}')")

    # ---- products side: moments_methods.py:12-182 -----------------------------------------------
    @staticmethod
    def moments(*, moment_0, moments, multiplicity, attr_data, cell_id, idx, length, ranks, min_x,
                max_x, x_attr, weighting_attribute, weighting_rank, skip_division_by_m0):
        m.moments(moment_0.data, moments.data, multiplicity.data, attr_data.data, cell_id.data,
                  idx.data, length, ranks.to_ndarray(), min_x, max_x, x_attr.data,
                  weighting_attribute.data, weighting_rank, skip_division_by_m0)

    @staticmethod
    def spectrum_moments(*, moment_0, moments, multiplicity, attr_data, cell_id, idx, length, rank,
                         x_bins, x_attr, weighting_attribute, weighting_rank):
        assert moments.shape[0] == x_bins.shape[0] - 1
        assert moment_0.shape == moments.shape
        m.spectrum_moments(moment_0.data, moments.data, multiplicity.data, attr_data.data,
                           cell_id.data, idx.data, length, rank, x_bins.data, x_attr.data,
                           weighting_attribute.data, weighting_rank)

    # ---- transport side: displacement_methods.py:14-260 -----------------------------------------
    def calculate_displacement(self, *, dim, displacement, courant, cell_origin, position_in_cell,
                               n_substeps):
        scheme = getattr(self.formulae.particle_advection, "__name__", "ImplicitInSpace")
        if scheme not in ("ImplicitInSpace", "ExplicitInSpace"):
            raise NotImplementedError(f"particle_advection={scheme}")
        if len(courant.shape) > 3:
            raise NotImplementedError()
        m.calculate_displacement(dim, displacement.data, courant.data, cell_origin.data,
                                 position_in_cell.data, n_substeps,
                                 implicit_in_space=scheme == "ImplicitInSpace")

    @staticmethod
    def flag_precipitated(*, cell_origin, position_in_cell, water_mass, multiplicity, idx, length,
                          healthy, precipitation_counting_level_index, displacement) -> float:
        return m.flag_precipitated(cell_origin.data, position_in_cell.data, water_mass.data,
                                   multiplicity.data, idx.data, length, healthy.data,
                                   precipitation_counting_level_index, displacement.data)

    @staticmethod
    def flag_out_of_column(cell_origin, position_in_cell, idx, length, healthy,
                           domain_top_level_index):
        m.flag_out_of_column(cell_origin.data, position_in_cell.data, idx.data, length,
                             healthy.data, domain_top_level_index)

    # ---- particle injection: seeding_methods.py:46-65 -------------------------------------------
    @staticmethod
    def seeding(*, idx, multiplicity, extensive_attributes, seeded_particle_index,
                seeded_particle_multiplicity, seeded_particle_extensive_attributes,
                number_of_super_particles_to_inject: int):
        m.seeding(idx.data, multiplicity.data, extensive_attributes.data,
                  seeded_particle_index.data, seeded_particle_multiplicity.data,
                  seeded_particle_extensive_attributes.data,
                  number_of_super_particles_to_inject)

    # ---- physics (physics_methods.py:107-133) -----------------------------------------------
    def _liquid_spheres_only(self):
        """the reference dispatches these through formulae.particle_shape_and_density
        (physics_methods.py:107-133); only LiquidSpheres (m = rho_w v) is implemented here --
        MixedPhaseSpheres (rho_i for negative, i.e. ice, masses) must not silently get it"""
        name = getattr(self.formulae.particle_shape_and_density, "__name__", "LiquidSpheres")
        if name != "LiquidSpheres":
            raise NotImplementedError(f"B200 backend: particle_shape_and_density={name}")

    def volume_of_water_mass(self, volume, mass):
        self._liquid_spheres_only()
        entry = RECORDER.current.get(volume.data.data_ptr())
        if entry is not None and entry[0] == mass.data.data_ptr():
            return  # a fused launch kept it current and nothing wrote mass or volume since
        if RECORDER.busy():
            RECORDER.sync([volume.data], [mass.data])
        m.volume_of_water_mass(volume.data, mass.data, self.rho_w)
        self._volume_of[volume.data.data_ptr()] = weakref.ref(mass.data)

    def mass_of_water_volume(self, mass, volume):
        self._liquid_spheres_only()
        m.mass_of_water_volume(mass.data, volume.data, self.rho_w)


# every method that is not recorded itself first lets the recorder replay what it has, if the
# call touches memory the record reads or writes (pysdm_b200/recorder.py)
_RECORDED = {"shuffle_local", "find_pairs", "sort_within_pair_by_attr", "_pair", "distance_pair",
             "max_pair", "min_pair", "sum_pair", "multiply_pair", "normalize", "compute_gamma",
             "collision_coalescence", "linear_collection_efficiency", "_record", "volume_source",
             "volume_of_water_mass",
             "version_of", "sort_facts", "rho_w", "_liquid_spheres_only",
             "_need_fragmentation_function"}


_READ_ONLY = {"volume_of_water_mass": ("mass",), "mass_of_water_volume": ("volume",),
              "interpolation": ("radius", "b", "c"), "terminal_velocity": ("radius",),
              "power_series": ("radius",)}


def _guarded(function):
    import functools  # pylint: disable=import-outside-toplevel

    # (only this method reads the permutation through `multiplicity.idx`)
    with_idx = function.__name__ == "scale_prob_for_adaptive_sdm_gamma"
    # the recalculations of derived attributes the orchestration triggers in the middle of a
    # step only READ these arguments (everything else counts as written)
    read_only = _READ_ONLY.get(function.__name__, ())
    names = function.__code__.co_varnames[:function.__code__.co_argcount
                                          + function.__code__.co_kwonlyargcount]

    @functools.wraps(function)
    def wrapper(*args, **kwargs):
        if RECORDER.busy():
            writes, reads = [], []
            for name, a in list(zip(names, args)) + list(kwargs.items()):
                tensors_of(a, reads if name in read_only else writes, with_idx)
            RECORDER.sync(writes, reads)
        return function(*args, **kwargs)

    return wrapper


for _name, _member in list(vars(B200).items()):
    if _name.startswith("__") or _name in _RECORDED or isinstance(_member, property):
        continue
    if isinstance(_member, staticmethod):
        setattr(B200, _name, staticmethod(_guarded(_member.__func__)))
    elif callable(_member) and not isinstance(_member, type):
        setattr(B200, _name, _guarded(_member))
del _name, _member

# everything outside the collision path: explicit NotImplementedError
for _name in (
        "condensation", "make_condensation_solver", "freeze_singular",
        "freeze_time_dependent", "freeze_time_dependent_homogeneous", "thaw_instantaneous",
        "oxidation", "dissolution", "chem_recalculate_drop_data", "chem_recalculate_cell_data",
        "equilibrate_H", "isotopic_delta", "isotopic_fractionation", "deposition",
        "critical_volume", "temperature_pressure_rh", "a_w_ice", "air_density",
        "air_dynamic_viscosity", "reynolds_number", "explicit_euler",
        "record_freezing_temperatures"):
    setattr(B200, _name, _unsupported(_name))
del _name
