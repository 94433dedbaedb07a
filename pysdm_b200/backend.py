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

import numpy as np
import torch

from . import _lib
from . import methods as m
from .random import HostReplayRandom, Random
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

    def __init__(self, formulae=None, double_precision=True, device="cuda", host_rng=False):
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
        BackendMethods.__init__(self)

    @property
    def rho_w(self):
        return float(self.formulae.constants.rho_w)

    # ---- index methods (index_methods.py) ------------------------------------------------
    @staticmethod
    def identity_index(idx):  # :13-20
        m.identity_index(idx)

    @staticmethod
    def shuffle_global(idx, length, u01):  # :22-30
        m.shuffle_global(idx, int(length), u01)

    @staticmethod
    def shuffle_local(idx, u01, cell_start):  # :32-44
        m.shuffle_local(idx, u01, cell_start)

    @staticmethod
    def sort_by_key(idx, attr):  # :46-48
        m.sort_by_key(idx.data, attr.data)

    @staticmethod
    def remove_zero_n_or_flagged(multiplicity, idx, length):  # collisions_methods.py:664-680
        return m.remove_zero_n_or_flagged(multiplicity, idx, int(length))

    @staticmethod
    def cell_id(cell_id, cell_origin, strides):  # collisions_methods.py:407-416
        m.cell_id(cell_id.data, cell_origin.data, strides.data)

    @staticmethod
    def make_cell_caretaker(idx_shape, idx_dtype, cell_start_len, scheme="default"):
        """collisions_methods.py:587-631: owns tmp_idx, swaps the .data of idx and tmp_idx"""
        class CellCaretaker:  # pylint: disable=too-few-public-methods
            def __init__(self):
                self.scheme = scheme
                self.tmp_idx = Storage.empty(idx_shape, idx_dtype)

            def __call__(self, cell_id, cell_idx, cell_start, idx):
                m.counting_sort_by_cell_id(self.tmp_idx.data, idx.data, cell_id.data,
                                           cell_idx.data, len(idx), cell_start.data)
                idx.data, self.tmp_idx.data = self.tmp_idx.data, idx.data

        assert cell_start_len >= 1
        return CellCaretaker()

    # ---- pair methods (pair_methods.py) ---------------------------------------------------
    @staticmethod
    def find_pairs(cell_start, is_first_in_pair, cell_id, cell_idx, idx):  # :34-55
        m.find_pairs(cell_start.data, is_first_in_pair.indicator.data, cell_id.data,
                     cell_idx.data, idx.data, len(idx))

    @staticmethod
    def sort_within_pair_by_attr(idx, is_first_in_pair, attr):  # :126-140
        m.sort_within_pair_by_attr(idx.data, len(idx), is_first_in_pair.indicator.data, attr.data)

    @staticmethod
    def _pair(op, data_out, data_in, is_first_in_pair, idx):
        m.pair_op(op, data_out.data, data_in.data, is_first_in_pair.indicator.data, idx.data,
                  len(idx))

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
    @staticmethod
    def normalize(prob, cell_id, cell_idx, cell_start, norm_factor, timestep, dv):  # :633-662
        m.normalize(prob.data, cell_id.data, cell_idx.data, cell_start.data, norm_factor.data,
                    timestep, dv)

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

    @staticmethod
    def compute_gamma(*, prob, rand, multiplicity, cell_id, collision_rate_deficit,
                      collision_rate, is_first_in_pair, out):  # :522-585
        m.compute_gamma(prob.data, rand.data, multiplicity.idx.data, len(multiplicity),
                        multiplicity.data, cell_id.data, collision_rate_deficit.data,
                        collision_rate.data, is_first_in_pair.indicator.data, out.data)

    @staticmethod
    def collision_coalescence(*, multiplicity, idx, attributes, gamma, healthy, cell_id,
                              coalescence_rate, is_first_in_pair):  # :418-477
        m.collision_coalescence(multiplicity.data, idx.data, len(idx), attributes.data,
                                gamma.data, healthy.data, cell_id.data, coalescence_rate.data,
                                is_first_in_pair.indicator.data)

    # ---- physics (physics_methods.py:107-133) -----------------------------------------------
    def volume_of_water_mass(self, volume, mass):
        m.volume_of_water_mass(volume.data, mass.data, self.rho_w)

    def mass_of_water_volume(self, mass, volume):
        m.mass_of_water_volume(mass.data, volume.data, self.rho_w)


# everything outside the collision path: explicit NotImplementedError
for _name in (
        "condensation", "make_condensation_solver", "calculate_displacement", "flag_precipitated",
        "flag_out_of_column", "moments", "spectrum_moments", "seeding", "freeze_singular",
        "freeze_time_dependent", "freeze_time_dependent_homogeneous", "thaw_instantaneous",
        "oxidation", "dissolution", "chem_recalculate_drop_data", "chem_recalculate_cell_data",
        "equilibrate_H", "isotopic_delta", "isotopic_fractionation", "deposition",
        "critical_volume", "temperature_pressure_rh", "a_w_ice", "air_density",
        "air_dynamic_viscosity", "reynolds_number", "explicit_euler",
        "record_freezing_temperatures"):
    setattr(B200, _name, _unsupported(_name))
del _name
