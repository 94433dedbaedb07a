"""ctypes/numpy front end of the CPU oracle (oracle/sdm_oracle.c).

TEST INFRASTRUCTURE, NOT PRODUCT CODE: may be imported only by tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs.
Function names and argument order follow the reference's Numba backend methods
(/root/reference/PySDM/backends/impl_numba/methods/*.py; each C function cites
its file:line), operating in place on numpy arrays.
"""

import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libsdm_oracle.so")

i64 = ctypes.c_int64
f64 = ctypes.c_double
ptr = ctypes.c_void_p


def build(force=False):
    """compile the C restatement with the system gcc (seconds)"""
    srcs = [os.path.join(_HERE, f) for f in ("sdm_oracle.c", "sdm_physics.c", "sdm_transport.c",
                                             "Makefile")]
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < max(map(os.path.getmtime, srcs)):
        subprocess.check_call(["make", "-C", _HERE, "-s", "-B"])
    return _SO


_lib = None


def lib():
    global _lib  # pylint: disable=global-statement
    if _lib is None:
        build()
        _lib = ctypes.CDLL(_SO)
        _lib.oracle_remove_zero_n_or_flagged.restype = i64
        _lib.oracle_adaptive_sdm_end.restype = i64
        _lib.oracle_collision_coalescence_breakup.restype = i64
        _lib.oracle_flag_precipitated.restype = f64
    return _lib


def set_num_threads(n):
    lib().oracle_set_num_threads(ctypes.c_int(int(n)))


def get_max_threads():
    return int(lib().oracle_get_max_threads())


def _p(a, dtype=None):
    """pointer to a C-contiguous numpy array (checked, never copied)"""
    assert isinstance(a, np.ndarray), type(a)
    assert a.flags["C_CONTIGUOUS"]
    if dtype is not None:
        assert a.dtype == dtype, (a.dtype, dtype)
    return ptr(a.ctypes.data)


def _pi(a):
    return _p(a, np.int64)


def _pf(a):
    return _p(a, np.float64)


def _pb(a):
    assert a.dtype in (np.bool_, np.uint8)
    return _p(a)


# ---- index methods -----------------------------------------------------------
def pair_indices(idx, is_first_in_pair, prob_like):
    n = len(prob_like)
    j = np.empty(n, np.int64)
    k = np.empty(n, np.int64)
    skip = np.empty(n, np.uint8)
    lib().oracle_pair_indices(i64(n), _pi(idx), _pb(is_first_in_pair), _pf(prob_like),
                              _pi(j), _pi(k), _pb(skip))
    return j, k, skip.astype(bool)


def identity_index(idx):
    lib().oracle_identity_index(_pi(idx), i64(len(idx)))


def shuffle_global(idx, length, u01):
    lib().oracle_shuffle_global(_pi(idx), i64(length), _pf(u01))


def shuffle_local(idx, u01, cell_start):
    lib().oracle_shuffle_local(_pi(idx), _pf(u01), _pi(cell_start), i64(len(cell_start) - 1))


def sort_by_key(idx, attr):
    lib().oracle_sort_by_key(_pi(idx), _pf(attr), i64(len(attr)))


def remove_zero_n_or_flagged(multiplicity, idx, length):
    return int(lib().oracle_remove_zero_n_or_flagged(_pi(multiplicity), _pi(idx), i64(length),
                                                     i64(len(idx))))


def counting_sort_by_cell_id(new_idx, idx, cell_id, cell_idx, length, cell_start):
    lib().oracle_counting_sort_by_cell_id(_pi(new_idx), _pi(idx), _pi(cell_id), _pi(cell_idx),
                                          i64(length), _pi(cell_start), i64(len(cell_start)))


def cell_id(cell_id_out, cell_origin, strides):
    n_dims, n_sd = cell_origin.shape
    strides = np.ascontiguousarray(np.asarray(strides, dtype=np.int64).ravel())
    lib().oracle_cell_id(_pi(cell_id_out), _pi(cell_origin), _pi(strides), i64(n_dims), i64(n_sd))


# ---- pair methods --------------------------------------------------------------
def find_pairs(cell_start, is_first_in_pair, cell_id_, cell_idx, idx, length):
    lib().oracle_find_pairs(_pi(cell_start), _pb(is_first_in_pair), _pi(cell_id_), _pi(cell_idx),
                            _pi(idx), i64(length))


def sort_within_pair_by_attr(idx, length, is_first_in_pair, attr):
    if attr.dtype == np.int64:
        lib().oracle_sort_within_pair_by_attr_i64(_pi(idx), i64(length), _pb(is_first_in_pair),
                                                  _pi(attr))
    else:
        lib().oracle_sort_within_pair_by_attr_f64(_pi(idx), i64(length), _pb(is_first_in_pair),
                                                  _pf(attr))


PAIR_OPS = {"distance": 0, "max": 1, "min": 2, "sum": 3, "multiply": 4}


def pair_op(op, data_out, data_in, is_first_in_pair, idx, length):
    fun = lib().oracle_pair_op_i64 if data_in.dtype == np.int64 else lib().oracle_pair_op_f64
    fun(ctypes.c_int(PAIR_OPS[op]), _pf(data_out), i64(len(data_out)), _p(data_in),
        _pb(is_first_in_pair), _pi(idx), i64(length))


def sort_pair(data_out, data_in, is_first_in_pair, idx, length):
    lib().oracle_sort_pair_f64(_pf(data_out), i64(len(data_out)), _pf(data_in),
                               _pb(is_first_in_pair), _pi(idx), i64(length))


# ---- physics ---------------------------------------------------------------------
def volume_of_water_mass(volume, mass, rho_w):
    lib().oracle_volume_of_water_mass(_pf(volume), _pf(mass), i64(len(volume)), f64(rho_w))


def mass_of_water_volume(mass, volume, rho_w):
    lib().oracle_mass_of_water_volume(_pf(mass), _pf(volume), i64(len(volume)), f64(rho_w))


# ---- collisions --------------------------------------------------------------------
def normalize(prob, cell_id_, cell_idx, cell_start, norm_factor, timestep, dv):
    lib().oracle_normalize(_pf(prob), i64(len(prob)), _pi(cell_id_), _pi(cell_idx),
                           _pi(cell_start), i64(len(cell_start) - 1), _pf(norm_factor),
                           f64(timestep), f64(dv))


def compute_gamma(prob, rand, idx, length, multiplicity, cell_id_, collision_rate_deficit,
                  collision_rate, is_first_in_pair, out):
    lib().oracle_compute_gamma(_pf(prob), _pf(rand), _pi(idx), i64(length), _pi(multiplicity),
                               _pi(cell_id_), _pi(collision_rate_deficit), _pi(collision_rate),
                               _pb(is_first_in_pair), _pf(out))


def collision_coalescence(multiplicity, idx, length, attributes, gamma, healthy, cell_id_,
                          coalescence_rate, is_first_in_pair):
    n_attr, n_sd = attributes.shape
    lib().oracle_collision_coalescence(_pi(multiplicity), _pi(idx), i64(length), _pf(attributes),
                                       i64(n_attr), i64(n_sd), _pf(gamma), _pi(healthy),
                                       _pi(cell_id_), _pi(coalescence_rate),
                                       _pb(is_first_in_pair))


def adaptive_sdm_end(dt_left, cell_start):
    return int(lib().oracle_adaptive_sdm_end(_pf(dt_left), i64(len(dt_left)), _pi(cell_start)))


def scale_prob_for_adaptive_sdm_gamma(prob, idx, length, multiplicity, cell_id_, dt_left, dt,
                                      dt_range, is_first_in_pair, stats_n_substep, stats_dt_min):
    lib().oracle_scale_prob_for_adaptive_sdm_gamma(
        _pf(prob), _pi(idx), i64(length), _pi(multiplicity), _pi(cell_id_), _pf(dt_left),
        i64(len(dt_left)), f64(dt), f64(dt_range[0]), f64(dt_range[1]), _pb(is_first_in_pair),
        _pi(stats_n_substep), _pf(stats_dt_min))


def collision_coalescence_breakup(multiplicity, idx, length, attributes, gamma, rand, Ec, Eb,
                                  fragment_mass, healthy, cell_id_, coalescence_rate,
                                  breakup_rate, breakup_rate_deficit, is_first_in_pair,
                                  max_multiplicity, particle_mass, handle_all_breakups=False):
    """returns the number of pairs that raised the overflow flag"""
    n_attr, n_sd = attributes.shape
    return int(lib().oracle_collision_coalescence_breakup(
        _pi(multiplicity), _pi(idx), i64(length), _pf(attributes), i64(n_attr), i64(n_sd),
        _pf(gamma), _pf(rand), _pf(Ec), _pf(Eb), _pf(fragment_mass), _pi(healthy), _pi(cell_id_),
        _pi(coalescence_rate), _pi(breakup_rate), _pi(breakup_rate_deficit),
        _pb(is_first_in_pair), f64(max_multiplicity), _pf(particle_mass),
        ctypes.c_int(int(handle_all_breakups))))


# ---- physics rows (oracle/sdm_physics.c) ---------------------------------------------
def power(x, exponent):
    """Storage.__ipow__: sign(x) * |x| ** exponent, in place"""
    lib().oracle_power(_pf(x), i64(len(x)), f64(exponent))


def product_scalar(out, in_, scalar):
    lib().oracle_product_scalar(_pf(out), _pf(in_), f64(scalar), i64(len(out)))


def interpolation(output, radius, factor, b, c):
    lib().oracle_interpolation(_pf(output), _pf(radius), i64(len(radius)), f64(factor), _pf(b),
                               _pf(c))


def terminal_velocity_rogers_yau(values, radius, params):
    params = np.ascontiguousarray(params, dtype=np.float64)
    lib().oracle_terminal_velocity_rogers_yau(_pf(values), _pf(radius), i64(len(values)),
                                              _pf(params))


def power_series(values, radius, prefactors, powers):
    prefactors = np.ascontiguousarray(prefactors, dtype=np.float64)
    powers = np.ascontiguousarray(powers, dtype=np.float64)
    lib().oracle_power_series(_pf(values), _pf(radius), i64(len(values)), i64(len(powers)),
                              _pf(prefactors), _pf(powers))


def int_mask_of(params):
    """bit e set when params[e] is a Python/numpy integer (see sdm_physics.c)"""
    return sum(1 << e for e, v in enumerate(params) if isinstance(v, (int, np.integer)))


def linear_collection_efficiency(params, int_mask, output, radii, is_first_in_pair, idx, length,
                                 unit):
    params = np.ascontiguousarray(params, dtype=np.float64)
    lib().oracle_linear_collection_efficiency(
        _pf(params), i64(int_mask), _pf(output), i64(len(output)), _pf(radii),
        _pb(is_first_in_pair), _pi(idx), i64(length), f64(unit))


def fragmentation_limiters(n_fragment, frag_volume, vmin, nfmax, x_plus_y):
    lib().oracle_fragmentation_limiters(
        _pf(n_fragment), _pf(frag_volume), i64(len(frag_volume)), f64(vmin),
        ctypes.c_int(nfmax is not None), f64(0.0 if nfmax is None else nfmax), _pf(x_plus_y))


def slams_fragmentation(n_fragment, frag_volume, x_plus_y, probs, rand, vmin, nfmax):
    lib().oracle_slams_fragmentation(_pf(n_fragment), _pf(frag_volume), _pf(x_plus_y), _pf(probs),
                                     _pf(rand), i64(len(rand)))
    fragmentation_limiters(n_fragment, frag_volume, vmin, nfmax, x_plus_y)


def exp_fragmentation(n_fragment, scale, frag_volume, x_plus_y, rand, vmin, nfmax, tol=1e-5):
    lib().oracle_exp_fragmentation(f64(scale), _pf(frag_volume), _pf(rand), i64(len(rand)),
                                   f64(tol))
    fragmentation_limiters(n_fragment, frag_volume, vmin, nfmax, x_plus_y)


def feingold1988_fragmentation(n_fragment, scale, frag_volume, x_plus_y, rand, fragtol, vmin,
                               nfmax):
    lib().oracle_feingold1988_fragmentation(f64(scale), _pf(frag_volume), _pf(x_plus_y),
                                            _pf(rand), i64(len(rand)), f64(fragtol))
    fragmentation_limiters(n_fragment, frag_volume, vmin, nfmax, x_plus_y)


def gauss_fragmentation(n_fragment, mu, sigma, frag_volume, x_plus_y, rand, vmin, nfmax):
    lib().oracle_gauss_fragmentation(f64(mu), f64(sigma), _pf(frag_volume), _pf(rand),
                                     i64(len(rand)))
    fragmentation_limiters(n_fragment, frag_volume, vmin, nfmax, x_plus_y)


def straub_fragmentation(n_fragment, CW, gam, ds, frag_volume, v_max, x_plus_y, rand, vmin,
                         nfmax, Nr1, Nr2, Nr3, Nr4, Nrt, d34):
    lib().oracle_straub_fragmentation(_pf(CW), _pf(gam), _pf(ds), _pf(v_max), _pf(frag_volume),
                                      _pf(rand), _pf(Nr1), _pf(Nr2), _pf(Nr3), _pf(Nr4),
                                      _pf(Nrt), _pf(d34), i64(len(rand)))
    fragmentation_limiters(n_fragment, frag_volume, vmin, nfmax, x_plus_y)


def ll82_fragmentation(n_fragment, CKE, W, W2, St, ds, dl, dcoal, frag_volume, x_plus_y, rand,
                       vmin, nfmax, Rf, Rs, Rd, tol=1e-8):
    lib().oracle_ll82_fragmentation(_pf(CKE), _pf(W), _pf(W2), _pf(St), _pf(ds), _pf(dl),
                                    _pf(dcoal), _pf(frag_volume), _pf(rand), _pf(Rf), _pf(Rs),
                                    _pf(Rd), i64(len(rand)), f64(tol))
    fragmentation_limiters(n_fragment, frag_volume, vmin, nfmax, x_plus_y)


def ll82_coalescence_check(Ec, dl):
    lib().oracle_ll82_coalescence_check(_pf(Ec), _pf(dl), i64(len(Ec)))


# ---- products and transport side (oracle/sdm_transport.c) ---------------------------------------
def _is_int(x):
    return int(isinstance(x, (int, np.integer)) and not isinstance(x, bool))


def moments(moment_0, moments_, multiplicity, attr_data, cell_id_, idx, length, ranks, min_x,
            max_x, x_attr, weighting_attribute, weighting_rank, skip_division_by_m0):
    ranks = np.ascontiguousarray(ranks, dtype=np.float64)
    lib().oracle_moments(
        _p(moment_0), _p(moments_), _p(multiplicity), _p(attr_data) if len(ranks) else None,
        _p(cell_id_), _p(idx), i64(length), _p(ranks), i64(len(ranks)), f64(min_x), f64(max_x),
        _p(x_attr), _p(weighting_attribute), f64(weighting_rank),
        ctypes.c_int(_is_int(weighting_rank)), ctypes.c_int(int(bool(skip_division_by_m0))),
        i64(len(moment_0)))


def spectrum_moments(moment_0, moments_, multiplicity, attr_data, cell_id_, idx, length, rank,
                     x_bins, x_attr, weighting_attribute, weighting_rank):
    n_bins, n_cell = moments_.shape
    lib().oracle_spectrum_moments(
        _p(moment_0), _p(moments_), _p(multiplicity), _p(attr_data), _p(cell_id_), _p(idx),
        i64(length), f64(rank), ctypes.c_int(_is_int(rank)), _p(x_bins), i64(n_bins), _p(x_attr),
        _p(weighting_attribute), f64(weighting_rank), ctypes.c_int(_is_int(weighting_rank)),
        i64(n_cell))


def calculate_displacement(dim, displacement, courant, cell_origin, position_in_cell, n_substeps,
                           implicit_in_space=True):
    n_dims, n_sd = displacement.shape
    shape = np.asarray(courant.shape, dtype=np.int64)
    lib().oracle_calculate_displacement(
        ctypes.c_int(dim), ctypes.c_int(n_dims), ctypes.c_int(int(implicit_in_space)),
        _p(displacement), _p(courant), _p(shape), _p(cell_origin), _p(position_in_cell), i64(n_sd),
        f64(n_substeps))


def flag_precipitated(cell_origin, position_in_cell, water_mass, multiplicity, idx, length,
                      healthy, precipitation_counting_level_index, displacement):
    return float(lib().oracle_flag_precipitated(
        _p(np.ascontiguousarray(cell_origin[-1])), _p(np.ascontiguousarray(position_in_cell[-1])),
        _p(water_mass), _p(multiplicity), _p(idx), i64(length), i64(len(idx)), _p(healthy),
        f64(precipitation_counting_level_index), _p(np.ascontiguousarray(displacement[-1]))))


def flag_out_of_column(cell_origin, position_in_cell, idx, length, healthy,
                       domain_top_level_index):
    lib().oracle_flag_out_of_column(
        _p(np.ascontiguousarray(cell_origin[-1])), _p(np.ascontiguousarray(position_in_cell[-1])),
        _p(idx), i64(length), i64(len(idx)), _p(healthy), f64(domain_top_level_index))


def update_position(cell_origin, position_in_cell, displacement, floor_of_position, grid,
                    apply_boundary=True):
    """Displacement.update_position + update_cell_origin (+ boundary_condition)"""
    n_dims, n_sd = position_in_cell.shape
    lib().oracle_update_position(_p(cell_origin), _p(position_in_cell), _p(displacement),
                                 _p(floor_of_position), _p(grid), i64(n_dims), i64(n_sd),
                                 ctypes.c_int(int(apply_boundary)))


def seeding(idx, multiplicity, extensive_attributes, seeded_particle_index,
            seeded_particle_multiplicity, seeded_particle_extensive_attributes,
            number_of_super_particles_to_inject):
    """seeding_methods.py:12-65; raises like the reference's assert when fewer than the
    requested number of empty slots exist"""
    n_attr, n_sd = extensive_attributes.shape
    lib().oracle_seeding.restype = i64
    done = lib().oracle_seeding(
        _p(idx), _p(multiplicity), _p(extensive_attributes), i64(n_attr), i64(n_sd),
        _p(seeded_particle_index), _p(seeded_particle_multiplicity),
        _p(seeded_particle_extensive_attributes), i64(len(seeded_particle_multiplicity)),
        i64(number_of_super_particles_to_inject))
    assert done == number_of_super_particles_to_inject
