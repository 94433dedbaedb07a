"""numpy-in / numpy-out adapter around pysdm_b200.methods (the C-ABI bindings) with the
call signatures of oracle/sdm_oracle.py, so that one test body checks both the CPU
oracle and the CUDA path against the same golden vectors.  Arrays are uploaded, the
CUDA kernels run, and the (in-place) results are copied back into the caller's arrays."""

import numpy as np
import torch

from pysdm_b200 import methods as dev

DEVICE = "cuda"


class _Mirror:
    """uploads numpy arrays on entry; copies the mutated ones back on exit"""

    def __init__(self, *arrays):
        self.host = arrays
        self.dev = [torch.from_numpy(np.ascontiguousarray(a)).to(DEVICE) for a in arrays]

    def back(self, *which):
        torch.cuda.synchronize()
        for i in which:
            h = self.host[i]
            d = self.dev[i].cpu().numpy()
            h[...] = d.view(h.dtype) if d.dtype != h.dtype else d


def _run(fun, arrays, mutated, *extra_before, **kw):
    m = _Mirror(*arrays)
    out = fun(*m.dev, **kw)
    m.back(*mutated)
    return out


def identity_index(idx):
    _run(dev.identity_index, (idx,), (0,))


def shuffle_global(idx, length, u01):
    m = _Mirror(idx, u01)
    dev.shuffle_global(m.dev[0], length, m.dev[1])
    m.back(0)


def shuffle_local(idx, u01, cell_start):
    m = _Mirror(idx, u01, cell_start)
    dev.shuffle_local(m.dev[0], m.dev[1], m.dev[2], int(cell_start[-1]))
    m.back(0)


def sort_by_key(idx, attr):
    _run(dev.sort_by_key, (idx, attr), (0,))


def remove_zero_n_or_flagged(multiplicity, idx, length):
    m = _Mirror(multiplicity, idx)
    n = dev.remove_zero_n_or_flagged(m.dev[0], m.dev[1], length)
    m.back(1)
    return n


def counting_sort_by_cell_id(new_idx, idx, cell_id, cell_idx, length, cell_start):
    m = _Mirror(new_idx, idx, cell_id, cell_idx, cell_start)
    dev.counting_sort_by_cell_id(m.dev[0], m.dev[1], m.dev[2], m.dev[3], length, m.dev[4])
    m.back(0, 4)


def cell_id(cell_id_out, cell_origin, strides):
    m = _Mirror(cell_id_out, np.ascontiguousarray(cell_origin, dtype=np.int64),
                np.ascontiguousarray(np.asarray(strides, dtype=np.int64).ravel()))
    dev.cell_id(m.dev[0], m.dev[1], m.dev[2])
    m.back(0)


def find_pairs(cell_start, is_first_in_pair, cell_id_, cell_idx, idx, length):
    m = _Mirror(cell_start, is_first_in_pair, cell_id_, cell_idx, idx)
    dev.find_pairs(m.dev[0], m.dev[1], m.dev[2], m.dev[3], m.dev[4], length)
    m.back(1)


def sort_within_pair_by_attr(idx, length, is_first_in_pair, attr):
    m = _Mirror(idx, is_first_in_pair, attr)
    dev.sort_within_pair_by_attr(m.dev[0], length, m.dev[1], m.dev[2])
    m.back(0)


def pair_op(op, data_out, data_in, is_first_in_pair, idx, length):
    m = _Mirror(data_out, data_in, is_first_in_pair, idx)
    dev.pair_op(op, m.dev[0], m.dev[1], m.dev[2], m.dev[3], length)
    m.back(0)


def sort_pair(data_out, data_in, is_first_in_pair, idx, length):
    m = _Mirror(data_out, data_in, is_first_in_pair, idx)
    dev.sort_pair(m.dev[0], m.dev[1], m.dev[2], m.dev[3], length)
    m.back(0)


def volume_of_water_mass(volume, mass, rho_w):
    m = _Mirror(volume, mass)
    dev.volume_of_water_mass(m.dev[0], m.dev[1], rho_w)
    m.back(0)


def mass_of_water_volume(mass, volume, rho_w):
    m = _Mirror(mass, volume)
    dev.mass_of_water_volume(m.dev[0], m.dev[1], rho_w)
    m.back(0)


def normalize(prob, cell_id_, cell_idx, cell_start, norm_factor, timestep, dv):
    m = _Mirror(prob, cell_id_, cell_idx, cell_start, norm_factor)
    dev.normalize(m.dev[0], m.dev[1], m.dev[2], m.dev[3], m.dev[4], timestep, dv)
    m.back(0, 4)


def compute_gamma(prob, rand, idx, length, multiplicity, cell_id_, collision_rate_deficit,
                  collision_rate, is_first_in_pair, out):
    alias = out is prob
    m = _Mirror(prob, rand, idx, multiplicity, cell_id_, collision_rate_deficit, collision_rate,
                is_first_in_pair, out)
    d = m.dev
    dev.compute_gamma(d[0], d[1], d[2], length, d[3], d[4], d[5], d[6], d[7],
                      d[0] if alias else d[8])
    if alias:
        m.back(0, 5, 6)
    else:
        m.back(5, 6, 8)


def collision_coalescence(multiplicity, idx, length, attributes, gamma, healthy, cell_id_,
                          coalescence_rate, is_first_in_pair):
    m = _Mirror(multiplicity, idx, attributes, gamma, healthy, cell_id_, coalescence_rate,
                is_first_in_pair)
    d = m.dev
    dev.collision_coalescence(d[0], d[1], length, d[2], d[3], d[4], d[5], d[6], d[7])
    m.back(0, 2, 4, 6)


def adaptive_sdm_end(dt_left, cell_start):
    m = _Mirror(dt_left, cell_start)
    return dev.adaptive_sdm_end(m.dev[0], m.dev[1])


def scale_prob_for_adaptive_sdm_gamma(prob, idx, length, multiplicity, cell_id_, dt_left, dt,
                                      dt_range, is_first_in_pair, stats_n_substep, stats_dt_min):
    m = _Mirror(prob, idx, multiplicity, cell_id_, dt_left, is_first_in_pair, stats_n_substep,
                stats_dt_min)
    d = m.dev
    dev.scale_prob_for_adaptive_sdm_gamma(d[0], d[1], length, d[2], d[3], d[4], dt, dt_range,
                                          d[5], d[6], d[7])
    m.back(0, 4, 6, 7)


# ---- physics rows (tests/test_physics_golden.py) ---------------------------------------------
def power(x, exponent):
    m = _Mirror(x)
    dev.elementwise("pow", m.dev[0], a=m.dev[0], scalar=exponent)
    m.back(0)


def product_scalar(out, in_, scalar):
    m = _Mirror(out, in_)
    dev.elementwise("mul", m.dev[0], a=m.dev[1], scalar=scalar)
    m.back(0)


def interpolation(output, radius, factor, b, c):
    m = _Mirror(output, radius, b, c)
    dev.interpolation(m.dev[0], m.dev[1], factor, m.dev[2], m.dev[3])
    m.back(0)


def terminal_velocity_rogers_yau(values, radius, params):
    m = _Mirror(values, radius)
    dev.terminal_velocity_rogers_yau(m.dev[0], m.dev[1], params)
    m.back(0)


def power_series(values, radius, prefactors, powers):
    m = _Mirror(values, radius)
    dev.power_series(m.dev[0], m.dev[1], prefactors, powers)
    m.back(0)


def linear_collection_efficiency(params, int_mask, output, radii, is_first_in_pair, idx, length,
                                 unit):
    m = _Mirror(output, radii, is_first_in_pair, idx)
    dev.linear_collection_efficiency(params, int_mask, m.dev[0], m.dev[1], m.dev[2], m.dev[3],
                                     length, unit)
    m.back(0)


def slams_fragmentation(n_fragment, frag_volume, x_plus_y, probs, rand, vmin, nfmax):
    m = _Mirror(n_fragment, frag_volume, x_plus_y, probs, rand)
    dev.slams_fragmentation(*m.dev, vmin, nfmax)
    m.back(0, 1, 3)


def exp_fragmentation(n_fragment, scale, frag_volume, x_plus_y, rand, vmin, nfmax, tol=1e-5):
    m = _Mirror(n_fragment, frag_volume, x_plus_y, rand)
    d = m.dev
    dev.exp_fragmentation(d[0], scale, d[1], d[2], d[3], vmin, nfmax, tol)
    m.back(0, 1)


def feingold1988_fragmentation(n_fragment, scale, frag_volume, x_plus_y, rand, fragtol, vmin,
                               nfmax):
    m = _Mirror(n_fragment, frag_volume, x_plus_y, rand)
    d = m.dev
    dev.feingold1988_fragmentation(d[0], scale, d[1], d[2], d[3], fragtol, vmin, nfmax)
    m.back(0, 1)


def gauss_fragmentation(n_fragment, mu, sigma, frag_volume, x_plus_y, rand, vmin, nfmax):
    m = _Mirror(n_fragment, frag_volume, x_plus_y, rand)
    d = m.dev
    dev.gauss_fragmentation(d[0], mu, sigma, d[1], d[2], d[3], vmin, nfmax)
    m.back(0, 1)


def straub_fragmentation(n_fragment, CW, gam, ds, frag_volume, v_max, x_plus_y, rand, vmin,
                         nfmax, Nr1, Nr2, Nr3, Nr4, Nrt, d34):
    m = _Mirror(n_fragment, CW, gam, ds, frag_volume, v_max, x_plus_y, rand, Nr1, Nr2, Nr3, Nr4,
                Nrt, d34)
    d = m.dev
    dev.straub_fragmentation(*d[:8], vmin, nfmax, *d[8:])
    m.back(0, 4, 8, 9, 10, 11, 12, 13)


def ll82_fragmentation(n_fragment, CKE, W, W2, St, ds, dl, dcoal, frag_volume, x_plus_y, rand,
                       vmin, nfmax, Rf, Rs, Rd, tol=1e-8):
    m = _Mirror(n_fragment, CKE, W, W2, St, ds, dl, dcoal, frag_volume, x_plus_y, rand, Rf, Rs,
                Rd)
    d = m.dev
    dev.ll82_fragmentation(*d[:11], vmin, nfmax, *d[11:], tol)
    m.back(0, 8, 10, 11, 12, 13)


def ll82_coalescence_check(Ec, dl):
    m = _Mirror(Ec, dl)
    dev.ll82_coalescence_check(*m.dev)
    m.back(0)


def collision_coalescence_breakup(multiplicity, idx, length, attributes, gamma, rand, Ec, Eb,
                                  fragment_mass, healthy, cell_id_, coalescence_rate,
                                  breakup_rate, breakup_rate_deficit, is_first_in_pair,
                                  max_multiplicity, particle_mass, handle_all_breakups=False):
    m = _Mirror(multiplicity, idx, attributes, gamma, rand, Ec, Eb, fragment_mass, healthy,
                cell_id_, coalescence_rate, breakup_rate, breakup_rate_deficit,
                is_first_in_pair, particle_mass)
    d = m.dev
    n = dev.collision_coalescence_breakup(d[0], d[1], length, d[2], d[3], d[4], d[5], d[6], d[7],
                                          d[8], d[9], d[10], d[11], d[12], d[13],
                                          max_multiplicity, d[14], handle_all_breakups,
                                          count_overflows=True)
    m.back(0, 2, 8, 10, 11, 12)
    return n


# ---- products and transport side ---------------------------------------------------------------
def moments(moment_0, moments_, multiplicity, attr_data, cell_id_, idx, length, ranks, min_x,
            max_x, x_attr, weighting_attribute, weighting_rank, skip_division_by_m0):
    m = _Mirror(moment_0, moments_, multiplicity, attr_data, cell_id_, idx, x_attr,
                weighting_attribute)
    d = m.dev
    dev.moments(d[0], d[1], d[2], d[3], d[4], d[5], length, ranks, min_x, max_x, d[6], d[7],
                weighting_rank, skip_division_by_m0)
    m.back(0, 1)


def spectrum_moments(moment_0, moments_, multiplicity, attr_data, cell_id_, idx, length, rank,
                     x_bins, x_attr, weighting_attribute, weighting_rank):
    m = _Mirror(moment_0, moments_, multiplicity, attr_data, cell_id_, idx, x_bins, x_attr,
                weighting_attribute)
    d = m.dev
    dev.spectrum_moments(d[0], d[1], d[2], d[3], d[4], d[5], length, rank, d[6], d[7], d[8],
                         weighting_rank)
    m.back(0, 1)


def calculate_displacement(dim, displacement, courant, cell_origin, position_in_cell, n_substeps,
                           implicit_in_space=True):
    m = _Mirror(displacement, courant, cell_origin, position_in_cell)
    dev.calculate_displacement(dim, *m.dev, n_substeps, implicit_in_space=implicit_in_space)
    m.back(0)


def flag_precipitated(cell_origin, position_in_cell, water_mass, multiplicity, idx, length,
                      healthy, precipitation_counting_level_index, displacement):
    m = _Mirror(cell_origin, position_in_cell, water_mass, multiplicity, idx, healthy,
                displacement)
    d = m.dev
    rain = dev.flag_precipitated(d[0], d[1], d[2], d[3], d[4], length, d[5],
                                 precipitation_counting_level_index, d[6])
    m.back(4, 5)
    return rain


def flag_out_of_column(cell_origin, position_in_cell, idx, length, healthy,
                       domain_top_level_index):
    m = _Mirror(cell_origin, position_in_cell, idx, healthy)
    dev.flag_out_of_column(*m.dev[:3], length, m.dev[3], domain_top_level_index)
    m.back(2, 3)


def update_position(cell_origin, position_in_cell, displacement, floor_of_position, grid,
                    apply_boundary=True):
    """the Storage operator sequence of Displacement (dynamics/displacement.py:138-153)
    through pysdm_b200.storage.Storage"""
    from pysdm_b200.storage import Storage  # pylint: disable=import-outside-toplevel

    s_origin, s_pos = Storage.from_ndarray(cell_origin), Storage.from_ndarray(position_in_cell)
    s_disp, s_tmp = Storage.from_ndarray(displacement), Storage.from_ndarray(floor_of_position)
    s_pos += s_disp
    s_tmp.floor(s_pos)
    s_origin += s_tmp
    s_pos -= s_tmp
    if apply_boundary:
        s_origin %= Storage.from_ndarray(grid)
    cell_origin[...] = s_origin.to_ndarray()
    position_in_cell[...] = s_pos.to_ndarray()
    floor_of_position[...] = s_tmp.to_ndarray()


def seeding(idx, multiplicity, extensive_attributes, seeded_particle_index,
            seeded_particle_multiplicity, seeded_particle_extensive_attributes,
            number_of_super_particles_to_inject):
    m = _Mirror(idx, multiplicity, extensive_attributes, seeded_particle_index,
                seeded_particle_multiplicity, seeded_particle_extensive_attributes)
    dev.seeding(*m.dev, number_of_super_particles_to_inject)
    m.back(0, 1, 2)
