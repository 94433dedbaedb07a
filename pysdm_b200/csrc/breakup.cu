// breakup.cu -- collision_coalescence_breakup: per candidate pair, bounce / coalesce / break
// up, updating multiplicities and extensive attributes in place.
//
// Reference: PySDM/backends/impl_numba/methods/collisions_methods.py
//   :62-93   compute_transfer_multiplicities      :96-114  get_new_multiplicities_and_update_attributes
//   :117-132 round_multiplicities_to_ints_...      :135-175 break_up
//   :178-243 break_up_while (Formulae(handle_all_breakups=True))
//   :246-311 collision_coalescence_breakup body    :479-520 host wrapper
//
// `particle_mass` normally IS a row of `attributes` (the water-mass attribute,
// particulator.py:179-209), so neither pointer is __restrict__: break_up_while re-reads the
// masses the previous transfer wrote.
//
// One thread per pair slot; a super-droplet belongs to at most one pair, so threads never
// touch each other's particles.  Rates go to per-cell int64 counters (warp-aggregated
// atomics).  HBM traffic per pair: gamma, rand, Ec, Eb, fragment_mass (40 B) + for pairs
// that collide the gathers/scatters of multiplicity, particle mass and A attribute rows.
//
// Floating point: the reference's helpers are compiled by Numba with fast-math whatever
// Formulae(fastmath=...) says (decorators use conf.JIT_FLAGS), so LLVM reassociates and
// contracts them.  The forms below (explicit fma, reciprocal multiplies) are the ones
// that reproduce the reference's output on tests/golden/physics.npz: integer results
// (multiplicities, rates, deficits) bit-exact, attributes bit-exact in > 99 % of the
// entries and within 1e-15 relative otherwise.  Rounding is half-to-even (quirk A8).
#include "pysdm_b200.h"
#include <math.h>

#include "sdm_device.cuh"

namespace {

struct Transfer {
    double take_from_j, new_mult_k;
    i64 gamma_j_k;
    bool overflow;
};

// :62-93
__device__ __forceinline__ Transfer compute_transfer(double gamma, double n_j, double n_k,
                                                     double m_j, double m_k, double fragment_mass,
                                                     double max_multiplicity)
{
    Transfer t;
    t.overflow = false;
    t.gamma_j_k = 0;
    double take_test = n_k;
    t.take_from_j = 0;
    double new_k_test = ((m_j + m_k) * n_k) / fragment_mass;
    t.new_mult_k = n_k;
    const double growth = m_j / fragment_mass;
    const i64 g = (i64)gamma;
    for (i64 m = 0; m < g; ++m) {
        if (new_k_test > max_multiplicity) {
            t.overflow = true;
            break;
        }
        if (take_test > n_j) break;
        t.take_from_j = take_test;
        t.new_mult_k = new_k_test;
        t.gamma_j_k = m + 1;
        take_test += new_k_test;
        new_k_test = fma(new_k_test, growth, new_k_test);
    }
    return t;
}

// :96-114; returns the (float) new multiplicities through nj, nk
__device__ __forceinline__ void new_multiplicities_and_attributes(
    i64 j, i64 k, double *attributes, i64 n_attr, i64 n_sd, double n_j, double n_k,
    double take_from_j, double new_mult_k, double &nj, double &nk)
{
    const double recip = 1.0 / new_mult_k;
    for (i64 a = 0; a < n_attr; ++a) {
        double *row = attributes + a * n_sd;
        row[k] = fma(take_from_j, row[j], n_k * row[k]) * recip;
    }
    if (n_j > take_from_j) {
        nj = n_j - take_from_j;
        nk = new_mult_k;
    } else {
        nj = new_mult_k / 2;
        nk = nj;
        for (i64 a = 0; a < n_attr; ++a) {
            double *row = attributes + a * n_sd;
            row[j] = row[k];
        }
    }
}

// :117-132
__device__ __forceinline__ void round_multiplicities(i64 j, i64 k, double nj, double nk,
                                                     double *attributes, i64 n_attr,
                                                     i64 n_sd, i64 *__restrict__ mult)
{
    i64 rj = (i64)rint(nj), rk = (i64)rint(nk);
    rj = rj > 1 ? rj : 1;
    rk = rk > 1 ? rk : 1;
    mult[j] = rj;
    mult[k] = rk;
    const double rcp_j = 1.0 / (double)rj, rcp_k = 1.0 / (double)rk;
    for (i64 a = 0; a < n_attr; ++a) {
        double *row = attributes + a * n_sd;
        row[k] = (row[k] * nk) * rcp_k;
        row[j] = (row[j] * nj) * rcp_j;
    }
}

template <bool HandleAll>
__global__ void k_collision_coalescence_breakup(
    i64 *__restrict__ mult, const i64 *__restrict__ idx, i64 n_pairs,
    double *attributes, i64 n_attr, i64 n_sd, const double *__restrict__ gamma,
    const double *__restrict__ rand, const double *__restrict__ Ec, const double *__restrict__ Eb,
    const double *__restrict__ fragment_mass, i64 *__restrict__ healthy,
    const i64 *__restrict__ cell_id, i64 *__restrict__ coalescence_rate,
    i64 *__restrict__ breakup_rate, i64 *__restrict__ breakup_rate_deficit,
    const u8 *__restrict__ flag, double max_multiplicity, const double *particle_mass,
    i64 *__restrict__ n_overflow)
{
    const i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    i64 cid = 0, coal_add = 0, break_add = 0, deficit_add = 0;
    bool active = false, overflow = false;
    if (i < n_pairs) {
        const double g = gamma[i];
        PairJK p = pair_indices(i, idx, flag, g);
        const double u = rand[i], ec = Ec[i];
        // :262-266 (bouncing pairs are left untouched)
        if (!p.skip && !(u - (ec + (1 - ec) * Eb[i]) > 0)) {
            i64 j = p.j, k = p.k;
            cid = cell_id[j];
            active = true;
            if (u - ec < 0) {
                coal_add = sdm_coalesce(j, k, g, mult, attributes, n_attr, n_sd);
            } else if (!HandleAll) {  // break_up, :135-175
                const double n_j = (double)mult[j], n_k = (double)mult[k];
                Transfer t = compute_transfer(g, n_j, n_k, particle_mass[j], particle_mass[k],
                                              fragment_mass[i], max_multiplicity);
                overflow = t.overflow;
                double nj, nk;
                new_multiplicities_and_attributes(j, k, attributes, n_attr, n_sd, n_j, n_k,
                                                  t.take_from_j, t.new_mult_k, nj, nk);
                break_add = t.gamma_j_k * mult[k];
                deficit_add = (i64)((g - (double)t.gamma_j_k) * n_k);
                round_multiplicities(j, k, nj, nk, attributes, n_attr, n_sd, mult);
            } else {  // break_up_while, :178-243
                double deficit = g;
                const double fm = fragment_mass[i];
                while (deficit > 0) {
                    double n_j = (double)mult[j], n_k = (double)mult[k];
                    double take, new_k, gamma_j_k;
                    if (mult[k] == mult[j]) {
                        take = n_j;
                        new_k = (particle_mass[j] + particle_mass[k]) / fm * n_k;
                        if (new_k > max_multiplicity) {
                            deficit_add += (i64)(deficit * n_k);
                            overflow = true;
                            break;
                        }
                        gamma_j_k = deficit;
                    } else {
                        if (mult[k] > mult[j]) {
                            i64 t = j;
                            j = k;
                            k = t;
                            n_j = (double)mult[j];
                            n_k = (double)mult[k];
                        }
                        Transfer t = compute_transfer(deficit, n_j, n_k, particle_mass[j],
                                                      particle_mass[k], fm, max_multiplicity);
                        take = t.take_from_j;
                        new_k = t.new_mult_k;
                        gamma_j_k = (double)t.gamma_j_k;
                        overflow = t.overflow;
                        // the reference never leaves this loop when no transfer is possible
                        // (gamma_j_k == 0 changes nothing); stop instead of hanging the GPU
                        if (t.gamma_j_k == 0) break;
                    }
                    double nj, nk;
                    new_multiplicities_and_attributes(j, k, attributes, n_attr, n_sd, n_j, n_k,
                                                      take, new_k, nj, nk);
                    break_add += (i64)(gamma_j_k * n_k);
                    deficit -= gamma_j_k;
                    round_multiplicities(j, k, nj, nk, attributes, n_attr, n_sd, mult);
                }
                deficit_add += (i64)(deficit * (double)mult[k]);
            }
            if (sdm_any_zero(p.j, p.k, mult)) healthy[0] = 0;  // :38-41
        }
    }
    cell_counter_add(coalescence_rate, cid, coal_add, active);
    cell_counter_add(breakup_rate, cid, break_add, active);
    cell_counter_add(breakup_rate_deficit, cid, deficit_add, active);
    if (overflow && n_overflow) atomicAdd((unsigned long long *)n_overflow, 1ull);
}

}  // namespace

PYSDM_EXPORT int pysdm_b200_collision_coalescence_breakup(
    i64 *multiplicity, const i64 *idx, i64 length, double *attributes, i64 n_attr, i64 n_sd,
    const double *gamma, const double *rand, const double *Ec, const double *Eb,
    const double *fragment_mass, i64 *healthy, const i64 *cell_id, i64 *coalescence_rate,
    i64 *breakup_rate, i64 *breakup_rate_deficit, const u8 *is_first_in_pair,
    double max_multiplicity, const double *particle_mass, int handle_all_breakups,
    i64 *n_overflow, void *stream)
{
    const i64 n_pairs = length / 2;
    if (n_pairs <= 0) return 0;
    cudaStream_t st = (cudaStream_t)stream;
    if (handle_all_breakups)
        k_collision_coalescence_breakup<true><<<grid_for(n_pairs), kThreads, 0, st>>>(
            multiplicity, idx, n_pairs, attributes, n_attr, n_sd, gamma, rand, Ec, Eb,
            fragment_mass, healthy, cell_id, coalescence_rate, breakup_rate, breakup_rate_deficit,
            is_first_in_pair, max_multiplicity, particle_mass, n_overflow);
    else
        k_collision_coalescence_breakup<false><<<grid_for(n_pairs), kThreads, 0, st>>>(
            multiplicity, idx, n_pairs, attributes, n_attr, n_sd, gamma, rand, Ec, Eb,
            fragment_mass, healthy, cell_id, coalescence_rate, breakup_rate, breakup_rate_deficit,
            is_first_in_pair, max_multiplicity, particle_mass, n_overflow);
    return LAUNCH_CHECK();
}
