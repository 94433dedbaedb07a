// breakup.cuh -- one candidate pair of collision_coalescence_breakup as a device function, shared
// by the method-level kernel (breakup.cu) and the fused breakup kernel (breakup_fused.cu).
// Reference: PySDM/backends/impl_numba/methods/collisions_methods.py
//   :62-93   compute_transfer_multiplicities      :96-114  get_new_multiplicities_and_update_attributes
//   :117-132 round_multiplicities_to_ints_...      :135-175 break_up
//   :178-243 break_up_while (Formulae(handle_all_breakups=True))
//   :246-311 collision_coalescence_breakup body
#pragma once
#include <math.h>

#include "sdm_device.cuh"

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

struct PairOutcome {
    bool active, overflow;  // active: the pair did not bounce (its cell's counters get the adds)
    i64 coal_add, break_add, deficit_add;
};

// :262-311 for one pair (j, k) with gamma g != 0: bounce / coalesce / break up
template <bool HandleAll>
__device__ __forceinline__ PairOutcome collide_pair(i64 pj, i64 pk, double g, double u, double ec,
                                                    double eb, double fm_i, i64 *mult,
                                                    double *attributes, i64 n_attr, i64 n_sd,
                                                    const double *particle_mass,
                                                    double max_multiplicity, i64 *healthy)
{
    PairOutcome o = {false, false, 0, 0, 0};
    if (u - (ec + (1 - ec) * eb) > 0) return o;  // bouncing pairs are left untouched (:262-266)
    i64 j = pj, k = pk;
    i64 coal_add = 0, break_add = 0, deficit_add = 0;
    bool overflow = false;
    o.active = true;
    {
    if (u - ec < 0) {
        coal_add = sdm_coalesce(j, k, g, mult, attributes, n_attr, n_sd);
    } else if (!HandleAll) {  // break_up, :135-175
        const double n_j = (double)mult[j], n_k = (double)mult[k];
        Transfer t = compute_transfer(g, n_j, n_k, particle_mass[j], particle_mass[k],
                                      fm_i, max_multiplicity);
        overflow = t.overflow;
        double nj, nk;
        new_multiplicities_and_attributes(j, k, attributes, n_attr, n_sd, n_j, n_k,
                                          t.take_from_j, t.new_mult_k, nj, nk);
        break_add = t.gamma_j_k * mult[k];
        deficit_add = (i64)((g - (double)t.gamma_j_k) * n_k);
        round_multiplicities(j, k, nj, nk, attributes, n_attr, n_sd, mult);
    } else {  // break_up_while, :178-243
        double deficit = g;
        const double fm = fm_i;
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
    if (sdm_any_zero(pj, pk, mult)) healthy[0] = 0;  // :38-41
    }
    o.coal_add = coal_add;
    o.break_add = break_add;
    o.deficit_add = deficit_add;
    o.overflow = overflow;
    return o;
}
