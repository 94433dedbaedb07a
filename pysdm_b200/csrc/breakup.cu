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

#include "breakup.cuh"

namespace {

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
        if (!p.skip) {
            PairOutcome o = collide_pair<HandleAll>(p.j, p.k, g, u, ec, Eb[i], fragment_mass[i], mult,
                                                    attributes, n_attr, n_sd, particle_mass,
                                                    max_multiplicity, healthy);
            if (o.active) {
                cid = cell_id[p.j];
                active = true;
                coal_add = o.coal_add;
                break_add = o.break_add;
                deficit_add = o.deficit_add;
                overflow = o.overflow;
            }
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
