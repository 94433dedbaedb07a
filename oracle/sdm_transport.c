/* sdm_transport.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE (see sdm_oracle.c).
 *
 * Plain-C restatement of the reference's Numba methods either side of the collision step
 * (SURVEY.md 8f), each function citing the file:line it follows under
 * /root/reference/PySDM/backends/impl_numba/methods/:
 *   moments, spectrum_moments                       moments_methods.py:12-182
 *   calculate_displacement (1/2/3-D)                displacement_methods.py:14-153
 *   flag_precipitated, flag_out_of_column           displacement_methods.py:155-260
 * and the Storage operators Displacement strings together
 * (/root/reference/PySDM/dynamics/displacement.py:138-153; impl_numba/storage_impl.py).
 * Serial loops in the reference's iteration order, so sums are bit-identical to a
 * NUMBA_NUM_THREADS=1 run (pinned by tests/golden/transport.npz). */
#include <math.h>
#include <stdint.h>

typedef int64_t i64;
#define ORACLE_API __attribute__((visibility("default")))

/* Numba: `x ** n` with an integer-typed n is square-and-multiply (numba/cpython/numbers.py
 * int_power_impl, exponents up to 2^16), `x ** y` with a float y is libm pow */
static double power_rank(double x, double rank, int rank_is_int)
{
    if (rank_is_int && fabs(rank) <= 65536.0) {
        i64 e = (i64)fabs(rank);
        double r = 1.0, a = x;
        while (e != 0) {
            if (e & 1) r *= a;
            e >>= 1;
            a *= a;
        }
        return rank < 0 ? 1.0 / r : r;
    }
    return pow(x, rank);
}

/* moments_methods.py:34-61 */
ORACLE_API void oracle_moments(double *moment_0, double *moments, const i64 *multiplicity,
                               const double *attr_data, const i64 *cell_id, const i64 *idx,
                               i64 length, const double *ranks, i64 n_ranks, double min_x,
                               double max_x, const double *x_attr,
                               const double *weighting_attribute, double weighting_rank,
                               int weighting_rank_is_int, int skip_division_by_m0, i64 n_cell)
{
    for (i64 c = 0; c < n_cell; ++c) moment_0[c] = 0;
    for (i64 t = 0; t < n_ranks * n_cell; ++t) moments[t] = 0;
    for (i64 p = 0; p < length; ++p) {
        const i64 i = idx[p];
        if (min_x <= x_attr[i] && x_attr[i] < max_x) {
            const double w = (double)multiplicity[i] *
                             power_rank(weighting_attribute[i], weighting_rank, weighting_rank_is_int);
            moment_0[cell_id[i]] += w;
            for (i64 k = 0; k < n_ranks; ++k)
                moments[k * n_cell + cell_id[i]] += w * pow(attr_data[i], ranks[k]);
        }
    }
    if (!skip_division_by_m0)
        for (i64 c = 0; c < n_cell; ++c)
            for (i64 k = 0; k < n_ranks; ++k)
                moments[k * n_cell + c] = moment_0[c] != 0 ? moments[k * n_cell + c] / moment_0[c] : 0;
}

/* moments_methods.py:118-149 */
ORACLE_API void oracle_spectrum_moments(double *moment_0, double *moments, const i64 *multiplicity,
                                        const double *attr_data, const i64 *cell_id,
                                        const i64 *idx, i64 length, double rank, int rank_is_int,
                                        const double *x_bins, i64 n_bins, const double *x_attr,
                                        const double *weighting_attribute, double weighting_rank,
                                        int weighting_rank_is_int, i64 n_cell)
{
    for (i64 t = 0; t < n_bins * n_cell; ++t) moment_0[t] = moments[t] = 0;
    for (i64 p = 0; p < length; ++p) {
        const i64 i = idx[p];
        for (i64 k = 0; k < n_bins; ++k)
            if (x_bins[k] <= x_attr[i] && x_attr[i] < x_bins[k + 1]) {
                const double w = (double)multiplicity[i] * power_rank(weighting_attribute[i],
                                                                      weighting_rank,
                                                                      weighting_rank_is_int);
                moment_0[k * n_cell + cell_id[i]] += w;
                moments[k * n_cell + cell_id[i]] += w * power_rank(attr_data[i], rank, rank_is_int);
                break;
            }
    }
    for (i64 t = 0; t < n_bins * n_cell; ++t)
        moments[t] = moment_0[t] != 0 ? moments[t] / moment_0[t] : 0;
}

/* displacement_methods.py:14-153; schemes: PySDM/physics/particle_advection/*.py.
 * courant: component `dim`, row-major, shape[n_dims] (one more entry along `dim`) */
ORACLE_API void oracle_calculate_displacement(int dim, int n_dims, int implicit_in_space,
                                              double *displacement, const double *courant,
                                              const i64 *shape, const i64 *cell_origin,
                                              const double *position_in_cell, i64 n_sd,
                                              double n_substeps)
{
    i64 stride[3] = {1, 1, 1};
    for (int d = n_dims - 2; d >= 0; --d) stride[d] = stride[d + 1] * shape[d + 1];
    for (i64 s = 0; s < n_sd; ++s) {
        i64 l = 0;
        for (int d = 0; d < n_dims; ++d) l += cell_origin[d * n_sd + s] * stride[d];
        const double c_l = courant[l] / n_substeps, c_r = courant[l + stride[dim]] / n_substeps;
        const double x = position_in_cell[dim * n_sd + s];
        double out = c_l * (1 - x) + c_r * x;
        if (implicit_in_space) out = out / (1 - c_r + c_l);
        displacement[dim * n_sd + s] = out;
    }
}

/* displacement_methods.py:155-186: *_z = last rows; returns the rainfall mass */
ORACLE_API double oracle_flag_precipitated(const i64 *cell_origin_z, const double *position_z,
                                           const double *water_mass, const i64 *multiplicity,
                                           i64 *idx, i64 length, i64 idx_len, i64 *healthy,
                                           double level, const double *displacement_z)
{
    double rainfall_mass = 0.0;
    for (i64 p = 0; p < length; ++p) {
        const i64 i = idx[p];
        const double z = (double)cell_origin_z[i] + position_z[i];
        if (displacement_z[i] < 0 && z < level) {
            rainfall_mass += fabs(water_mass[i]) * (double)multiplicity[i];
            idx[p] = idx_len;
            healthy[0] = 0;
        }
    }
    return rainfall_mass;
}

/* displacement_methods.py:188-206 */
ORACLE_API void oracle_flag_out_of_column(const i64 *cell_origin_z, const double *position_z,
                                          i64 *idx, i64 length, i64 idx_len, i64 *healthy,
                                          double top)
{
    for (i64 p = 0; p < length; ++p) {
        const i64 i = idx[p];
        const double z = (double)cell_origin_z[i] + position_z[i];
        if (z < 0 || z > top) {
            idx[p] = idx_len;
            healthy[0] = 0;
        }
    }
}

/* Displacement.update_position / update_cell_origin / boundary_condition,
 * dynamics/displacement.py:138-153 (Storage `+=`, `floor`, `+=`, `-=`, `%=`) */
ORACLE_API void oracle_update_position(i64 *cell_origin, double *position_in_cell,
                                       const double *displacement, i64 *floor_of_position,
                                       const i64 *grid, i64 n_dims, i64 n_sd, int apply_boundary)
{
    for (i64 t = 0; t < n_dims * n_sd; ++t) {
        position_in_cell[t] += displacement[t];
        floor_of_position[t] = (i64)floor(position_in_cell[t]);
        cell_origin[t] += floor_of_position[t];
        position_in_cell[t] -= (double)floor_of_position[t];
        if (apply_boundary) {
            const i64 d = grid[t / n_sd];
            i64 r = cell_origin[t] % d;
            if (r != 0 && ((r < 0) != (d < 0))) r += d;
            cell_origin[t] = r;
        }
    }
}

/* seeding (particle injection), impl_numba/methods/seeding_methods.py:12-44: scan ALL slots
 * in physical order; every slot with multiplicity 0, until n_inject have been filled, gets
 * idx[i] = -1, the multiplicity and the A extensive attributes of seed
 * seeded_particle_index[k] (k = 0, 1, ... in order of filling).  Returns the number injected
 * (the reference asserts it equals n_inject). */
ORACLE_API i64 oracle_seeding(i64 *idx, i64 *multiplicity, double *extensive_attributes,
                              i64 n_attr, i64 n_sd, const i64 *seeded_particle_index,
                              const i64 *seeded_particle_multiplicity,
                              const double *seeded_particle_extensive_attributes, i64 n_seeds,
                              i64 n_inject)
{
    i64 done = 0;
    for (i64 i = 0; i < n_sd; ++i) {
        if (done == n_inject) break;
        if (multiplicity[i] == 0) {
            idx[i] = -1;
            const i64 s = seeded_particle_index[done++];
            multiplicity[i] = seeded_particle_multiplicity[s];
            for (i64 a = 0; a < n_attr; ++a)
                extensive_attributes[a * n_sd + i] = seeded_particle_extensive_attributes[a * n_seeds + s];
        }
    }
    return done;
}
