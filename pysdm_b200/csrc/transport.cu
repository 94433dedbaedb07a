// transport.cu -- the backend methods either side of the collision step:
//   products side:  moments, spectrum_moments        impl_numba/methods/moments_methods.py:12-182
//   transport side: calculate_displacement (1/2/3-D), flag_precipitated, flag_out_of_column
//                                                     impl_numba/methods/displacement_methods.py:14-260
// (paths relative to /root/reference/PySDM/backends/).
//
// All kernels are one thread per working position of `idx` (coalesced idx read, gathers of
// the attributes through it) and HBM-bound.  Per-cell sums are fp64 atomics, reduced first
// inside the warp over runs of equal cell id (after the sort by cell a warp usually sees one
// or two cells), so the summation order -- and with it the last bits -- differs from the
// reference's serial loop: parity is 1e-12 relative, not bit-exact (tests/test_transport_golden.py).
#include "pysdm_b200.h"
#include <math.h>

#include "common.cuh"

namespace {

constexpr int kMaxRanks = PYSDM_B200_MAX_MOMENT_RANKS;

struct Ranks {
    double r[kMaxRanks];
    int n;
};

// Numba lowers `x ** n` with an integer-typed n as square-and-multiply (exponents up to 2^16),
// `x ** y` with a float y as libm pow
__device__ __forceinline__ double power_rank(double x, double rank, bool rank_is_int)
{
    if (rank_is_int && fabs(rank) <= 65536.0) {
        long long e = (long long)fabs(rank);
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

// out[key] += v for the active lanes: one atomic per run of equal keys inside the warp.
// Must be called by all 32 lanes.
__device__ __forceinline__ void run_sum_add(double *out, i64 key, double v, bool active)
{
    const int lane = threadIdx.x & 31;
    if (!active) {
        key = -1 - lane;  // distinct from every real key and from the neighbours
        v = 0.0;
    }
    const i64 prev = __shfl_up_sync(0xffffffffu, key, 1);
    const unsigned heads = __ballot_sync(0xffffffffu, lane == 0 || prev != key);
    const int start = 31 - __clz(heads & (0xffffffffu >> (31 - lane)));
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const double t = __shfl_up_sync(0xffffffffu, v, o);
        if (lane - o >= start) v += t;
    }
    const bool tail = lane == 31 || ((heads >> (lane + 1)) & 1u);
    if (tail && active) atomicAdd(&out[key], v);
}

// moments_methods.py:34-53
__global__ void k_moments(double *__restrict__ moment_0, double *__restrict__ moments,
                          const i64 *__restrict__ multiplicity,
                          const double *__restrict__ attr_data, const i64 *__restrict__ cell_id,
                          const i64 *__restrict__ idx, i64 length, Ranks ranks, double min_x,
                          double max_x, const double *__restrict__ x_attr,
                          const double *__restrict__ weighting_attribute, double weighting_rank,
                          bool weighting_rank_is_int, i64 n_cell)
{
    const i64 p = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    bool active = false;
    i64 cid = 0, i = 0;
    double w = 0.0;
    if (p < length) {
        i = idx[p];
        const double x = x_attr[i];
        if (min_x <= x && x < max_x) {
            active = true;
            cid = cell_id[i];
            w = (double)multiplicity[i] *
                power_rank(weighting_attribute[i], weighting_rank, weighting_rank_is_int);
        }
    }
    run_sum_add(moment_0, cid, w, active);
    for (int k = 0; k < ranks.n; ++k) {
        const double v = active ? w * pow(attr_data[i], ranks.r[k]) : 0.0;
        run_sum_add(moments + (i64)k * n_cell, cid, v, active);
    }
}

// moments_methods.py:54-61 and :142-149
__global__ void k_divide_by_m0(double *__restrict__ moments, const double *__restrict__ moment_0,
                               i64 n_rows, i64 n_cell, bool m0_per_row)
{
    const i64 t = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n_rows * n_cell) return;
    const double m0 = m0_per_row ? moment_0[t] : moment_0[t % n_cell];
    moments[t] = m0 != 0 ? moments[t] / m0 : 0.0;
}

// moments_methods.py:121-141
__global__ void k_spectrum_moments(double *__restrict__ moment_0, double *__restrict__ moments,
                                   const i64 *__restrict__ multiplicity,
                                   const double *__restrict__ attr_data,
                                   const i64 *__restrict__ cell_id, const i64 *__restrict__ idx,
                                   i64 length, double rank, bool rank_is_int,
                                   const double *__restrict__ x_bins, i64 n_bins,
                                   const double *__restrict__ x_attr,
                                   const double *__restrict__ weighting_attribute,
                                   double weighting_rank, bool weighting_rank_is_int, i64 n_cell)
{
    const i64 p = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    bool active = false;
    i64 slot = 0;
    double w = 0.0, v = 0.0;
    if (p < length) {
        const i64 i = idx[p];
        const double x = x_attr[i];
        for (i64 k = 0; k < n_bins; ++k)
            if (x_bins[k] <= x && x < x_bins[k + 1]) {
                active = true;
                slot = k * n_cell + cell_id[i];
                w = (double)multiplicity[i] *
                    power_rank(weighting_attribute[i], weighting_rank, weighting_rank_is_int);
                v = w * power_rank(attr_data[i], rank, rank_is_int);
                break;
            }
    }
    run_sum_add(moment_0, slot, w, active);
    run_sum_add(moments, slot, v, active);
}

// displacement_methods.py:14-23 + physics/particle_advection/{implicit,explicit}_in_space.py
template <int kDims>
__global__ void k_displacement(int dim, bool implicit, double *__restrict__ displacement,
                               const double *__restrict__ courant, i64 s0, i64 s1,
                               const i64 *__restrict__ cell_origin,
                               const double *__restrict__ position_in_cell, i64 n_sd,
                               double n_substeps)
{
    const i64 d = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (d >= n_sd) return;
    // Arakawa-C: the courant component `dim` lives on the cell faces normal to `dim`;
    // row-major strides (s0, s1, 1) of its array, which has one more entry along `dim`
    i64 l = 0, step = 0;
    if (kDims == 1) {
        l = cell_origin[d];
        step = 1;
    } else if (kDims == 2) {
        l = cell_origin[d] * s0 + cell_origin[n_sd + d];
        step = dim == 0 ? s0 : 1;
    } else {
        l = cell_origin[d] * s0 + cell_origin[n_sd + d] * s1 + cell_origin[2 * n_sd + d];
        step = dim == 0 ? s0 : (dim == 1 ? s1 : 1);
    }
    const double c_l = courant[l] / n_substeps, c_r = courant[l + step] / n_substeps;
    const double x = position_in_cell[(i64)dim * n_sd + d];
    double out = c_l * (1 - x) + c_r * x;
    if (implicit) out = out / (1 - c_r + c_l);
    displacement[(i64)dim * n_sd + d] = out;
}

// displacement_methods.py:155-186.  rain[0] += |m| n of the flagged SDs (block-reduced)
__global__ void k_flag_precipitated(const i64 *__restrict__ cell_origin_z,
                                    const double *__restrict__ position_z,
                                    const double *__restrict__ water_mass,
                                    const i64 *__restrict__ multiplicity, i64 *__restrict__ idx,
                                    i64 length, i64 flag, i64 *__restrict__ healthy, double level,
                                    const double *__restrict__ displacement_z,
                                    double *__restrict__ rain)
{
    const i64 p = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    double mine = 0.0;
    if (p < length) {
        const i64 i = idx[p];
        const double z = (double)cell_origin_z[i] + position_z[i];
        if (displacement_z[i] < 0 && z < level) {
            mine = fabs(water_mass[i]) * (double)multiplicity[i];
            idx[p] = flag;
            healthy[0] = 0;
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mine += __shfl_xor_sync(0xffffffffu, mine, o);
    __shared__ double part[kThreads / 32];
    if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = mine;
    __syncthreads();
    if (threadIdx.x == 0) {
        double s = 0.0;
        for (int w = 0; w < kThreads / 32; ++w) s += part[w];
        if (s != 0.0) atomicAdd(rain, s);
    }
}

// displacement_methods.py:188-206
__global__ void k_flag_out_of_column(const i64 *__restrict__ cell_origin_z,
                                     const double *__restrict__ position_z,
                                     i64 *__restrict__ idx, i64 length, i64 flag,
                                     i64 *__restrict__ healthy, double top)
{
    const i64 p = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= length) return;
    const i64 i = idx[p];
    const double z = (double)cell_origin_z[i] + position_z[i];
    if (z < 0 || z > top) {
        idx[p] = flag;
        healthy[0] = 0;
    }
}

}  // namespace

PYSDM_EXPORT int pysdm_b200_moments(
    double *moment_0, double *moments, const i64 *multiplicity, const double *attr_data,
    const i64 *cell_id, const i64 *idx, i64 length, const double *host_ranks, i64 n_ranks,
    double min_x, double max_x, const double *x_attr, const double *weighting_attribute,
    double weighting_rank, int weighting_rank_is_int, int skip_division_by_m0, i64 n_cell,
    void *stream)
{
    if (n_ranks < 0 || n_ranks > kMaxRanks || n_cell < 0) return (int)cudaErrorInvalidValue;
    cudaStream_t st = (cudaStream_t)stream;
    int err = (int)cudaMemsetAsync(moment_0, 0, sizeof(double) * n_cell, st);
    if (!err && n_ranks) err = (int)cudaMemsetAsync(moments, 0, sizeof(double) * n_ranks * n_cell, st);
    if (err) return err;
    Ranks ranks;
    ranks.n = (int)n_ranks;
    for (i64 k = 0; k < n_ranks; ++k) ranks.r[k] = host_ranks[k];
    if (length > 0)
        k_moments<<<grid_for(length), kThreads, 0, st>>>(
            moment_0, moments, multiplicity, attr_data, cell_id, idx, length, ranks, min_x, max_x,
            x_attr, weighting_attribute, weighting_rank, weighting_rank_is_int != 0, n_cell);
    if (!skip_division_by_m0 && n_ranks * n_cell > 0)
        k_divide_by_m0<<<grid_for(n_ranks * n_cell), kThreads, 0, st>>>(moments, moment_0, n_ranks,
                                                                         n_cell, false);
    return LAUNCH_CHECK();
}

PYSDM_EXPORT int pysdm_b200_spectrum_moments(
    double *moment_0, double *moments, const i64 *multiplicity, const double *attr_data,
    const i64 *cell_id, const i64 *idx, i64 length, double rank, int rank_is_int,
    const double *x_bins, i64 n_bins, const double *x_attr, const double *weighting_attribute,
    double weighting_rank, int weighting_rank_is_int, i64 n_cell, void *stream)
{
    if (n_bins < 0 || n_cell < 0) return (int)cudaErrorInvalidValue;
    cudaStream_t st = (cudaStream_t)stream;
    const i64 n = n_bins * n_cell;
    if (n == 0) return 0;
    int err = (int)cudaMemsetAsync(moment_0, 0, sizeof(double) * n, st);
    if (!err) err = (int)cudaMemsetAsync(moments, 0, sizeof(double) * n, st);
    if (err) return err;
    if (length > 0)
        k_spectrum_moments<<<grid_for(length), kThreads, 0, st>>>(
            moment_0, moments, multiplicity, attr_data, cell_id, idx, length, rank, rank_is_int != 0,
            x_bins, n_bins, x_attr, weighting_attribute, weighting_rank,
            weighting_rank_is_int != 0, n_cell);
    k_divide_by_m0<<<grid_for(n), kThreads, 0, st>>>(moments, moment_0, n_bins, n_cell, true);
    return LAUNCH_CHECK();
}

PYSDM_EXPORT int pysdm_b200_calculate_displacement(
    int dim, int n_dims, int implicit_in_space, double *displacement, const double *courant,
    const i64 *courant_shape, const i64 *cell_origin, const double *position_in_cell, i64 n_sd,
    double n_substeps, void *stream)
{
    if (n_dims < 1 || n_dims > 3 || dim < 0 || dim >= n_dims) return (int)cudaErrorInvalidValue;
    if (n_sd <= 0) return 0;
    cudaStream_t st = (cudaStream_t)stream;
    const bool imp = implicit_in_space != 0;
    if (n_dims == 1)
        k_displacement<1><<<grid_for(n_sd), kThreads, 0, st>>>(
            dim, imp, displacement, courant, 0, 0, cell_origin, position_in_cell, n_sd, n_substeps);
    else if (n_dims == 2)
        k_displacement<2><<<grid_for(n_sd), kThreads, 0, st>>>(
            dim, imp, displacement, courant, courant_shape[1], 0, cell_origin, position_in_cell,
            n_sd, n_substeps);
    else
        k_displacement<3><<<grid_for(n_sd), kThreads, 0, st>>>(
            dim, imp, displacement, courant, courant_shape[1] * courant_shape[2], courant_shape[2],
            cell_origin, position_in_cell, n_sd, n_substeps);
    return LAUNCH_CHECK();
}

// ---- one advection step of the particle positions in one launch -------------------------------
// Displacement.__call__ without sedimentation (PySDM/dynamics/displacement.py:103-153) issues
// calculate_displacement per dimension, `position += displacement`, `floor`, `cell_origin +=`,
// `position -=` and `cell_origin %= grid` as seven passes over the [n_dims, n_sd] arrays.  Here
// every super-droplet does all of it in registers, in the same order with the same operations:
// the displacement of EVERY dimension is taken from the position before the update (as the
// separate calls do), then per dimension the add, the floor, the integer add, the subtraction of
// the converted floor and numpy's modulo.  `displacement` is still written (flag_precipitated
// reads it afterwards).
struct CourantField {
    const double *data[3];
    i64 s0[3], s1[3];  // row-major strides of component `dim` (it has one more entry along `dim`)
    i64 grid[3];
};

template <int kDims>
__global__ void k_displacement_step(bool implicit, double *__restrict__ displacement,
                                    const CourantField f, i64 *__restrict__ cell_origin,
                                    double *__restrict__ position_in_cell, i64 n_sd,
                                    double n_substeps)
{
    const i64 d = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (d >= n_sd) return;
    i64 org[kDims];
    double pos[kDims], disp[kDims];
#pragma unroll
    for (int k = 0; k < kDims; ++k) {
        org[k] = cell_origin[(i64)k * n_sd + d];
        pos[k] = position_in_cell[(i64)k * n_sd + d];
    }
#pragma unroll
    for (int dim = 0; dim < kDims; ++dim) {  // k_displacement, for each component
        i64 l, step;
        if (kDims == 1) {
            l = org[0];
            step = 1;
        } else if (kDims == 2) {
            l = org[0] * f.s0[dim] + org[1];
            step = dim == 0 ? f.s0[dim] : 1;
        } else {
            l = org[0] * f.s0[dim] + org[1] * f.s1[dim] + org[2];
            step = dim == 0 ? f.s0[dim] : (dim == 1 ? f.s1[dim] : 1);
        }
        const double c_l = f.data[dim][l] / n_substeps, c_r = f.data[dim][l + step] / n_substeps;
        double out = c_l * (1 - pos[dim]) + c_r * pos[dim];
        if (implicit) out = out / (1 - c_r + c_l);
        disp[dim] = out;
    }
#pragma unroll
    for (int k = 0; k < kDims; ++k) {
        const double moved = pos[k] + disp[k];        // update_position
        const i64 fl = (i64)floor(moved);             // update_cell_origin: floor -> int64
        i64 o = org[k] + fl;
        const double p = moved - (double)fl;
        const i64 g = f.grid[k];                      // boundary_condition: numpy's %
        i64 r = o % g;
        if (r != 0 && ((r < 0) != (g < 0))) r += g;
        displacement[(i64)k * n_sd + d] = disp[k];
        position_in_cell[(i64)k * n_sd + d] = p;
        cell_origin[(i64)k * n_sd + d] = r;
    }
}

PYSDM_EXPORT int pysdm_b200_displacement_step(
    int n_dims, int implicit_in_space, double *displacement, const double *const *host_courant,
    const i64 *host_courant_shapes, const i64 *host_grid, i64 *cell_origin,
    double *position_in_cell, i64 n_sd, double n_substeps, void *stream)
{
    if (n_dims < 1 || n_dims > 3 || !host_courant || !host_courant_shapes || !host_grid)
        return (int)cudaErrorInvalidValue;
    if (n_sd <= 0) return 0;
    CourantField f = {};
    for (int dim = 0; dim < n_dims; ++dim) {
        const i64 *shape = host_courant_shapes + (i64)dim * n_dims;  // shape of component `dim`
        f.data[dim] = host_courant[dim];
        if (f.data[dim] == nullptr || host_grid[dim] == 0) return (int)cudaErrorInvalidValue;
        f.s0[dim] = n_dims == 2 ? shape[1] : (n_dims == 3 ? shape[1] * shape[2] : 0);
        f.s1[dim] = n_dims == 3 ? shape[2] : 0;
        f.grid[dim] = host_grid[dim];
    }
    cudaStream_t st = (cudaStream_t)stream;
    const bool imp = implicit_in_space != 0;
    if (n_dims == 1)
        k_displacement_step<1><<<grid_for(n_sd), kThreads, 0, st>>>(
            imp, displacement, f, cell_origin, position_in_cell, n_sd, n_substeps);
    else if (n_dims == 2)
        k_displacement_step<2><<<grid_for(n_sd), kThreads, 0, st>>>(
            imp, displacement, f, cell_origin, position_in_cell, n_sd, n_substeps);
    else
        k_displacement_step<3><<<grid_for(n_sd), kThreads, 0, st>>>(
            imp, displacement, f, cell_origin, position_in_cell, n_sd, n_substeps);
    return LAUNCH_CHECK();
}

PYSDM_EXPORT int pysdm_b200_flag_precipitated(
    const i64 *cell_origin_z, const double *position_in_cell_z, const double *water_mass,
    const i64 *multiplicity, i64 *idx, i64 length, i64 idx_len, i64 *healthy,
    double precipitation_counting_level_index, const double *displacement_z,
    double *rainfall_mass, void *stream)
{
    cudaStream_t st = (cudaStream_t)stream;
    int err = (int)cudaMemsetAsync(rainfall_mass, 0, sizeof(double), st);
    if (err) return err;
    if (length > 0)
        k_flag_precipitated<<<grid_for(length), kThreads, 0, st>>>(
            cell_origin_z, position_in_cell_z, water_mass, multiplicity, idx, length, idx_len,
            healthy, precipitation_counting_level_index, displacement_z, rainfall_mass);
    return LAUNCH_CHECK();
}

PYSDM_EXPORT int pysdm_b200_flag_out_of_column(const i64 *cell_origin_z,
                                               const double *position_in_cell_z, i64 *idx,
                                               i64 length, i64 idx_len, i64 *healthy,
                                               double domain_top_level_index, void *stream)
{
    if (length <= 0) return 0;
    k_flag_out_of_column<<<grid_for(length), kThreads, 0, (cudaStream_t)stream>>>(
        cell_origin_z, position_in_cell_z, idx, length, idx_len, healthy, domain_top_level_index);
    return LAUNCH_CHECK();
}
