// sdm_device.cuh -- per-pair device functions of the SDM collision step, shared by the
// method-level kernels (pairs.cu) and the fused step (fused.cu), so that both paths
// execute literally the same arithmetic.  Each function names the reference lines it
// implements (paths relative to PySDM/backends/impl_numba/methods/).
#pragma once
#include "common.cuh"

// find_pairs body, pair_methods.py:38-41
__device__ __forceinline__ u8 sdm_is_first_in_pair(i64 i, i64 sd_i, i64 sd_next,
                                                   const i64 *__restrict__ cell_id,
                                                   const i64 *__restrict__ cell_idx,
                                                   const i64 *__restrict__ cell_start)
{
    i64 ci = cell_id[sd_i];
    bool same = ci == cell_id[sd_next];
    i64 off = i - cell_start[cell_idx[ci]];
    return (u8)(same && ((off & 1) == 0));  // Python's (off % 2 == 0) also for negative off
}

// pairwise reductions, pair_methods.py:14-32,57-95,142-180
// op: 0 distance, 1 max, 2 min, 3 sum, 4 multiply
template <typename T>
__device__ __forceinline__ double sdm_pair_reduce(int op, T a, T b)
{
    switch (op) {
    case 0: {
        T d = a - b;
        return (double)(d < 0 ? -d : d);
    }
    case 1: return (double)(a > b ? a : b);   // Python max(a, b): b if b > a else a
    case 2: return (double)(b < a ? b : a);   // Python min(a, b): b if b < a else a
    case 3: return (double)(a + b);
    default: return (double)(a * b);
    }
}

// normalize, per-cell factor, collisions_methods.py:639-646 (left-to-right fp64)
__device__ __forceinline__ double sdm_norm_factor(i64 sd_num, double timestep, double dv)
{
    if (sd_num < 2) return 0.0;
    return timestep / dv * (double)sd_num * (double)(sd_num - 1) / 2.0 / (double)(sd_num / 2);
}

// compute_gamma body, collisions_methods.py:546-558.
// in: prob - rand already ceil'ed into `g_raw` by the caller (out[i] = ceil(prob-rand));
// returns the capped gamma and the two counter increments
struct GammaResult {
    double gamma;
    i64 rate_add, deficit_add;
};
__device__ __forceinline__ GammaResult sdm_gamma(double g_raw, i64 n_j, i64 n_k)
{
    GammaResult r;
    i64 prop = floordiv(n_j, n_k);
    i64 gi = (i64)g_raw;
    i64 g = gi < prop ? gi : prop;
    r.rate_add = g * n_k;
    r.deficit_add = (gi - g) * n_k;
    r.gamma = (double)g;
    return r;
}

// coalesce, collisions_methods.py:44-59.  Float arithmetic stored into int64 (quirk A7).
// The attribute update is a FUSED multiply-add: the reference compiles this helper with
// fastmath=True regardless of Formulae(fastmath) (decorator uses conf.JIT_FLAGS) and LLVM
// contracts gamma*a_j + a_k on FMA hardware -- checked bit-for-bit against the reference's
// outputs (tests/golden), see DESIGN.md "floating-point contract".
// Returns the coalescence_rate increment.
__device__ __forceinline__ i64 sdm_coalesce(i64 j, i64 k, double gamma, i64 *__restrict__ mult,
                                            double *__restrict__ attributes, i64 n_attr, i64 n_sd)
{
    i64 n_j = mult[j], n_k = mult[k];
    i64 rate_add = (i64)(gamma * (double)n_k);
    double new_n = (double)n_j - gamma * (double)n_k;
    if (new_n > 0) {
        mult[j] = (i64)new_n;
        for (i64 a = 0; a < n_attr; ++a) {
            double *row = attributes + a * n_sd;
            row[k] = fma(gamma, row[j], row[k]);
        }
    } else {
        i64 half = floordiv(n_k, 2);
        mult[j] = half;
        mult[k] = n_k - half;
        for (i64 a = 0; a < n_attr; ++a) {
            double *row = attributes + a * n_sd;
            double v = fma(gamma, row[j], row[k]);
            row[j] = v;
            row[k] = v;
        }
    }
    return rate_add;
}

// flag_zero_multiplicity, collisions_methods.py:38-41
__device__ __forceinline__ bool sdm_any_zero(i64 j, i64 k, const i64 *__restrict__ mult)
{
    return mult[k] == 0 || mult[j] == 0;
}
