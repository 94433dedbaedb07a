// pairs.cu -- method-level kernels of the SDM collision step on candidate pairs:
// find_pairs, sort_within_pair_by_attr, pairwise reductions, normalize,
// scale_prob_for_adaptive_sdm_gamma, adaptive_sdm_end, compute_gamma,
// collision_coalescence(_breakup).  One launch per reference backend method
// (drop-in granularity); the fused step lives in fused.cu and shares sdm_device.cuh.
#include "pysdm_b200.h"
#include "sdm_device.cuh"

// ---------------------------------------------------------------------------------
// find_pairs  (reference: impl_numba/methods/pair_methods.py:34-55)
// ---------------------------------------------------------------------------------
__global__ void k_find_pairs(const i64 *__restrict__ cell_start, u8 *__restrict__ flag,
                             const i64 *__restrict__ cell_id, const i64 *__restrict__ cell_idx,
                             const i64 *__restrict__ idx, i64 length)
{
    i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= length) return;
    if (i == length - 1) {
        flag[i] = 0;
        return;
    }
    flag[i] = sdm_is_first_in_pair(i, idx[i], idx[i + 1], cell_id, cell_idx, cell_start);
}

PYSDM_EXPORT int pysdm_b200_find_pairs(const i64 *cell_start, u8 *is_first_in_pair,
                                       const i64 *cell_id, const i64 *cell_idx, const i64 *idx,
                                       i64 length, void *stream)
{
    if (length <= 0) return 0;
    k_find_pairs<<<grid_for(length), kThreads, 0, (cudaStream_t)stream>>>(
        cell_start, is_first_in_pair, cell_id, cell_idx, idx, length);
    return LAUNCH_CHECK();
}

// ---------------------------------------------------------------------------------
// sort_within_pair_by_attr  (pair_methods.py:126-140); ties keep their order
// ---------------------------------------------------------------------------------
template <typename T>
__global__ void k_sort_within_pair(i64 *__restrict__ idx, i64 length, const u8 *__restrict__ flag,
                                   const T *__restrict__ attr)
{
    i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= length - 1 || !flag[i]) return;
    i64 a = idx[i], b = idx[i + 1];
    if (attr[a] < attr[b]) {
        idx[i] = b;
        idx[i + 1] = a;
    }
}

PYSDM_EXPORT int pysdm_b200_sort_within_pair_by_attr(i64 *idx, i64 length, const u8 *flag,
                                                     const void *attr, int attr_is_int,
                                                     void *stream)
{
    if (length < 2) return 0;
    cudaStream_t st = (cudaStream_t)stream;
    if (attr_is_int)
        k_sort_within_pair<i64><<<grid_for(length), kThreads, 0, st>>>(idx, length, flag,
                                                                       (const i64 *)attr);
    else
        k_sort_within_pair<double><<<grid_for(length), kThreads, 0, st>>>(idx, length, flag,
                                                                          (const double *)attr);
    return LAUNCH_CHECK();
}

// ---------------------------------------------------------------------------------
// {distance,max,min,sum,multiply}_pair  (pair_methods.py:14-32,57-95,142-180)
// The WHOLE output (out_len = n_sd//2 slots) is zeroed first, like the reference
// (quirk A5); slot i//2 is then written for every flagged i.  One pass over pair
// slots: slot q is owned by the flagged position among {2q, 2q+1}.
// ---------------------------------------------------------------------------------
template <typename T>
__global__ void k_pair_op(int op, double *__restrict__ out, i64 out_len, const T *__restrict__ in,
                          const u8 *__restrict__ flag, const i64 *__restrict__ idx, i64 length)
{
    i64 q = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= out_len) return;
    double r = 0.0;
    i64 i = 2 * q;
    // flag[length-1] is False by construction; never read idx[length]
    if (i < length - 1 && flag[i]) {
    } else if (i + 1 < length - 1 && flag[i + 1]) {
        i = i + 1;
    } else {
        i = -1;
    }
    if (i >= 0) {
        T a = in[idx[i]], b = in[idx[i + 1]];
        r = sdm_pair_reduce<T>(op, a, b);
    }
    out[q] = r;
}

PYSDM_EXPORT int pysdm_b200_pair_op(int op, double *out, i64 out_len, const void *in, int in_is_int,
                                    const u8 *flag, const i64 *idx, i64 length, void *stream)
{
    if (out_len <= 0) return 0;
    cudaStream_t st = (cudaStream_t)stream;
    if (in_is_int)
        k_pair_op<i64><<<grid_for(out_len), kThreads, 0, st>>>(op, out, out_len, (const i64 *)in,
                                                               flag, idx, length);
    else
        k_pair_op<double><<<grid_for(out_len), kThreads, 0, st>>>(
            op, out, out_len, (const double *)in, flag, idx, length);
    return LAUNCH_CHECK();
}

// sort_pair (pair_methods.py:97-124): n-sized output, larger value first
__global__ void k_sort_pair(double *__restrict__ out, i64 out_len, const double *__restrict__ in,
                            const u8 *__restrict__ flag, const i64 *__restrict__ idx, i64 length)
{
    i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= out_len) return;
    double r = 0.0;
    if (i < length - 1 && flag[i]) {
        double a = in[idx[i]], b = in[idx[i + 1]];
        r = a < b ? b : a;
    } else if (i >= 1 && i - 1 < length - 1 && flag[i - 1]) {
        double a = in[idx[i - 1]], b = in[idx[i]];
        r = a < b ? a : b;
    }
    out[i] = r;
}

PYSDM_EXPORT int pysdm_b200_sort_pair(double *out, i64 out_len, const double *in, const u8 *flag,
                                      const i64 *idx, i64 length, void *stream)
{
    if (out_len <= 0) return 0;
    k_sort_pair<<<grid_for(out_len), kThreads, 0, (cudaStream_t)stream>>>(out, out_len, in, flag,
                                                                         idx, length);
    return LAUNCH_CHECK();
}

// ---------------------------------------------------------------------------------
// normalize  (collisions_methods.py:633-662), quirk A1 reproduced:
//   prob[d] *= norm_factor[cell_idx[cell_id[d]]] with d the PAIR SLOT number
// ---------------------------------------------------------------------------------
__global__ void k_norm_factor(const i64 *__restrict__ cell_start, i64 n_cell,
                              double *__restrict__ norm_factor, double timestep, double dv)
{
    i64 c = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (c < n_cell) norm_factor[c] = sdm_norm_factor(cell_start[c + 1] - cell_start[c], timestep, dv);
}

__global__ void k_normalize(double *__restrict__ prob, i64 prob_len, const i64 *__restrict__ cell_id,
                            const i64 *__restrict__ cell_idx,
                            const double *__restrict__ norm_factor)
{
    i64 d = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (d < prob_len) prob[d] *= norm_factor[cell_idx[cell_id[d]]];
}

PYSDM_EXPORT int pysdm_b200_normalize(double *prob, i64 prob_len, const i64 *cell_id,
                                      const i64 *cell_idx, const i64 *cell_start, i64 n_cell,
                                      double *norm_factor, double timestep, double dv, void *stream)
{
    cudaStream_t st = (cudaStream_t)stream;
    if (n_cell > 0)
        k_norm_factor<<<grid_for(n_cell), kThreads, 0, st>>>(cell_start, n_cell, norm_factor,
                                                             timestep, dv);
    if (prob_len > 0)
        k_normalize<<<grid_for(prob_len), kThreads, 0, st>>>(prob, prob_len, cell_id, cell_idx,
                                                             norm_factor);
    return LAUNCH_CHECK();
}

// ---------------------------------------------------------------------------------
// compute_gamma  (collisions_methods.py:522-585); `out` may alias `prob`
// ---------------------------------------------------------------------------------
__global__ void k_compute_gamma(const double *prob, const double *__restrict__ rand,
                                const i64 *__restrict__ idx, i64 n_pairs,
                                const i64 *__restrict__ mult, const i64 *__restrict__ cell_id,
                                i64 *__restrict__ rate_deficit, i64 *__restrict__ rate,
                                const u8 *__restrict__ flag, double *out)
{
    i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    bool active = false;
    i64 cid = 0;
    GammaResult g = {0.0, 0, 0};
    if (i < n_pairs) {
        double g_raw = ceil(prob[i] - rand[i]);
        PairJK p = pair_indices(i, idx, flag, g_raw);
        if (p.skip) {
            out[i] = g_raw;
        } else {
            g = sdm_gamma(g_raw, mult[p.j], mult[p.k]);
            cid = cell_id[p.j];
            out[i] = g.gamma;
            active = true;
        }
    }
    cell_counter_add(rate, cid, g.rate_add, active);
    cell_counter_add(rate_deficit, cid, g.deficit_add, active);
}

PYSDM_EXPORT int pysdm_b200_compute_gamma(const double *prob, const double *rand, const i64 *idx,
                                          i64 length, const i64 *multiplicity, const i64 *cell_id,
                                          i64 *collision_rate_deficit, i64 *collision_rate,
                                          const u8 *is_first_in_pair, double *out, void *stream)
{
    i64 n_pairs = length / 2;
    if (n_pairs <= 0) return 0;
    k_compute_gamma<<<grid_for(n_pairs), kThreads, 0, (cudaStream_t)stream>>>(
        prob, rand, idx, n_pairs, multiplicity, cell_id, collision_rate_deficit, collision_rate,
        is_first_in_pair, out);
    return LAUNCH_CHECK();
}

// ---------------------------------------------------------------------------------
// scale_prob_for_adaptive_sdm_gamma  (collisions_methods.py:330-405)
// The reference's serial middle loop is a per-cell min: done with atomics here.
// ---------------------------------------------------------------------------------
__global__ void k_adaptive_dt_todo_init(const double *__restrict__ dt_left, double dt_max,
                                        double *__restrict__ dt_todo, i64 n_cell)
{
    i64 c = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (c < n_cell) dt_todo[c] = py_min(dt_left[c], dt_max);
}

__global__ void k_adaptive_dt_optimal(const double *__restrict__ prob, const i64 *__restrict__ idx,
                                      i64 n_pairs, const i64 *__restrict__ mult,
                                      const i64 *__restrict__ cell_id, double dt, double dt_min,
                                      const u8 *__restrict__ flag, double *dt_todo,
                                      double *stats_dt_min)
{
    i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_pairs) return;
    double pr = prob[i];
    PairJK p = pair_indices(i, idx, flag, pr);
    if (p.skip) return;
    i64 prop = floordiv(mult[p.j], mult[p.k]);
    double dt_optimal = dt * (double)prop / pr;
    i64 cid = cell_id[p.j];
    dt_optimal = py_max(dt_optimal, dt_min);
    // min(cur, x) = x if x < cur else cur: a NaN `cur` stays NaN (quirk A9)
    double cur = dt_todo[cid];
    if (cur == cur && dt_optimal < cur) atomic_min_f64(&dt_todo[cid], dt_optimal);
    cur = stats_dt_min[cid];
    if (cur == cur && dt_optimal < cur) atomic_min_f64(&stats_dt_min[cid], dt_optimal);
}

__global__ void k_adaptive_scale(double *__restrict__ prob, const i64 *__restrict__ idx,
                                 i64 n_pairs, const i64 *__restrict__ cell_id, double dt,
                                 const u8 *__restrict__ flag, const double *__restrict__ dt_todo)
{
    i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_pairs) return;
    double pr = prob[i];
    PairJK p = pair_indices(i, idx, flag, pr);
    if (p.skip) return;
    prob[i] = pr * (dt_todo[cell_id[p.j]] / dt);
}

__global__ void k_adaptive_finish(double *__restrict__ dt_left, const double *__restrict__ dt_todo,
                                  i64 *__restrict__ n_substep, i64 n_cell)
{
    i64 c = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= n_cell) return;
    double t = dt_todo[c];
    dt_left[c] -= t;
    if (t > 0) n_substep[c] += 1;
}

// dt_todo: caller-provided scratch of n_cell doubles (the reference allocates it per call)
PYSDM_EXPORT int pysdm_b200_scale_prob_for_adaptive_sdm_gamma(
    double *prob, const i64 *idx, i64 length, const i64 *multiplicity, const i64 *cell_id,
    double *dt_left, i64 n_cell, double dt, double dt_min, double dt_max,
    const u8 *is_first_in_pair, i64 *stats_n_substep, double *stats_dt_min, double *dt_todo,
    void *stream)
{
    cudaStream_t st = (cudaStream_t)stream;
    if (n_cell <= 0) return 0;
    const i64 n_pairs = length / 2;
    k_adaptive_dt_todo_init<<<grid_for(n_cell), kThreads, 0, st>>>(dt_left, dt_max, dt_todo, n_cell);
    if (n_pairs > 0) {
        k_adaptive_dt_optimal<<<grid_for(n_pairs), kThreads, 0, st>>>(
            prob, idx, n_pairs, multiplicity, cell_id, dt, dt_min, is_first_in_pair, dt_todo,
            stats_dt_min);
        k_adaptive_scale<<<grid_for(n_pairs), kThreads, 0, st>>>(prob, idx, n_pairs, cell_id, dt,
                                                                 is_first_in_pair, dt_todo);
    }
    k_adaptive_finish<<<grid_for(n_cell), kThreads, 0, st>>>(dt_left, dt_todo, stats_n_substep,
                                                             n_cell);
    return LAUNCH_CHECK();
}

// ---------------------------------------------------------------------------------
// adaptive_sdm_end  (collisions_methods.py:313-328): last cell (by position) with
// dt_left != 0 -> cell_start[i+1], else 0.  Host result: stream is synchronised.
// ---------------------------------------------------------------------------------
__global__ void k_adaptive_end(const double *__restrict__ dt_left, i64 n_cell,
                               const i64 *__restrict__ cell_start, i64 *__restrict__ result)
{
    // result (zeroed by the launcher) = max over cells i with dt_left[i] != 0 of cell_start[i+1]
    // = cell_start[last such i + 1], because cell_start never decreases
    __shared__ i64 best[kThreads / 32];
    i64 m = -1;
    for (i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x; i < n_cell;
         i += (i64)gridDim.x * blockDim.x)
        if (dt_left[i] != 0) m = i;  // i increases: the last hit is the largest for this thread
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        i64 t = __shfl_xor_sync(0xffffffffu, m, o);
        m = t > m ? t : m;
    }
    if ((threadIdx.x & 31) == 0) best[threadIdx.x >> 5] = m;
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < kThreads / 32; ++w) m = best[w] > m ? best[w] : m;
        if (m >= 0) atomicMax((long long *)result, (long long)cell_start[m + 1]);
    }
}

PYSDM_EXPORT int pysdm_b200_adaptive_sdm_end(const double *dt_left, i64 n_cell,
                                             const i64 *cell_start, i64 *host_result,
                                             i64 *dev_scratch, void *stream)
{
    cudaStream_t st = (cudaStream_t)stream;
    int err = (int)cudaMemsetAsync(dev_scratch, 0, sizeof(i64), st);
    if (err) return err;
    int blocks = grid_for(n_cell);
    if (blocks > 148) blocks = 148;
    k_adaptive_end<<<blocks, kThreads, 0, st>>>(dt_left, n_cell, cell_start, dev_scratch);
    err = LAUNCH_CHECK();
    if (err) return err;
    err = (int)cudaMemcpyAsync(host_result, dev_scratch, sizeof(i64), cudaMemcpyDeviceToHost, st);
    if (err) return err;
    return (int)cudaStreamSynchronize(st);
}

// ---------------------------------------------------------------------------------
// collision_coalescence  (collisions_methods.py:418-477)
// ---------------------------------------------------------------------------------
__global__ void k_collision_coalescence(i64 *__restrict__ mult, const i64 *__restrict__ idx,
                                        i64 n_pairs, double *__restrict__ attributes, i64 n_attr,
                                        i64 n_sd, const double *__restrict__ gamma,
                                        i64 *__restrict__ healthy, const i64 *__restrict__ cell_id,
                                        i64 *__restrict__ coalescence_rate,
                                        const u8 *__restrict__ flag)
{
    i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    bool active = false;
    i64 cid = 0, add = 0;
    if (i < n_pairs) {
        double g = gamma[i];
        PairJK p = pair_indices(i, idx, flag, g);
        if (!p.skip) {
            cid = cell_id[p.j];
            add = sdm_coalesce(p.j, p.k, g, mult, attributes, n_attr, n_sd);
            if (sdm_any_zero(p.j, p.k, mult)) healthy[0] = 0;
            active = true;
        }
    }
    cell_counter_add(coalescence_rate, cid, add, active);
}

PYSDM_EXPORT int pysdm_b200_collision_coalescence(i64 *multiplicity, const i64 *idx, i64 length,
                                                  double *attributes, i64 n_attr, i64 n_sd,
                                                  const double *gamma, i64 *healthy,
                                                  const i64 *cell_id, i64 *coalescence_rate,
                                                  const u8 *is_first_in_pair, void *stream)
{
    i64 n_pairs = length / 2;
    if (n_pairs <= 0) return 0;
    k_collision_coalescence<<<grid_for(n_pairs), kThreads, 0, (cudaStream_t)stream>>>(
        multiplicity, idx, n_pairs, attributes, n_attr, n_sd, gamma, healthy, cell_id,
        coalescence_rate, is_first_in_pair);
    return LAUNCH_CHECK();
}
