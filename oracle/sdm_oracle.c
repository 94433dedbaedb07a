/*
 * sdm_oracle.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * Plain-C restatement (CPU, host pointers) of the reference's Numba backend
 * methods on the SDM collision / coalescence / breakup path.  Only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
 * may load this library.  The product (pysdm_b200) never does.
 *
 * Every function cites the reference file:line it restates (paths relative to
 * /root/reference/PySDM/backends/impl_numba/).  Pinned against the reference
 * itself: oracle/gen_golden.py runs the reference Numba backend
 * (Formulae(fastmath=False)) in the build container and stores input/output
 * vectors in tests/golden/ (npz files); tests/test_oracle_golden.py replays them.
 *
 * Semantics are the sequential ones.  `#pragma omp parallel for` appears only
 * where the reference itself uses numba.prange, so the multi-threaded timing
 * of this port (bench.py cpu_baseline, kind "port") parallelises exactly what
 * the reference parallelises.
 *
 * All floats fp64, all integers int64, flags 1-byte bool (storage.py:16-19).
 * Build: see oracle/Makefile (-O2 -ffp-contract=off: no FMA contraction, no
 * fast-math, mirroring Formulae(fastmath=False)).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

typedef int64_t i64;
typedef uint8_t u8;

#define ORACLE_API __attribute__((visibility("default")))

/* thread control for the timed CPU baseline (bench.py sweeps the thread count) */
#ifdef _OPENMP
#include <omp.h>
ORACLE_API void oracle_set_num_threads(int n) { omp_set_num_threads(n); }
ORACLE_API int oracle_get_max_threads(void) { return omp_get_max_threads(); }
#else
ORACLE_API void oracle_set_num_threads(int n) { (void)n; }
ORACLE_API int oracle_get_max_threads(void) { return 1; }
#endif

/* ---- pair_indices: methods/collisions_methods.py:16-35 ------------------ */
static inline int pair_indices(i64 i, const i64 *idx, const u8 *is_first_in_pair,
                               const double *prob_like, i64 *j, i64 *k)
{
    if (prob_like[i] == 0) {
        *j = -1;
        *k = -1;
        return 1;
    }
    i64 offset = 1 - (i64)is_first_in_pair[2 * i];
    *j = idx[2 * i + offset];
    *k = idx[2 * i + 1 + offset];
    return 0;
}

ORACLE_API void oracle_pair_indices(i64 n_pairs, const i64 *idx, const u8 *flag,
                                    const double *prob_like, i64 *j_out, i64 *k_out,
                                    u8 *skip_out)
{
    for (i64 i = 0; i < n_pairs; ++i)
        skip_out[i] = (u8)pair_indices(i, idx, flag, prob_like, &j_out[i], &k_out[i]);
}

/* ---- index methods: methods/index_methods.py ---------------------------- */

/* identity_index, index_methods.py:13-20 */
ORACLE_API void oracle_identity_index(i64 *idx, i64 n)
{
    for (i64 i = 0; i < n; ++i) idx[i] = i;
}

/* shuffle_global, index_methods.py:22-30 (serial in the reference) */
ORACLE_API void oracle_shuffle_global(i64 *idx, i64 length, const double *u01)
{
    for (i64 i = length - 1; i > 0; --i) {
        i64 j = (i64)(u01[i] * (double)(i + 1));
        i64 t = idx[i];
        idx[i] = idx[j];
        idx[j] = t;
    }
}

/* shuffle_local, index_methods.py:32-44 (prange over cells) */
ORACLE_API void oracle_shuffle_local(i64 *idx, const double *u01, const i64 *cell_start,
                                     i64 n_cell)
{
#pragma omp parallel for schedule(static)
    for (i64 c = 0; c < n_cell; ++c) {
        i64 s = cell_start[c], e = cell_start[c + 1];
        for (i64 i = e - 1; i > s; --i) {
            i64 j = (i64)((double)s + u01[i] * (double)(e - s));
            i64 t = idx[i];
            idx[i] = idx[j];
            idx[j] = t;
        }
    }
}

/* sort_by_key, index_methods.py:46-48:
 *   idx[:] = argsort(attr, kind="stable")[::-1]
 * numpy sorts NaN last; the reversed stable order is reproduced with a
 * stable merge sort on (key, position). */
static const double *g_keys;
static int cmp_key_stable(const void *a, const void *b)
{
    i64 ia = *(const i64 *)a, ib = *(const i64 *)b;
    double ka = g_keys[ia], kb = g_keys[ib];
    int na = isnan(ka), nb = isnan(kb);
    if (na != nb) return na - nb; /* NaN last */
    if (!na) {
        if (ka < kb) return -1;
        if (ka > kb) return 1;
    }
    return (ia > ib) - (ia < ib);
}
ORACLE_API void oracle_sort_by_key(i64 *idx, const double *attr, i64 n)
{
    i64 *tmp = (i64 *)malloc(sizeof(i64) * (size_t)(n > 0 ? n : 1));
    for (i64 i = 0; i < n; ++i) tmp[i] = i;
    g_keys = attr;
    qsort(tmp, (size_t)n, sizeof(i64), cmp_key_stable);
    for (i64 i = 0; i < n; ++i) idx[i] = tmp[n - 1 - i];
    free(tmp);
}

/* remove_zero_n_or_flagged, collisions_methods.py:664-680 */
ORACLE_API i64 oracle_remove_zero_n_or_flagged(const i64 *multiplicity, i64 *idx, i64 length,
                                               i64 n_sd)
{
    i64 flag = n_sd; /* len(idx) */
    i64 new_length = length;
    i64 i = 0;
    while (i < new_length) {
        if (idx[i] == flag || multiplicity[idx[i]] == 0) {
            new_length -= 1;
            idx[i] = idx[new_length];
            idx[new_length] = flag;
        } else {
            i += 1;
        }
    }
    return new_length;
}

/* _counting_sort_by_cell_id_and_update_cell_start, collisions_methods.py:682-697.
 * The "parallel" scheme (:699-741) yields the same stable order. */
ORACLE_API void oracle_counting_sort_by_cell_id(i64 *new_idx, const i64 *idx, const i64 *cell_id,
                                                const i64 *cell_idx, i64 length, i64 *cell_start,
                                                i64 cell_start_len)
{
    i64 *cell_end = cell_start;
    for (i64 i = 0; i < cell_start_len; ++i) cell_end[i] = 0;
    for (i64 i = 0; i < length; ++i) cell_end[cell_idx[cell_id[idx[i]]]] += 1;
    for (i64 i = 1; i < cell_start_len; ++i) cell_end[i] += cell_end[i - 1];
    for (i64 i = length - 1; i >= 0; --i) {
        i64 c = cell_idx[cell_id[idx[i]]];
        cell_end[c] -= 1;
        new_idx[cell_end[c]] = idx[i];
    }
}

/* cell_id, collisions_methods.py:407-416: cell_id[:] = dot(strides, cell_origin) */
ORACLE_API void oracle_cell_id(i64 *cell_id, const i64 *cell_origin, const i64 *strides,
                               i64 n_dims, i64 n_sd)
{
    for (i64 i = 0; i < n_sd; ++i) {
        i64 acc = 0;
        for (i64 d = 0; d < n_dims; ++d) acc += strides[d] * cell_origin[d * n_sd + i];
        cell_id[i] = acc;
    }
}

/* ---- pair methods: methods/pair_methods.py ------------------------------ */

/* find_pairs, pair_methods.py:34-55 */
ORACLE_API void oracle_find_pairs(const i64 *cell_start, u8 *is_first_in_pair, const i64 *cell_id,
                                  const i64 *cell_idx, const i64 *idx, i64 length)
{
    if (length < 1) return;
#pragma omp parallel for schedule(static)
    for (i64 i = 0; i < length - 1; ++i) {
        i64 ci = cell_id[idx[i]];
        int same = ci == cell_id[idx[i + 1]];
        i64 off = i - cell_start[cell_idx[ci]];
        /* Python % : result has the sign of the divisor, so (x % 2 == 0) <=> x even */
        int even = (off % 2) == 0;
        is_first_in_pair[i] = (u8)(same && even);
    }
    is_first_in_pair[length - 1] = 0;
}

/* sort_within_pair_by_attr, pair_methods.py:126-140 (attr is int64 multiplicity
 * in the collision path; a float variant is provided for completeness) */
ORACLE_API void oracle_sort_within_pair_by_attr_i64(i64 *idx, i64 length, const u8 *flag,
                                                    const i64 *attr)
{
#pragma omp parallel for schedule(static)
    for (i64 i = 0; i < length - 1; ++i)
        if (flag[i] && attr[idx[i]] < attr[idx[i + 1]]) {
            i64 t = idx[i];
            idx[i] = idx[i + 1];
            idx[i + 1] = t;
        }
}
ORACLE_API void oracle_sort_within_pair_by_attr_f64(i64 *idx, i64 length, const u8 *flag,
                                                    const double *attr)
{
#pragma omp parallel for schedule(static)
    for (i64 i = 0; i < length - 1; ++i)
        if (flag[i] && attr[idx[i]] < attr[idx[i + 1]]) {
            i64 t = idx[i];
            idx[i] = idx[i + 1];
            idx[i + 1] = t;
        }
}

/* {distance,max,min,sum,multiply}_pair, pair_methods.py:14-32,57-95,142-180.
 * op: 0 distance, 1 max, 2 min, 3 sum, 4 multiply.  The whole output (out_len,
 * i.e. n_sd//2) is zeroed first (quirk A5).  min/sum iterate to `length` in the
 * reference and rely on flag[length-1]==False; same set of pairs. */
#define PAIR_OP_BODY(T)                                                         \
    memset(out, 0, sizeof(double) * (size_t)out_len);                           \
    _Pragma("omp parallel for schedule(static)")                                \
    for (i64 i = 0; i < length - 1; ++i)                                        \
        if (flag[i]) {                                                          \
            T a = in[idx[i]], b = in[idx[i + 1]];                               \
            double r;                                                           \
            switch (op) {                                                       \
            case 0: r = fabs((double)(a - b)); break;                           \
            case 1: r = (double)(a > b ? a : b); break;                         \
            case 2: r = (double)(a < b ? a : b); break;                         \
            case 3: r = (double)(a + b); break;                                 \
            default: r = (double)(a * b); break;                                \
            }                                                                   \
            out[i / 2] = r;                                                     \
        }

ORACLE_API void oracle_pair_op_f64(int op, double *out, i64 out_len, const double *in,
                                   const u8 *flag, const i64 *idx, i64 length)
{
    PAIR_OP_BODY(double)
}
ORACLE_API void oracle_pair_op_i64(int op, double *out, i64 out_len, const i64 *in, const u8 *flag,
                                   const i64 *idx, i64 length)
{
    PAIR_OP_BODY(i64)
}

/* sort_pair, pair_methods.py:97-124 (n-sized output: larger first) */
ORACLE_API void oracle_sort_pair_f64(double *out, i64 out_len, const double *in, const u8 *flag,
                                     const i64 *idx, i64 length)
{
    memset(out, 0, sizeof(double) * (size_t)out_len);
    for (i64 i = 0; i < length - 1; ++i)
        if (flag[i]) {
            double a = in[idx[i]], b = in[idx[i + 1]];
            if (a < b) {
                out[i] = b;
                out[i + 1] = a;
            } else {
                out[i] = a;
                out[i + 1] = b;
            }
        }
}

/* ---- physics: methods/physics_methods.py:107-133 with
 * physics/particle_shape_and_density/liquid_spheres.py:17-23 --------------- */
ORACLE_API void oracle_volume_of_water_mass(double *volume, const double *mass, i64 n,
                                            double rho_w)
{
#pragma omp parallel for schedule(static)
    for (i64 i = 0; i < n; ++i) volume[i] = mass[i] / rho_w;
}
ORACLE_API void oracle_mass_of_water_volume(double *mass, const double *volume, i64 n,
                                            double rho_w)
{
#pragma omp parallel for schedule(static)
    for (i64 i = 0; i < n; ++i) mass[i] = rho_w * volume[i];
}

/* ---- normalize, collisions_methods.py:633-662 ---------------------------
 * note quirk A1: prob[d] *= norm_factor[cell_idx[cell_id[d]]] with d = pair slot */
ORACLE_API void oracle_normalize(double *prob, i64 prob_len, const i64 *cell_id,
                                 const i64 *cell_idx, const i64 *cell_start, i64 n_cell,
                                 double *norm_factor, double timestep, double dv)
{
    for (i64 i = 0; i < n_cell; ++i) {
        i64 sd_num = cell_start[i + 1] - cell_start[i];
        if (sd_num < 2)
            norm_factor[i] = 0;
        else
            norm_factor[i] = timestep / dv * (double)sd_num * (double)(sd_num - 1) / 2 /
                             (double)(sd_num / 2);
    }
#pragma omp parallel for schedule(static)
    for (i64 d = 0; d < prob_len; ++d) prob[d] *= norm_factor[cell_idx[cell_id[d]]];
}

/* ---- compute_gamma, collisions_methods.py:522-585 ------------------------ */
static inline i64 floordiv_i64(i64 a, i64 b)
{
    i64 q = a / b;
    if ((a % b != 0) && ((a < 0) != (b < 0))) q -= 1;
    return q;
}

ORACLE_API void oracle_compute_gamma(const double *prob, const double *rand, const i64 *idx,
                                     i64 length, const i64 *multiplicity, const i64 *cell_id,
                                     i64 *collision_rate_deficit, i64 *collision_rate,
                                     const u8 *flag, double *out)
{
#pragma omp parallel for schedule(static)
    for (i64 i = 0; i < length / 2; ++i) {
        out[i] = ceil(prob[i] - rand[i]);
        i64 j, k;
        if (pair_indices(i, idx, flag, out, &j, &k)) continue;
        i64 prop = floordiv_i64(multiplicity[j], multiplicity[k]);
        i64 gi = (i64)out[i];
        i64 g = gi < prop ? gi : prop;
        i64 cid = cell_id[j];
#pragma omp atomic
        collision_rate[cid] += g * multiplicity[k];
#pragma omp atomic
        collision_rate_deficit[cid] += (gi - g) * multiplicity[k];
        out[i] = (double)g;
    }
}

/* ---- coalesce + collision_coalescence, collisions_methods.py:38-59,418-477 - */
static inline void flag_zero_multiplicity(i64 j, i64 k, const i64 *multiplicity, i64 *healthy)
{
    if (multiplicity[k] == 0 || multiplicity[j] == 0) healthy[0] = 0;
}

static inline void coalesce(i64 i, i64 j, i64 k, i64 cid, i64 *multiplicity, const double *gamma,
                            double *attributes, i64 n_attr, i64 n_sd, i64 *coalescence_rate)
{
    /* float arithmetic stored into int64 (quirk A7).
     * FMA: this helper is compiled by the reference with fastmath=True whatever
     * Formulae(fastmath=...) says (decorator uses conf.JIT_FLAGS,
     * collisions_methods.py:44 + impl_numba/conf.py:5-10), and LLVM contracts
     * gamma*a_j + a_k into one fused multiply-add on every FMA-capable x86.
     * Verified on all golden vectors: fma reproduces the reference bit for bit,
     * separate mul+add does not (1-ulp differences in ~5 % of the pairs). */
    i64 add = (i64)(gamma[i] * (double)multiplicity[k]);
#pragma omp atomic
    coalescence_rate[cid] += add;
    double new_n = (double)multiplicity[j] - gamma[i] * (double)multiplicity[k];
    if (new_n > 0) {
        multiplicity[j] = (i64)new_n;
        for (i64 a = 0; a < n_attr; ++a)
            attributes[a * n_sd + k] =
                fma(gamma[i], attributes[a * n_sd + j], attributes[a * n_sd + k]);
    } else {
        multiplicity[j] = floordiv_i64(multiplicity[k], 2);
        multiplicity[k] = multiplicity[k] - multiplicity[j];
        for (i64 a = 0; a < n_attr; ++a) {
            attributes[a * n_sd + j] =
                fma(gamma[i], attributes[a * n_sd + j], attributes[a * n_sd + k]);
            attributes[a * n_sd + k] = attributes[a * n_sd + j];
        }
    }
}

ORACLE_API void oracle_collision_coalescence(i64 *multiplicity, const i64 *idx, i64 length,
                                             double *attributes, i64 n_attr, i64 n_sd,
                                             const double *gamma, i64 *healthy,
                                             const i64 *cell_id, i64 *coalescence_rate,
                                             const u8 *flag)
{
#pragma omp parallel for schedule(static)
    for (i64 i = 0; i < length / 2; ++i) {
        i64 j, k;
        if (pair_indices(i, idx, flag, gamma, &j, &k)) continue;
        coalesce(i, j, k, cell_id[j], multiplicity, gamma, attributes, n_attr, n_sd,
                 coalescence_rate);
        flag_zero_multiplicity(j, k, multiplicity, healthy);
    }
}

/* ---- adaptive time-stepping ---------------------------------------------- */

/* adaptive_sdm_end, collisions_methods.py:313-328 */
ORACLE_API i64 oracle_adaptive_sdm_end(const double *dt_left, i64 n_cell, const i64 *cell_start)
{
    i64 end = 0;
    for (i64 i = n_cell - 1; i >= 0; --i) {
        if (dt_left[i] == 0) continue;
        end = cell_start[i + 1];
        break;
    }
    return end;
}

/* Python/numba min(a, b) on floats: b if b < a else a  (NaN in `a` persists, quirk A9) */
static inline double py_min(double a, double b) { return b < a ? b : a; }
static inline double py_max(double a, double b) { return b > a ? b : a; }

/* scale_prob_for_adaptive_sdm_gamma, collisions_methods.py:330-405 */
ORACLE_API void oracle_scale_prob_for_adaptive_sdm_gamma(
    double *prob, const i64 *idx, i64 length, const i64 *multiplicity, const i64 *cell_id,
    double *dt_left, i64 n_cell, double dt, double dt_min, double dt_max, const u8 *flag,
    i64 *stats_n_substep, double *stats_dt_min)
{
    double *dt_todo = (double *)malloc(sizeof(double) * (size_t)(n_cell > 0 ? n_cell : 1));
    for (i64 c = 0; c < n_cell; ++c) dt_todo[c] = py_min(dt_left[c], dt_max);
    for (i64 i = 0; i < length / 2; ++i) { /* serial in the reference (TODO #571) */
        i64 j, k;
        if (pair_indices(i, idx, flag, prob, &j, &k)) continue;
        i64 prop = floordiv_i64(multiplicity[j], multiplicity[k]);
        double dt_optimal = dt * (double)prop / prob[i];
        i64 cid = cell_id[j];
        dt_optimal = py_max(dt_optimal, dt_min);
        dt_todo[cid] = py_min(dt_todo[cid], dt_optimal);
        stats_dt_min[cid] = py_min(stats_dt_min[cid], dt_optimal);
    }
#pragma omp parallel for schedule(static)
    for (i64 i = 0; i < length / 2; ++i) {
        i64 j, k;
        if (pair_indices(i, idx, flag, prob, &j, &k)) continue;
        prob[i] *= dt_todo[cell_id[j]] / dt;
    }
    for (i64 c = 0; c < n_cell; ++c) {
        dt_left[c] -= dt_todo[c];
        if (dt_todo[c] > 0) stats_n_substep[c] += 1;
    }
    free(dt_todo);
}

/* ---- breakup, collisions_methods.py:62-311 ------------------------------- */

/* compute_transfer_multiplicities, :62-93 */
static inline void compute_transfer_multiplicities(double gamma, i64 j, i64 k,
                                                   const i64 *multiplicity,
                                                   const double *particle_mass,
                                                   double fragment_mass_i, double max_multiplicity,
                                                   double *take_from_j_out, double *new_mult_k_out,
                                                   i64 *gamma_j_k_out, int *overflow_out)
{
    int overflow_flag = 0;
    i64 gamma_j_k = 0;
    double take_from_j_test = (double)multiplicity[k];
    double take_from_j = 0;
    double new_mult_k_test =
        ((particle_mass[j] + particle_mass[k]) * (double)multiplicity[k]) / fragment_mass_i;
    double new_mult_k = (double)multiplicity[k];
    i64 g = (i64)gamma;
    for (i64 m = 0; m < g; ++m) {
        if (new_mult_k_test > max_multiplicity) {
            overflow_flag = 1;
            break;
        }
        if (take_from_j_test > (double)multiplicity[j]) break;
        take_from_j = take_from_j_test;
        new_mult_k = new_mult_k_test;
        gamma_j_k = m + 1;
        take_from_j_test += new_mult_k_test;
        new_mult_k_test = fma(new_mult_k_test, particle_mass[j] / fragment_mass_i, new_mult_k_test);
    }
    *take_from_j_out = take_from_j;
    *new_mult_k_out = new_mult_k;
    *gamma_j_k_out = gamma_j_k;
    *overflow_out = overflow_flag;
}

/* get_new_multiplicities_and_update_attributes, :96-114 */
static inline void get_new_multiplicities_and_update_attributes(i64 j, i64 k, double *attributes,
                                                                i64 n_attr, i64 n_sd,
                                                                const i64 *multiplicity,
                                                                double take_from_j,
                                                                double new_mult_k, double *nj,
                                                                double *nk)
{
    const double recip = 1.0 / new_mult_k;
    for (i64 a = 0; a < n_attr; ++a) {
        double *ak = &attributes[a * n_sd + k];
        *ak = fma(take_from_j, attributes[a * n_sd + j], (double)multiplicity[k] * *ak) * recip;
    }
    if ((double)multiplicity[j] > take_from_j) {
        *nj = (double)multiplicity[j] - take_from_j;
        *nk = new_mult_k;
    } else {
        *nj = new_mult_k / 2;
        *nk = *nj;
        for (i64 a = 0; a < n_attr; ++a) attributes[a * n_sd + j] = attributes[a * n_sd + k];
    }
}

/* round_multiplicities_to_ints_and_update_attributes, :117-132.
 * Python round() = round-half-to-even = rint() in the default rounding mode (A8) */
static inline void round_multiplicities_to_ints_and_update_attributes(i64 j, i64 k, double nj,
                                                                      double nk,
                                                                      double *attributes,
                                                                      i64 n_attr, i64 n_sd,
                                                                      i64 *multiplicity)
{
    i64 rj = (i64)rint(nj), rk = (i64)rint(nk);
    multiplicity[j] = rj > 1 ? rj : 1;
    multiplicity[k] = rk > 1 ? rk : 1;
    double rcp_j = 1.0 / (double)multiplicity[j];
    double rcp_k = 1.0 / (double)multiplicity[k];
    for (i64 a = 0; a < n_attr; ++a) {
        attributes[a * n_sd + k] = (attributes[a * n_sd + k] * nk) * rcp_k;
        attributes[a * n_sd + j] = (attributes[a * n_sd + j] * nj) * rcp_j;
    }
}

/* break_up, :135-175.  Returns the overflow flag. */
static inline int break_up(i64 i, i64 j, i64 k, i64 cid, i64 *multiplicity, const double *gamma,
                           double *attributes, i64 n_attr, i64 n_sd, const double *fragment_mass,
                           double max_multiplicity, i64 *breakup_rate, i64 *breakup_rate_deficit,
                           const double *particle_mass)
{
    double take_from_j, new_mult_k, nj, nk;
    i64 gamma_j_k;
    int overflow;
    compute_transfer_multiplicities(gamma[i], j, k, multiplicity, particle_mass, fragment_mass[i],
                                    max_multiplicity, &take_from_j, &new_mult_k, &gamma_j_k,
                                    &overflow);
    double gamma_deficit = gamma[i] - (double)gamma_j_k;
    get_new_multiplicities_and_update_attributes(j, k, attributes, n_attr, n_sd, multiplicity,
                                                 take_from_j, new_mult_k, &nj, &nk);
    i64 add1 = gamma_j_k * multiplicity[k];
    i64 add2 = (i64)(gamma_deficit * (double)multiplicity[k]);
#pragma omp atomic
    breakup_rate[cid] += add1;
#pragma omp atomic
    breakup_rate_deficit[cid] += add2;
    round_multiplicities_to_ints_and_update_attributes(j, k, nj, nk, attributes, n_attr, n_sd,
                                                       multiplicity);
    return overflow;
}

/* break_up_while, :178-243 (Formulae(handle_all_breakups=True)) */
static inline int break_up_while(i64 i, i64 j, i64 k, i64 cid, i64 *multiplicity,
                                 const double *gamma, double *attributes, i64 n_attr, i64 n_sd,
                                 const double *fragment_mass, double max_multiplicity,
                                 i64 *breakup_rate, i64 *breakup_rate_deficit,
                                 const double *particle_mass)
{
    double gamma_deficit = gamma[i];
    int overflow_flag = 0;
    while (gamma_deficit > 0) {
        double take_from_j, new_mult_k, nj, nk;
        double gamma_j_k;
        if (multiplicity[k] == multiplicity[j]) {
            take_from_j = (double)multiplicity[j];
            new_mult_k = (particle_mass[j] + particle_mass[k]) / fragment_mass[i] *
                         (double)multiplicity[k];
            if (new_mult_k > max_multiplicity) {
                i64 add = (i64)(gamma_deficit * (double)multiplicity[k]);
#pragma omp atomic
                breakup_rate_deficit[cid] += add;
                overflow_flag = 1;
                break;
            }
            gamma_j_k = gamma_deficit;
        } else {
            if (multiplicity[k] > multiplicity[j]) {
                i64 t = j;
                j = k;
                k = t;
            }
            i64 g;
            compute_transfer_multiplicities(gamma_deficit, j, k, multiplicity, particle_mass,
                                            fragment_mass[i], max_multiplicity, &take_from_j,
                                            &new_mult_k, &g, &overflow_flag);
            gamma_j_k = (double)g;
            /* the reference loops forever when no transfer is possible (g == 0 changes
             * nothing); both the oracle and the CUDA kernel leave the loop instead */
            if (g == 0) break;
        }
        get_new_multiplicities_and_update_attributes(j, k, attributes, n_attr, n_sd, multiplicity,
                                                     take_from_j, new_mult_k, &nj, &nk);
        i64 add = (i64)(gamma_j_k * (double)multiplicity[k]);
#pragma omp atomic
        breakup_rate[cid] += add;
        gamma_deficit -= gamma_j_k;
        round_multiplicities_to_ints_and_update_attributes(j, k, nj, nk, attributes, n_attr,
                                                           n_sd, multiplicity);
    }
    i64 add = (i64)(gamma_deficit * (double)multiplicity[k]);
#pragma omp atomic
    breakup_rate_deficit[cid] += add;
    return overflow_flag;
}

/* collision_coalescence_breakup body, :246-311.  Returns the number of pairs
 * that raised the overflow flag (the reference prints a warning per pair). */
ORACLE_API i64 oracle_collision_coalescence_breakup(
    i64 *multiplicity, const i64 *idx, i64 length, double *attributes, i64 n_attr, i64 n_sd,
    const double *gamma, const double *rand, const double *Ec, const double *Eb,
    const double *fragment_mass, i64 *healthy, const i64 *cell_id, i64 *coalescence_rate,
    i64 *breakup_rate, i64 *breakup_rate_deficit, const u8 *flag, double max_multiplicity,
    const double *particle_mass, int handle_all_breakups)
{
    i64 n_overflow = 0;
    for (i64 i = 0; i < length / 2; ++i) {
        i64 j, k;
        if (pair_indices(i, idx, flag, gamma, &j, &k)) continue;
        int bouncing = rand[i] - (Ec[i] + (1 - Ec[i]) * (Eb[i])) > 0;
        if (bouncing) continue;
        if (rand[i] - Ec[i] < 0) {
            coalesce(i, j, k, cell_id[j], multiplicity, gamma, attributes, n_attr, n_sd,
                     coalescence_rate);
        } else {
            int ov = handle_all_breakups
                         ? break_up_while(i, j, k, cell_id[j], multiplicity, gamma, attributes,
                                          n_attr, n_sd, fragment_mass, max_multiplicity,
                                          breakup_rate, breakup_rate_deficit, particle_mass)
                         : break_up(i, j, k, cell_id[j], multiplicity, gamma, attributes, n_attr,
                                    n_sd, fragment_mass, max_multiplicity, breakup_rate,
                                    breakup_rate_deficit, particle_mass);
            n_overflow += ov;
        }
        flag_zero_multiplicity(j, k, multiplicity, healthy);
    }
    return n_overflow;
}
