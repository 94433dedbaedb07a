/*
 * pysdm_b200.h -- C-ABI of libpysdm_b200.so: the SDM collision / coalescence / breakup
 * path of PySDM as hand-written CUDA for sm_100a (NVIDIA B200).
 *
 * This is the drop-in boundary.  Each entry point replaces one method of the
 * reference's backend classes (the "L1 backend" of PySDM: the mix-ins listed in
 * PySDM/backends/numba.py:18-33); the reference method is cited next to it as
 * file:line, paths relative to PySDM/backends/impl_numba/ unless they start with PySDM/.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer unless its name starts with `host_`;
 *   - floats are fp64, integers int64, flags 1-byte bools (storage.py:16-19);
 *   - `stream` is a cudaStream_t passed as void* (NULL = default stream); calls are
 *     asynchronous unless they have a `host_` out-parameter, in which case the stream
 *     is synchronised before returning;
 *   - the return value is 0 or a cudaError_t;
 *   - no entry point allocates: scratch is caller-provided (`*_workspace_bytes`), or, for
 *     pysdm_b200_shuffle_local_static, a fixed device array inside the library;
 *   - nothing here touches host memory other than `host_*` out-parameters and there is
 *     no CPU implementation behind any symbol.
 */
#ifndef PYSDM_B200_H
#define PYSDM_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- index methods: methods/index_methods.py --------------------------------------- */

/* identity_index, index_methods.py:13-20 */
int pysdm_b200_identity_index(int64_t *idx, int64_t n, void *stream);

/* shuffle_global, index_methods.py:22-30 (serial swap chain i = length-1 .. 1,
 * j = int(u01[i] * (i + 1))); replayed bit-exactly in parallel. */
size_t pysdm_b200_shuffle_workspace_bytes(int64_t length);
int pysdm_b200_shuffle_global(int64_t *idx, int64_t length, const double *u01, void *ws,
                              size_t ws_bytes, void *stream);

/* shuffle_local, index_methods.py:32-44 (one swap chain per cell, j drawn over the
 * whole cell); `length` = cell_start[n_cell]. */
int pysdm_b200_shuffle_local(int64_t *idx, const double *u01, const int64_t *cell_start,
                             int64_t n_cell, int64_t length, void *ws, size_t ws_bytes,
                             void *stream);

/* shuffle_local with NO allocation and no caller-provided workspace.  It has no counterpart in
 * the reference and was added for callers that cannot
 * call cudaMalloc (code under CUDA graph capture, serving loops, harnesses that forbid it).
 * Same kernels and the same result, bit for bit, as pysdm_b200_shuffle_local; the scratch is a
 * file-scope `static __device__` array in index.cu, sized for at most
 * PYSDM_B200_SHUFFLE_STATIC_MAX_LENGTH positions (6 MiB, held by every GPU that loads the
 * library until the process exits).  A longer `length` is rejected with cudaErrorInvalidValue
 * and a message on stderr.  The first call on a device looks the array's device address up
 * (cudaGetSymbolAddress; a call with length 0 does only that) -- do it before a graph capture.
 * All calls share the array: calls that overlap on the GPU (different streams or threads) race. */
#define PYSDM_B200_SHUFFLE_STATIC_MAX_LENGTH 262144
int pysdm_b200_shuffle_local_static(int64_t *idx, const double *u01, const int64_t *cell_start,
                                    int64_t n_cell, int64_t length, void *stream);

/* CellCaretaker.__call__, methods/collisions_methods.py:608-629 with the counting sorts
 * :682-741: stable sort of idx[:length] by cell_idx[cell_id[idx[i]]] into new_idx and
 * cell_start[0..n_cell] (exclusive prefix sums).  The caller swaps idx <-> new_idx. */
size_t pysdm_b200_sort_by_cell_workspace_bytes(int64_t length);
int pysdm_b200_sort_by_cell_id(int64_t *new_idx, const int64_t *idx, const int64_t *cell_id,
                               const int64_t *cell_idx, int64_t length, int64_t *cell_start,
                               int64_t n_cell, void *ws, size_t ws_bytes, void *stream);

/* sort_by_key, index_methods.py:46-48: idx[:] = argsort(attr, stable)[::-1] */
size_t pysdm_b200_sort_by_key_workspace_bytes(int64_t n);
int pysdm_b200_sort_by_key(int64_t *idx, const double *attr, int64_t n, void *ws,
                           size_t ws_bytes, void *stream);

/* The same, queued speculatively in the adaptive sub-step loop
 * (PySDM/dynamics/collisions/collision.py:178-190, where `cell_idx.sort_by_key(dt_left)` opens
 * every sub-step): nothing happens when the device word *skip_if_zero is 0 -- the working
 * length the previous sub-step's adaptive_sdm_end left (pysdm_b200_fused_args.adaptive_end),
 * i.e. the loop has ended.  NULL = always sort.  Looked at only for n <= 65536 keys. */
int pysdm_b200_sort_by_key_if(int64_t *idx, const double *attr, int64_t n, void *ws,
                              size_t ws_bytes, const int64_t *skip_if_zero, void *stream);

/* remove_zero_n_or_flagged, methods/collisions_methods.py:664-680: the reference's
 * serial swap-with-last compaction, same surviving order; new length -> *host_new_length */
size_t pysdm_b200_remove_zero_workspace_bytes(int64_t length);
int pysdm_b200_remove_zero_n_or_flagged(const int64_t *multiplicity, int64_t *idx,
                                        int64_t length, int64_t n_sd,
                                        int64_t *host_new_length, void *ws, size_t ws_bytes,
                                        void *stream);

/* cell_id, methods/collisions_methods.py:407-416: cell_id = strides . cell_origin;
 * cell_origin is [n_dims, n_sd] row-major */
int pysdm_b200_cell_id(int64_t *cell_id, const int64_t *cell_origin, const int64_t *strides,
                       int64_t n_dims, int64_t n_sd, void *stream);

/* ---- pair methods: methods/pair_methods.py ----------------------------------------- */

/* find_pairs, pair_methods.py:34-55 */
int pysdm_b200_find_pairs(const int64_t *cell_start, uint8_t *is_first_in_pair,
                          const int64_t *cell_id, const int64_t *cell_idx, const int64_t *idx,
                          int64_t length, void *stream);

/* sort_within_pair_by_attr, pair_methods.py:126-140; attr is int64 (attr_is_int) or fp64 */
int pysdm_b200_sort_within_pair_by_attr(int64_t *idx, int64_t length,
                                        const uint8_t *is_first_in_pair, const void *attr,
                                        int attr_is_int, void *stream);

/* {distance,max,min,sum,multiply}_pair, pair_methods.py:14-32,57-95,142-180
 * op: 0 distance, 1 max, 2 min, 3 sum, 4 multiply; the whole out[0..out_len) is zeroed
 * first, like the reference does */
int pysdm_b200_pair_op(int op, double *out, int64_t out_len, const void *in, int in_is_int,
                       const uint8_t *is_first_in_pair, const int64_t *idx, int64_t length,
                       void *stream);

/* sort_pair, pair_methods.py:97-124 */
int pysdm_b200_sort_pair(double *out, int64_t out_len, const double *in,
                         const uint8_t *is_first_in_pair, const int64_t *idx, int64_t length,
                         void *stream);

/* ---- collision methods: methods/collisions_methods.py ------------------------------ */

/* normalize, collisions_methods.py:633-662 (caller PySDM/particulator.py:87-96).
 * prob[d] *= norm_factor[cell_idx[cell_id[d]]], d = pair slot (reference behaviour) */
int pysdm_b200_normalize(double *prob, int64_t prob_len, const int64_t *cell_id,
                         const int64_t *cell_idx, const int64_t *cell_start, int64_t n_cell,
                         double *norm_factor, double timestep, double dv, void *stream);

/* compute_gamma, collisions_methods.py:522-585; `out` may alias `prob` */
int pysdm_b200_compute_gamma(const double *prob, const double *rand, const int64_t *idx,
                             int64_t length, const int64_t *multiplicity,
                             const int64_t *cell_id, int64_t *collision_rate_deficit,
                             int64_t *collision_rate, const uint8_t *is_first_in_pair,
                             double *out, void *stream);

/* scale_prob_for_adaptive_sdm_gamma, collisions_methods.py:330-405;
 * dt_todo: scratch of n_cell doubles (the reference allocates it per call, :355) */
int pysdm_b200_scale_prob_for_adaptive_sdm_gamma(
    double *prob, const int64_t *idx, int64_t length, const int64_t *multiplicity,
    const int64_t *cell_id, double *dt_left, int64_t n_cell, double dt, double dt_min,
    double dt_max, const uint8_t *is_first_in_pair, int64_t *stats_n_substep,
    double *stats_dt_min, double *dt_todo, void *stream);

/* adaptive_sdm_end, collisions_methods.py:313-328 -> *host_result */
int pysdm_b200_adaptive_sdm_end(const double *dt_left, int64_t n_cell,
                                const int64_t *cell_start, int64_t *host_result,
                                int64_t *dev_scratch, void *stream);

/* collision_coalescence, collisions_methods.py:418-477 (+ coalesce :44-59,
 * flag_zero_multiplicity :38-41); attributes is [n_attr, n_sd] row-major */
int pysdm_b200_collision_coalescence(int64_t *multiplicity, const int64_t *idx, int64_t length,
                                     double *attributes, int64_t n_attr, int64_t n_sd,
                                     const double *gamma, int64_t *healthy,
                                     const int64_t *cell_id, int64_t *coalescence_rate,
                                     const uint8_t *is_first_in_pair, void *stream);

/* ---- Storage arithmetic: storage.py:54-100,203-213, storage_impl.py:11-83 ---------- */
/* op: 0 add 1 sub 2 mul 3 div 4 div_if_not_zero 5 pow(sign(x)|x|^s) 6 axpy(a + s*b)
 *     7 floor 8 exp 9 abs 10 mod; b == NULL selects the scalar operand */
int pysdm_b200_elementwise(int op, double *out, const double *a, const double *b,
                           double scalar, int64_t n, void *stream);
int pysdm_b200_row_modulo(double *out, const double *divisor, int64_t n_rows, int64_t n_cols,
                          void *stream);
/* Storage.amin / amax (storage.py:109-119): result -> host_result[0]; host_result[1] != 0
 * when a NaN was seen (numpy propagates NaN) */
size_t pysdm_b200_reduce_scratch_bytes(void);
int pysdm_b200_minmax_f64(const double *in, int64_t n, int is_max, double *host_result,
                          void *scratch, void *stream);
int pysdm_b200_minmax_i64(const int64_t *in, int64_t n, int is_max, int64_t *host_result,
                          void *scratch, void *stream);

/* ---- Random: random.py:9-19 (call contract); Philox4x32-10 counter-based generator --- */
int pysdm_b200_philox_uniform(double *out, int64_t n, uint64_t seed, uint64_t call,
                              void *stream);
int pysdm_b200_philox_raw(uint32_t *dev_out4, uint32_t c0, uint32_t c1, uint32_t c2,
                          uint32_t c3, uint32_t k0, uint32_t k1, void *stream);

/* ---- physics: methods/physics_methods.py:107-133 ----------------------------------- */
int pysdm_b200_volume_of_water_mass(double *volume, const double *mass, int64_t n,
                                    double rho_w, void *stream);
int pysdm_b200_mass_of_water_volume(double *mass, const double *volume, int64_t n,
                                    double rho_w, void *stream);

/* largest cell population max_c(cell_start[c+1]-cell_start[c]) -> *host_result */
int pysdm_b200_max_cell_size(const int64_t *cell_start, int64_t n_cell, int64_t *host_result,
                             int64_t *dev_scratch, void *stream);

/* ---- fused collision step --------------------------------------------------------- */
/* One launch = one Collision.step() (PySDM/dynamics/collisions/collision.py:196-234) of a
 * `Coalescence` dynamic with the "local" croupier on a population whose idx is sorted
 * by cell: shuffle_local, find_pairs, sort_within_pair_by_attr(multiplicity), collision
 * kernel, max_pair, normalize, /= substeps, compute_gamma, collision_coalescence --
 * with bit-identical results in every API-visible array.  Cells larger than
 * pysdm_b200_fused_max_cell_size() are refused (cudaErrorInvalidConfiguration): the
 * caller then runs the method-level entry points above. */
typedef struct pysdm_b200_fused_args {
    const int64_t *idx_in;      /* [n_sd] sorted by cell; working length `length` */
    int64_t *idx_out;           /* [n_sd] permuted order out (caller swaps the buffers) */
    const int64_t *cell_start;  /* [n_cell + 1] */
    const int64_t *cell_idx;    /* [n_cell] */
    const int64_t *cell_id;     /* [n_sd] */
    int64_t *multiplicity;      /* [n_sd] */
    double *attributes;         /* [n_attr, n_sd] extensive attributes, row-major */
    const double *u01;          /* [>= length] uniforms for the shuffle; NULL = Philox */
    const double *rand;         /* [n_sd / 2] uniforms for gamma; NULL = Philox */
    uint8_t *is_first_in_pair;  /* [n_sd] */
    double *gamma;              /* [n_sd / 2] */
    double *norm_factor;        /* [n_cell] */
    int64_t *collision_rate, *collision_rate_deficit, *coalescence_rate; /* [n_cell] */
    int64_t *healthy;           /* [1] */
    int64_t *status;            /* [1] set != 0 if a batch of cells did not fit */
    int64_t length, n_sd, n_cell, n_attr, mass_row;
    int64_t cells_per_batch;    /* filled in by the launcher */
    uint64_t seed, call_u01, call_rand; /* Philox stream: Random(seed), call numbers */
    double timestep, dv, rho_w, substeps;
    double kernel_a, kernel_b;  /* Golovin: b; ConstantK: a; Linear: a, b;
                                   Geometric: a = PI * collection_efficiency;
                                   Hydrodynamic: a = PI */
    int32_t kernel;             /* 0 Golovin, 1 ConstantK, 2 Linear, 3 Geometric (geometric.py:15-22),
                                   4 Hydrodynamic / Parameterized (impl/parameterized.py:13-30); 3 and 4
                                   gather `radius` and `fall_velocity` instead of the water mass */
    int32_t reserved;
    int64_t max_cell_size;      /* upper bound of cell_start[c+1]-cell_start[c] (host hint) */
    double *nf_slot;            /* [n_sd / 2] scratch: norm_factor[cell_idx[cell_id[d]]] per pair
                                   slot d (the factor `normalize` applies, :647-648).  Adaptive
                                   launches over cells of at most pysdm_b200_fused_max_cell_size()
                                   super-droplets follow that chain in the kernel instead
                                   (cell_idx changes with every sub-step) and leave it untouched */
    int64_t refresh_nf;         /* != 0: recompute norm_factor and nf_slot first.  Must be set
                                   whenever cell_start, cell_idx, cell_id, timestep or dv changed
                                   since the previous call with the same nf_slot buffer */
    int64_t *gate;              /* [1] or NULL.  Lets the caller queue step k+1 before it has read
                                   `healthy` of step k (the reference reads it after every step,
                                   PySDM/impl/particle_attributes.py:67-73): a launch does nothing
                                   at all when gate[0] is neither 0 nor `epoch`, and a launch that
                                   clears healthy[0] stores its `epoch` in gate[0].  The caller
                                   zeroes gate[0] when it has removed the zero-multiplicity SDs. */
    int64_t epoch;              /* != 0, different for consecutive launches */
    uint32_t philox_round_keys[20]; /* filled in by the launcher from `seed` */
    int64_t shuffle_only;       /* set by pysdm_b200_shuffle_local_batched: stop after the shuffle */
    const double *radius;        /* [n_sd] kernels 3, 4: the "radius" attribute */
    const double *fall_velocity; /* [n_sd] kernels 3, 4: the "relative fall velocity" attribute */
    double berry_params[13];     /* kernel 4: (A, B, D1, D2, E1, E2, F1, F2, G1, G2, G3, Mf, Mg) as */
    int64_t berry_int_mask;      /* in pysdm_b200_linear_collection_efficiency */
    double berry_unit;
    /* ---- adaptive time step (Collision.__call__ sub-step loop, collision.py:174-194; one launch =
     * one sub-step incl. scale_prob_for_adaptive_sdm_gamma, collisions_methods.py:330-405) ---- */
    int64_t adaptive;            /* != 0: the fields below are used, `substeps` is ignored */
    double *dt_left;             /* [n_cell] by cell id: -= dt_todo */
    int64_t *stats_n_substep;    /* [n_cell] by cell id: += (dt_todo > 0) */
    double *stats_dt_min;        /* [n_cell] by cell id: min with the pairs' dt_optimal (NaN sticks) */
    double dt_min, dt_max;       /* dt_coal_range */
    const int64_t *rank_to_cell; /* [n_cell] cell id held by rank r of cell_start, i.e. the inverse of
                                    the cell_idx the LAST sort by cell used (cell_idx itself is
                                    re-sorted every sub-step without a re-sort of idx, quirk A2) */
    int64_t *rank_info;          /* [n_cell] scratch: cell id | parity of cell_start[cell_idx[cell id]]
                                    << 62, rebuilt by every adaptive launch */
    int64_t *adaptive_end;       /* [2] or NULL: [0] = max over cells i with dt_left[i] != 0 after
                                    this sub-step of cell_start[i + 1] (adaptive_sdm_end, :313-328;
                                    zeroed by the launch), [1] = 1 if some stats_dt_min == dt_min */
    const int64_t *length_dev;   /* [1] or NULL: working length read on the device instead of
                                    `length` (the adaptive_end[0] of the previous sub-step); a
                                    launch that finds 0 there does nothing */
    /* ---- dependent attributes (kernels 3, 4): the orchestration derives volume, radius and fall
     * velocity from the water mass (attributes/physics/{volume,radius,terminal_velocity}.py) and
     * recomputes all of them after every step; with dep_radius != NULL the launch instead
     * refreshes them for exactly the super-droplets a coalescence changed -- same operation
     * sequence, so the arrays equal a full recomputation bit for bit */
    double *dep_volume;          /* [n_sd] or NULL ("volume" attribute) */
    double *dep_radius;          /* [n_sd] or NULL; normally == radius */
    double *dep_velocity;        /* [n_sd] or NULL; normally == fall_velocity */
    const double *dep_table_b, *dep_table_c; /* Gunn-Kinzer interpolation tables (`a`, `b` of
                                    dynamics/terminal_velocity/gunn_and_kinzer.py:108-125) */
    double dep_radius_scale, dep_radius_exponent; /* 1 / PI_4_3 and 1 / 3 as the host computes them */
    double dep_table_factor, dep_radius_max;      /* Interpolation.factor, .maximum_radius */
    int64_t *dep_status;         /* [1]: set to 1 if a refreshed radius exceeds dep_radius_max (the
                                    reference raises ValueError at the next recalculation) */
    /* ---- cells larger than pysdm_b200_fused_max_cell_size() (single-cell boxes: BASELINE
     * configs[0..2]): the step runs as the parallel replay of the cells' serial shuffle chains
     * (pysdm_b200_shuffle_local) followed by one streaming kernel over the pair slots (two for the
     * adaptive time step), all inside the one call; needs device scratch */
    void *ws;                    /* scratch of >= pysdm_b200_fused_workspace_bytes(n_sd, n_cell)
                                    bytes, or NULL: large cells are then refused as before */
    uint64_t ws_bytes;
    int64_t sorted_length;       /* cell_start[n_cell] when the working length was cut below it
                                    (adaptive loop); 0 = `length` */
    int64_t skip_coalescence;    /* != 0: stop after compute_gamma (breakup: pysdm_b200_fused_breakup
                                    takes over) */
    int64_t prepass_done;        /* adaptive, != 0: pysdm_b200_adaptive_presort sorted cell_idx for
                                    this sub-step and zeroed *adaptive_end, norm_factor is valid:
                                    no pre-pass launch here, rank_info is neither read nor written */
} pysdm_b200_fused_args;

int64_t pysdm_b200_fused_max_cell_size(void);
size_t pysdm_b200_fused_workspace_bytes(int64_t n_sd, int64_t n_cell);
/* shuffle_local (index_methods.py:32-44) for populations whose cells all hold at most
 * max_cell_size <= pysdm_b200_fused_max_cell_size() super-droplets: the staging + swap-chain
 * phases of the fused step alone (shared memory, one launch) instead of the general
 * five-kernel replay of pysdm_b200_shuffle_local.  In place; same result bit for bit. */
int pysdm_b200_shuffle_local_batched(int64_t *idx, const double *u01, const int64_t *cell_start,
                                     int64_t n_cell, int64_t n_sd, int64_t max_cell_size,
                                     int64_t *status, void *stream);
int pysdm_b200_fused_collision_step(const pysdm_b200_fused_args *host_args, void *stream);

/* Opens one sub-step of the adaptive loop (PySDM/dynamics/collisions/collision.py:181,
 * `cell_idx.sort_by_key(dt_left)`, index_methods.py:46-48) with ONE launch between two launches
 * of the fused step: the sort of the n_cell keys, which also zeroes *adaptive_end (the word the
 * fused sub-step accumulates adaptive_sdm_end in).  Follow it with
 * pysdm_b200_fused_collision_step(prepass_done = 1): that launch then derives, per rank of
 * cell_start, the cell id and the parity find_pairs counts from under the re-sorted cell_idx
 * (pair_methods.py:40) itself instead of reading rank_info, and takes norm_factor as it is.
 * norm_factor != NULL: the per-cell factors of normalize (collisions_methods.py:639-646) are
 * (re)computed from cell_start, timestep and dv first -- needed once after every sort by cell.
 * `skip_if_zero` as in pysdm_b200_sort_by_key_if (a skipped launch still zeroes *adaptive_end, so
 * the sub-steps queued behind it do nothing either);
 * ws: pysdm_b200_sort_by_key_workspace_bytes(n_cell). */
int pysdm_b200_adaptive_presort(int64_t *cell_idx, const double *dt_left, int64_t n_cell, void *ws,
                                size_t ws_bytes, const int64_t *skip_if_zero,
                                const int64_t *cell_start, double *norm_factor, double timestep,
                                double dv, int64_t *adaptive_end, void *stream);

/* ---- fused breakup: the second half of a Collision.step() with enable_breakup ---------------
 * (PySDM/dynamics/collisions/collision.py:206-234).  pysdm_b200_fused_collision_step with
 * skip_coalescence != 0 runs the step up to and including compute_gamma; this entry point then
 * handles, in ONE launch over the pair slots with gamma != 0: the coalescence efficiency
 * (ConstEc, Straub2010Ec: coalescence_efficiencies/straub2010.py:24-47 -- a chain of ~15
 * pairwise Storage operations in the reference --, or a Parameterized one such as Berry1967:
 * _parameterized.py:16-24), the breakup efficiency (ConstEb), the fragment mass (AlwaysN:
 * breakup_fragmentations/always_n.py:12-15; Straub2010Nf: straub2010.py:44-102 +
 * fragmentation_methods.py:218-260, another ~30 operations, + volume_based.py:15-17), the
 * uniforms of the two RandomGeneratorOptimizerNoPair streams, and collision_coalescence_breakup
 * itself (collisions_methods.py:246-311).  The operation sequences are the reference's. */
typedef struct pysdm_b200_breakup_args {
    int64_t *multiplicity;        /* [n_sd] */
    const int64_t *idx;           /* [n_sd] pairs as sort_within_pair_by_attr left them */
    double *attributes;           /* [n_attr, n_sd] */
    const double *gamma;          /* [n_sd / 2] after compute_gamma */
    const uint8_t *is_first_in_pair; /* [n_sd] */
    const int64_t *cell_id;       /* [n_sd] */
    int64_t *healthy;             /* [1] */
    int64_t *coalescence_rate, *breakup_rate, *breakup_rate_deficit; /* [n_cell] */
    int64_t *n_overflow;          /* [1] or NULL: counts pairs that hit max_multiplicity */
    int64_t length, n_sd, n_attr, mass_row;
    const double *rand_proc, *rand_frag; /* [n_sd / 2] each, or NULL = Philox (seed, call_*) */
    uint64_t seed, call_proc, call_frag;
    const double *volume, *radius, *fall_velocity; /* [n_sd] derived attributes (Straub, Berry) */
    int32_t ec_kind;              /* 0 ConstEc, 1 Straub2010Ec, 2 Parameterized (linear_collection_efficiency ** 2) */
    int32_t nf_kind;              /* 0 AlwaysN, 1 Straub2010Nf */
    double ec_const, eb_const, nf_n;
    double ec_params[13];         /* ec_kind 2, as in pysdm_b200_linear_collection_efficiency */
    int64_t ec_int_mask;
    double ec_unit;
    double vmin, nfmax;           /* fragmentation_limiters */
    int32_t has_nfmax, handle_all_breakups;
    double max_multiplicity;
    /* constants as the host's Formulae carries them (products are formed on the host exactly as
     * the reference's Python does) */
    double rho_w;                 /* const.rho_w */
    double six_over_pi;           /* 6 / const.PI */
    double pi_sgm_w;              /* const.PI * const.sgm_w */
    double two_thirds;            /* 2 / 3 */
    double nf_sc_factor;          /* const.PI * const.sgm_w * (6 / const.PI) ** (2 / 3) */
    double half_rho_w;            /* const.rho_w / 2 */
    double micro_joule;           /* si.uJ */
    /* derived attributes kept current for the super-droplets the launch changed (as dep_* of
     * pysdm_b200_fused_args); dep_radius == NULL: left to the caller */
    double *dep_volume, *dep_radius, *dep_velocity;
    const double *dep_table_b, *dep_table_c;
    double dep_radius_scale, dep_radius_exponent, dep_table_factor, dep_radius_max;
    int64_t *dep_status;
} pysdm_b200_breakup_args;
int pysdm_b200_fused_breakup(const pysdm_b200_breakup_args *host_args, void *stream);
/* out[in[c]] = c for c < n: the rank -> cell map (`rank_to_cell` above) of the cell -> rank map
 * `cell_idx` a sort by cell has just used (collisions_methods.py:692-697 consume it that way) */
int pysdm_b200_invert_index(int64_t *out, const int64_t *in, int64_t n, void *stream);
/* volume, radius, fall velocity of every super-droplet from the water mass in one pass
 * (volume_of_water_mass, Radius.recalculate = product + `**=`, Gunn-Kinzer `interpolation`; the
 * five-launch chain of attributes/physics/{volume,radius,terminal_velocity}.py); any output may be
 * NULL.  status[0] is set to 1 if a radius exceeds radius_max (no velocity is written for it). */
int pysdm_b200_dependent_attributes(const double *mass, int64_t n, double rho_w, double radius_scale,
                                    double radius_exponent, double table_factor, double radius_max,
                                    const double *table_b, const double *table_c, double *volume,
                                    double *radius, double *velocity, int64_t *status, void *stream);

/* ---- terminal velocities: methods/terminal_velocity_methods.py ----------------------- */

/* interpolation, terminal_velocity_methods.py:13-30 (Gunn-Kinzer table lookup,
 * PySDM/dynamics/terminal_velocity/gunn_and_kinzer.py:108-133): b, c = tables f64[601] */
int pysdm_b200_interpolation(double *output, const double *radius, int64_t n, double factor,
                             const double *b, const double *c, void *stream);

/* terminal_velocity, :32-44 with PySDM/physics/terminal_velocity/rogers_yau.py:13-23 */
int pysdm_b200_terminal_velocity_rogers_yau(double *values, const double *radius, int64_t n,
                                            double small_k, double medium_k, double large_k,
                                            double small_r_limit, double medium_r_limit,
                                            void *stream);

/* power_series, :46-64; num_terms <= 8, coefficient arrays on the HOST */
int pysdm_b200_power_series(double *values, const double *radius, int64_t n, int64_t num_terms,
                            const double *host_prefactors, const double *host_powers,
                            void *stream);

/* linear_collection_efficiency, methods/collisions_methods.py:743-782.  host_params13 =
 * (A, B, D1, D2, E1, E2, F1, F2, G1, G2, G3, Mf, Mg); bit e of int_mask set = params[e] was
 * passed as an integer (Numba lowers `x ** int` as square-and-multiply, `x ** float` as pow:
 * Hydrodynamic passes Mf = 4 and Mg = 8 as ints, collision_kernels/hydrodynamic.py:11-13). */
int pysdm_b200_linear_collection_efficiency(const double *host_params13, int64_t int_mask,
                                            double *output, int64_t out_len,
                                            const double *radii, const uint8_t *is_first_in_pair,
                                            const int64_t *idx, int64_t length, double unit,
                                            void *stream);

/* ---- breakup fragmentation functions: methods/fragmentation_methods.py ------------------
 * Every entry point = the function body + fragmentation_limiters (:78-95) in one launch,
 * as the reference's public methods do.  has_nfmax = 0 encodes `nfmax is None`. */

/* slams_fragmentation, :106-133 */
int pysdm_b200_slams_fragmentation(double *n_fragment, double *frag_volume,
                                   const double *x_plus_y, double *probs, const double *rand,
                                   int64_t n, double vmin, int has_nfmax, double nfmax,
                                   void *stream);
/* exp_fragmentation, :136-170 */
int pysdm_b200_exp_fragmentation(double *n_fragment, double scale, double *frag_volume,
                                 const double *x_plus_y, const double *rand, int64_t n,
                                 double vmin, int has_nfmax, double nfmax, double tol,
                                 void *stream);
/* feingold1988_fragmentation, :172-198, 486-497; PySDM/physics/fragmentation_function/feingold1988.py */
int pysdm_b200_feingold1988_fragmentation(double *n_fragment, double scale, double *frag_volume,
                                          const double *x_plus_y, const double *rand, int64_t n,
                                          double fragtol, double vmin, int has_nfmax,
                                          double nfmax, void *stream);
/* gauss_fragmentation, :200-216, 474-483 */
int pysdm_b200_gauss_fragmentation(double *n_fragment, double mu, double sigma,
                                   double *frag_volume, const double *x_plus_y,
                                   const double *rand, int64_t n, double vmin, int has_nfmax,
                                   double nfmax, void *stream);
/* straub_fragmentation, :12-50, 218-260, 321-378; PySDM/physics/fragmentation_function/straub2010nf.py.
 * Nr1..Nr3 are read-modify-write (the caller zero-fills them), Nr4, Nrt, d34 are outputs. */
int pysdm_b200_straub_fragmentation(double *n_fragment, const double *CW, const double *gam,
                                    const double *ds, double *frag_volume, const double *v_max,
                                    const double *x_plus_y, const double *rand, double *Nr1,
                                    double *Nr2, double *Nr3, double *Nr4, double *Nrt,
                                    double *d34, int64_t n, double vmin, int has_nfmax,
                                    double nfmax, void *stream);
/* ll82_fragmentation, :52-74, 262-303, 380-472; PySDM/physics/fragmentation_function/lowlist82.py.
 * rand is rescaled in place, as in the reference. */
int pysdm_b200_ll82_fragmentation(double *n_fragment, const double *CKE, const double *W,
                                  const double *W2, const double *St, const double *ds,
                                  const double *dl, const double *dcoal, double *frag_volume,
                                  const double *x_plus_y, double *rand, double *Rf, double *Rs,
                                  double *Rd, int64_t n, double tol, double vmin, int has_nfmax,
                                  double nfmax, void *stream);
/* ll82_coalescence_check, :305-319 */
int pysdm_b200_ll82_coalescence_check(double *Ec, const double *dl, int64_t n, void *stream);

/* ---- collision_coalescence_breakup, methods/collisions_methods.py:62-311, 479-520 ------
 * handle_all_breakups selects break_up_while (:178-243) over break_up (:135-175), as
 * Formulae(handle_all_breakups=...) does.  n_overflow (device int64[1], may be NULL) is
 * incremented once per pair that raised the overflow flag; the host prints the reference's
 * warning when warn_overflows is set. */
int pysdm_b200_collision_coalescence_breakup(
    int64_t *multiplicity, const int64_t *idx, int64_t length, double *attributes,
    int64_t n_attr, int64_t n_sd, const double *gamma, const double *rand, const double *Ec,
    const double *Eb, const double *fragment_mass, int64_t *healthy, const int64_t *cell_id,
    int64_t *coalescence_rate, int64_t *breakup_rate, int64_t *breakup_rate_deficit,
    const uint8_t *is_first_in_pair, double max_multiplicity, const double *particle_mass,
    int handle_all_breakups, int64_t *n_overflow, void *stream);

/* ---- rows either side of the collision step -------------------------------- */

/* Storage operators used by Displacement.update_cell_origin / boundary_condition
 * (PySDM/dynamics/displacement.py:141-153) on mixed int64 / fp64 storages:
 * `int_storage.floor(float_storage)`, `float_storage -= int_storage`,
 * `int_storage %= int_divisors` (impl_numba/storage.py:54-100, storage_impl.py:11-83) */
int pysdm_b200_floor_to_i64(int64_t *out, const double *in, int64_t n, void *stream);
int pysdm_b200_sub_i64(double *inout, const int64_t *b, int64_t n, void *stream);
int pysdm_b200_row_modulo_i64(int64_t *out, const int64_t *divisor, int64_t n_rows,
                              int64_t n_cols, void *stream);

/* moments, methods/moments_methods.py:12-98.  moment_0 f64[n_cell], moments f64[n_ranks, n_cell]
 * (both zero-filled first, as the reference does); host_ranks = the `ranks` storage on the
 * HOST, n_ranks <= PYSDM_B200_MAX_MOMENT_RANKS; *_is_int != 0 = the exponent was passed as a
 * Python int (Numba then multiplies instead of calling pow).  Sums are fp64 atomics:
 * within 1e-12 relative of the reference's serial sums, not bit-identical. */
#define PYSDM_B200_MAX_MOMENT_RANKS 16
int pysdm_b200_moments(double *moment_0, double *moments, const int64_t *multiplicity,
                       const double *attr_data, const int64_t *cell_id, const int64_t *idx,
                       int64_t length, const double *host_ranks, int64_t n_ranks, double min_x,
                       double max_x, const double *x_attr, const double *weighting_attribute,
                       double weighting_rank, int weighting_rank_is_int, int skip_division_by_m0,
                       int64_t n_cell, void *stream);
/* spectrum_moments, methods/moments_methods.py:100-182.  moment_0, moments f64[n_bins, n_cell],
 * x_bins f64[n_bins + 1] on the device */
int pysdm_b200_spectrum_moments(double *moment_0, double *moments, const int64_t *multiplicity,
                                const double *attr_data, const int64_t *cell_id,
                                const int64_t *idx, int64_t length, double rank, int rank_is_int,
                                const double *x_bins, int64_t n_bins, const double *x_attr,
                                const double *weighting_attribute, double weighting_rank,
                                int weighting_rank_is_int, int64_t n_cell, void *stream);

/* calculate_displacement, methods/displacement_methods.py:14-153: one call per `dim`, as
 * Particulator.calculate_displacement issues them (particulator.py:428-440).  courant = the
 * component `dim` of the Courant field on the Arakawa-C grid, row-major, shape
 * host_courant_shape[n_dims] (HOST array); cell_origin i64[n_dims, n_sd], position_in_cell and
 * displacement f64[n_dims, n_sd]; implicit_in_space selects
 * PySDM/physics/particle_advection/implicit_in_space.py over explicit_in_space.py */
int pysdm_b200_calculate_displacement(int dim, int n_dims, int implicit_in_space,
                                      double *displacement, const double *courant,
                                      const int64_t *host_courant_shape,
                                      const int64_t *cell_origin, const double *position_in_cell,
                                      int64_t n_sd, double n_substeps, void *stream);
/* Displacement.__call__ without sedimentation (PySDM/dynamics/displacement.py:103-153) in one
 * launch: calculate_displacement for every dimension (from the positions before the update, as
 * the separate calls see them), update_position (`position_in_cell += displacement`),
 * update_cell_origin (`floor`, `cell_origin +=`, `position_in_cell -=`) and boundary_condition
 * (`cell_origin %= grid`, numpy's modulo) -- per super-droplet, same operations in the same
 * order, so bit-identical to the seven separate calls; `displacement` is written as well.
 * host_courant[n_dims]: HOST array of the device pointers of the Courant components;
 * host_courant_shapes[n_dims * n_dims]: HOST, row `dim` = the shape of component `dim`;
 * host_grid[n_dims]: HOST, the grid. */
int pysdm_b200_displacement_step(int n_dims, int implicit_in_space, double *displacement,
                                 const double *const *host_courant,
                                 const int64_t *host_courant_shapes, const int64_t *host_grid,
                                 int64_t *cell_origin, double *position_in_cell, int64_t n_sd,
                                 double n_substeps, void *stream);
/* flag_precipitated, :155-186, 208-238: the *_z pointers are the LAST rows of the [n_dims, n_sd]
 * arrays; flagged positions of idx get idx_len (= n_sd); rainfall_mass = device f64[1] */
int pysdm_b200_flag_precipitated(const int64_t *cell_origin_z, const double *position_in_cell_z,
                                 const double *water_mass, const int64_t *multiplicity,
                                 int64_t *idx, int64_t length, int64_t idx_len, int64_t *healthy,
                                 double precipitation_counting_level_index,
                                 const double *displacement_z, double *rainfall_mass,
                                 void *stream);
/* flag_out_of_column, :188-206, 240-260 */
int pysdm_b200_flag_out_of_column(const int64_t *cell_origin_z, const double *position_in_cell_z,
                                  int64_t *idx, int64_t length, int64_t idx_len,
                                  int64_t *healthy, double domain_top_level_index, void *stream);

/* ---- particle injection ------------------------------------------------------------------
 * seeding, backends/impl_numba/methods/seeding_methods.py:12-65 (caller particulator.py:447-490,
 * dynamic dynamics/seeding.py:91-117): in physical order, every slot with multiplicity 0 --
 * until n_inject slots have been filled -- gets idx[i] = -1 and the multiplicity and the n_attr
 * extensive attributes ([n_attr, n_sd] / [n_attr, n_seeds] row-major) of seed
 * seeded_particle_index[k], k = 0, 1, ... counting filled slots.  *host_n_injected = number
 * filled (the reference asserts it equals n_inject); the stream is synchronised. */
size_t pysdm_b200_seeding_workspace_bytes(int64_t n_sd);
int pysdm_b200_seeding(int64_t *idx, int64_t *multiplicity, double *extensive_attributes,
                       int64_t n_attr, int64_t n_sd, const int64_t *seeded_particle_index,
                       const int64_t *seeded_particle_multiplicity,
                       const double *seeded_particle_extensive_attributes, int64_t n_seeds,
                       int64_t n_inject, int64_t *host_n_injected, void *ws, size_t ws_bytes,
                       void *stream);

/* ---- migration between the slabs of a multi-GPU run (no reference counterpart: PySDM is
 * single-process; data model of PySDM/impl/particle_attributes.py) ---------------------------
 * The first grid dimension is split into `world` balanced contiguous slabs.  A record is
 * R = 1 + n_attr + 2 n_dims int64 words: multiplicity, attributes (fp64 bits), cell origin
 * (global coordinates), position in cell (fp64 bits). */

/* for every working position: dest[i] = owner rank of SD i = idx[p], cell_id[i] = cell id local
 * to the owner's slab; host_grid = grid size per dimension (HOST array) */
int pysdm_b200_migration_destination(int64_t *dest, int64_t *cell_id, const int64_t *cell_origin,
                                     const int64_t *idx, int64_t length, int64_t n_sd, int n_dims,
                                     const int64_t *host_grid, int64_t world, void *stream);
/* records[r, :] <- SD order[q], q = r skipping [own_begin, own_begin + own_count) (`order` =
 * idx sorted by destination rank); the multiplicity of a packed SD is zeroed (slot vacant) */
int pysdm_b200_migration_pack(int64_t *records, int64_t n_rows, const int64_t *order,
                              int64_t own_begin, int64_t own_count, int64_t *multiplicity,
                              const double *attributes, int64_t n_attr,
                              const int64_t *cell_origin, const double *position_in_cell,
                              int n_dims, int64_t n_sd, void *stream);
/* SD slots[r] <- records[r, :]; idx_tail[r] = slots[r] (idx_tail = idx + working length) */
int pysdm_b200_migration_unpack(const int64_t *records, int64_t n_rows, const int64_t *slots,
                                int64_t *idx_tail, int64_t *multiplicity, double *attributes,
                                int64_t n_attr, int64_t *cell_origin, double *position_in_cell,
                                int n_dims, int64_t n_sd, void *stream);

/* ---- migration to the two ring-neighbour slabs without host round trips --------------------
 * (pysdm_b200/migration.py; csrc/migration.cu).  counters (i64[8], device): [0] stayers,
 * [1] free slots (in/out), [2] new working length (out), [3] status (0 ok; 1 a droplet moved
 * beyond a neighbour slab: nothing was modified, use the general path; 2 capacity exceeded),
 * [4]/[5] leavers to rank+1 / rank-1, [6]/[7] arrivals.  send / recv: [2, cap + 1, R] records,
 * row 0 of each buffer = header (word 0 = count). */
size_t pysdm_b200_migration_split_workspace_bytes(int64_t length);
int pysdm_b200_migration_split(const int64_t *idx, int64_t length, int64_t *idx_out, int64_t *dest,
                               int64_t *cell_id, const int64_t *cell_origin,
                               const double *position_in_cell, int64_t *multiplicity,
                               const double *attributes, int64_t n_attr, int64_t n_sd, int n_dims,
                               const int64_t *host_grid, int64_t rank, int64_t world, int64_t *send,
                               int64_t *leavers, int64_t cap, int64_t *free_stack,
                               int64_t *counters, void *ws, size_t ws_bytes, void *stream);
int pysdm_b200_migration_merge(const int64_t *recv, int64_t cap, int64_t *idx_out,
                               int64_t *multiplicity, double *attributes, int64_t n_attr,
                               int64_t *cell_origin, double *position_in_cell, int64_t *cell_id,
                               int64_t n_sd, int n_dims, const int64_t *host_grid, int64_t rank,
                               int64_t world, const int64_t *free_stack, int64_t *counters,
                               void *stream);

/* PYSDM_B200_H_END_OF_METHOD_LEVEL -- fused step and later additions follow */

#ifdef __cplusplus
}
#endif
#endif /* PYSDM_B200_H */
