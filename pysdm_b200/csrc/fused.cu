// fused.cu -- C-ABI of the one-launch SDM collision step (kernel template: fused_kernel.cuh,
// instantiations: fused_k*.cu): argument checks, the per-rank / per-slot pre-pass kernels,
// batch geometry, dispatch on (collision kernel, uniform source, group size, adaptive).
#include "fused_kernel.cuh"

#include <stdlib.h>

// ---- pre-pass: norm_factor per cell rank (normalize, collisions_methods.py:639-646) ---------
// adaptive launches also get, per rank r: the cell id it holds and the parity of
// cell_start[cell_idx[cell id]] -- the position find_pairs (pair_methods.py:40) counts "even"
// from.  cell_idx is re-sorted by dt_left before every sub-step while idx keeps the order of the
// last sort by cell (quirk A2), so this changes from sub-step to sub-step.
__global__ void k_fused_norm_factor(const i64 *__restrict__ cell_start, i64 n_cell,
                                    double *__restrict__ norm_factor, double timestep, double dv,
                                    bool refresh_nf, const i64 *__restrict__ rank_to_cell,
                                    const i64 *__restrict__ cell_idx, i64 *__restrict__ rank_info,
                                    i64 *__restrict__ adaptive_end)
{
    i64 r = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (r == 0 && adaptive_end != nullptr) adaptive_end[0] = 0;
    if (r >= n_cell) return;
    if (refresh_nf)
        norm_factor[r] = sdm_norm_factor(cell_start[r + 1] - cell_start[r], timestep, dv);
    if (rank_info != nullptr) {
        const i64 cid = rank_to_cell[r];
        rank_info[r] = cid | ((cell_start[cell_idx[cid]] & 1) << 62);
    }
}

int pysdm_fused_prepass(const i64 *cell_start, i64 n_cell, double *norm_factor, double timestep,
                        double dv, const i64 *rank_to_cell, const i64 *cell_idx, i64 *rank_info,
                        i64 *adaptive_end, cudaStream_t st)
{
    k_fused_norm_factor<<<grid_for(n_cell), kThreads, 0, st>>>(
        cell_start, n_cell, norm_factor, timestep, dv, true, rank_to_cell, cell_idx, rank_info,
        adaptive_end);
    return LAUNCH_CHECK();
}

// per pair slot d: the factor normalize() multiplies prob[d] with (collisions_methods.py:647-648:
// norm_factor[cell_idx[cell_id[d]]], d used as a PARTICLE number -- reference quirk A1).  It only
// changes when cell_start / cell_idx / cell_id / dt / dv change, so it is cached across steps.
__global__ void k_fused_nf_slot(const i64 *__restrict__ cell_id, const i64 *__restrict__ cell_idx,
                                const double *__restrict__ norm_factor, double *__restrict__ nf_slot,
                                i64 n_slots)
{
    i64 d = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (d < n_slots) nf_slot[d] = norm_factor[cell_idx[cell_id[d]]];
}

// rank_to_cell[cell_idx[c]] = c: the inverse of the cell -> rank map a sort by cell used
__global__ void k_invert_index(i64 *__restrict__ out, const i64 *__restrict__ in, i64 n)
{
    i64 c = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (c < n) {
        const i64 r = in[c];
        if (r >= 0 && r < n) out[r] = c;
    }
}

PYSDM_EXPORT int pysdm_b200_invert_index(i64 *out, const i64 *in, i64 n, void *stream)
{
    if (n <= 0) return 0;
    k_invert_index<<<grid_for(n), kThreads, 0, (cudaStream_t)stream>>>(out, in, n);
    return LAUNCH_CHECK();
}

__global__ void k_dependent_attributes(const double *__restrict__ mass, i64 n, DependentParams p,
                                       double *__restrict__ volume, double *__restrict__ radius,
                                       double *__restrict__ velocity, i64 *__restrict__ status)
{
    i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    double vol, rad, vel;
    const bool ok = dependents_of(mass[i], p, vol, rad, vel);
    if (volume != nullptr) volume[i] = vol;
    if (radius != nullptr) radius[i] = rad;
    if (!ok && status != nullptr) status[0] = 1;
    if (ok && velocity != nullptr) velocity[i] = vel;
}

PYSDM_EXPORT int pysdm_b200_dependent_attributes(const double *mass, i64 n, double rho_w,
                                                 double radius_scale, double radius_exponent,
                                                 double table_factor, double radius_max,
                                                 const double *table_b, const double *table_c,
                                                 double *volume, double *radius, double *velocity,
                                                 i64 *status, void *stream)
{
    if (n <= 0) return 0;
    if (velocity != nullptr && (table_b == nullptr || table_c == nullptr))
        return (int)cudaErrorInvalidValue;
    DependentParams p;
    p.rho_w = rho_w;
    p.radius_scale = radius_scale;
    p.radius_exponent = radius_exponent;
    p.table_factor = table_factor;
    p.radius_max = radius_max;
    p.table_b = table_b;
    p.table_c = table_c;
    k_dependent_attributes<<<grid_for(n), kThreads, 0, (cudaStream_t)stream>>>(
        mass, n, p, volume, radius, velocity, status);
    return LAUNCH_CHECK();
}

PYSDM_EXPORT int64_t pysdm_b200_fused_max_cell_size(void) { return kCtaPos; }

// stream.cu: the step for cells beyond kCtaPos (boxes)
int pysdm_stream_collision_step(const FusedArgs &a, cudaStream_t st);
size_t pysdm_stream_ws_bytes(i64 n_sd, i64 n_cell);
PYSDM_EXPORT size_t pysdm_b200_fused_workspace_bytes(int64_t n_sd, int64_t n_cell)
{
    return pysdm_stream_ws_bytes(n_sd, n_cell);
}

// precondition of div_by(): b normal, moderate, significand not all ones
static bool divisor_ok(double b)
{
    u64 bits;
    memcpy(&bits, &b, sizeof bits);
    const u64 mant = bits & 0x000FFFFFFFFFFFFFull;
    const double m = b < 0 ? -b : b;
    return m > 1e-10 && m < 1e10 && mant != 0x000FFFFFFFFFFFFFull;
}

// CTAs of k_fused_step that are resident at once on the current device
static i64 resident_ctas()
{
    static int sm_count[64];
    int device = 0;
    if (cudaGetDevice(&device) != cudaSuccess || device < 0 || device >= 64) return 148 * PYSDM_MIN_CTAS;
    if (sm_count[device] == 0) {
        int n = 0;
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, device) != cudaSuccess || n <= 0)
            n = 148;
        sm_count[device] = n;
    }
    return (i64)sm_count[device] * PYSDM_MIN_CTAS;
}

// Cells per batch (= per warp, or per CTA for cells beyond kWarpPos super-droplets).  The upper
// bound is what fits the group's shared memory.  Below ~10 waves of CTAs the choice also decides
// how full the LAST wave is: 2^22 super-droplets in cells of 128 are 512 CTAs of 8 cells per
// warp on 444 resident slots -- two waves, the second one 15 % full.  So the cost model
// waves(c) * (c + 1) -- a batch costs its cells plus about one cell of fixed work -- is minimised
// over the admissible c; results do not depend on the choice (cells are independent, uniforms
// are addressed by global position).
static i64 cells_per_batch(i64 max_cell_size, i64 n_cell, bool &warp)
{
    const i64 mcs = max_cell_size < 1 ? 1 : max_cell_size;
    warp = mcs <= kWarpPos;
    i64 cpb = (warp ? kWarpPos : kCtaPos) / mcs;
    const i64 cap = warp ? kWarpCells : kCtaCells;
    cpb = cpb > cap ? cap : cpb;
    static const char *forced = getenv("PYSDM_B200_CELLS_PER_BATCH");  // tuning experiments only
    if (forced != nullptr && atoll(forced) >= 1) return atoll(forced) < cpb ? atoll(forced) : cpb;
    const i64 groups = warp ? kFusedThreads / 32 : 1, slots = resident_ctas();
    i64 best = cpb, best_cost = -1;
    for (i64 c = cpb; c >= 1 && c >= cpb / 4; --c) {
        const i64 n_ctas = ((n_cell + c - 1) / c + groups - 1) / groups;
        const i64 waves = (n_ctas + slots - 1) / slots;
        const i64 cost = waves * (c + 1);
        if (best_cost < 0 || cost < best_cost) {
            best = c;
            best_cost = cost;
        }
        if (waves > 12) break;  // many waves: the tail no longer matters
    }
    return best;
}

PYSDM_EXPORT int pysdm_b200_fused_collision_step(const pysdm_b200_fused_args *host_args,
                                                 void *stream)
{
    FusedArgs a = *host_args;
    if (a.n_sd <= 0 || a.n_sd >= ((i64)1 << 32) || a.n_cell <= 0) return (int)cudaErrorInvalidValue;
    const bool big = a.max_cell_size > kCtaPos;  // -> stream.cu (needs scratch)
    if (big && a.ws == nullptr) return (int)cudaErrorInvalidConfiguration;
    if ((a.u01 == nullptr) != (a.rand == nullptr)) return (int)cudaErrorInvalidValue;
    cudaStream_t st = (cudaStream_t)stream;
    if (a.nf_slot == nullptr) return (int)cudaErrorInvalidValue;
    const bool adaptive = a.adaptive != 0;
    if (adaptive) {
        if (big) a.rank_info = nullptr;  // (cell ids come from cell_id[] there)
        if (!a.dt_left || !a.stats_n_substep || !a.stats_dt_min ||
            (!big && (!a.rank_to_cell || !a.rank_info)) || !(a.dt_min >= 0.0) ||
            !(a.timestep > 0.0))
            return (int)cudaErrorInvalidValue;
        a.substeps = 1.0;
        a.refresh_nf = 1;  // cell_idx was re-sorted since the previous sub-step
    } else {
        a.rank_info = nullptr;
        a.adaptive_end = nullptr;
        a.length_dev = nullptr;
    }
    // (prepass_done: pysdm_b200_adaptive_presort has left norm_factor, rank_info and the zeroed
    // adaptive_end behind, in the launch that sorted cell_idx)
    if ((a.refresh_nf || adaptive) && !(adaptive && !big && a.prepass_done)) {
        k_fused_norm_factor<<<grid_for(a.n_cell), kThreads, 0, st>>>(
            a.cell_start, a.n_cell, a.norm_factor, a.timestep, a.dv, a.refresh_nf != 0,
            a.rank_to_cell, a.cell_idx, a.rank_info, a.adaptive_end);
        // (adaptive sub-steps of the batched kernel follow the chain themselves: cell_idx changes
        // with every sub-step, so a per-slot cache would be rebuilt as often as it is read)
        if (a.n_sd / 2 > 0 && (!adaptive || big))
            k_fused_nf_slot<<<grid_for(a.n_sd / 2), kThreads, 0, st>>>(
                a.cell_id, a.cell_idx, a.norm_factor, a.nf_slot, a.n_sd / 2);
    }
    bool warp;
    const i64 cpb = cells_per_batch(a.max_cell_size, a.n_cell, warp);
    a.cells_per_batch = cpb;
    a.reserved = divisor_ok(a.rho_w) ? 1 : 0;
    for (int r = 0; r < 10; ++r) {  // Philox4x32 key schedule (philox.cuh)
        a.philox_round_keys[2 * r] = (u32)a.seed + (u32)r * 0x9E3779B9u;
        a.philox_round_keys[2 * r + 1] = (u32)(a.seed >> 32) + (u32)r * 0xBB67AE85u;
    }
    if (big) {
        if ((a.kernel == KERNEL_GEOMETRIC || a.kernel == KERNEL_HYDRODYNAMIC) &&
            (a.radius == nullptr || a.fall_velocity == nullptr))
            return (int)cudaErrorInvalidValue;
        return pysdm_stream_collision_step(a, st);
    }
    const i64 n_batches = (a.n_cell + cpb - 1) / cpb;
    const bool philox = a.u01 == nullptr;
    const bool aggregate = a.n_cell < 64;  // few counters: aggregate within the warp first
    switch (a.kernel) {
    case KERNEL_GOLOVIN:
        return dispatch_fused<KERNEL_GOLOVIN>(a, n_batches, philox, warp, aggregate, adaptive, st);
    case KERNEL_CONSTANT_K:
        return dispatch_fused<KERNEL_CONSTANT_K>(a, n_batches, philox, warp, aggregate, adaptive, st);
    case KERNEL_LINEAR:
        return dispatch_fused<KERNEL_LINEAR>(a, n_batches, philox, warp, aggregate, adaptive, st);
    case KERNEL_GEOMETRIC:
        if (a.radius == nullptr || a.fall_velocity == nullptr) return (int)cudaErrorInvalidValue;
        if (a.dep_radius != nullptr && a.dep_velocity != nullptr &&
            (a.dep_table_b == nullptr || a.dep_table_c == nullptr))
            return (int)cudaErrorInvalidValue;
        return dispatch_fused<KERNEL_GEOMETRIC>(a, n_batches, philox, warp, aggregate, adaptive, st);
    case KERNEL_HYDRODYNAMIC:
        if (a.radius == nullptr || a.fall_velocity == nullptr) return (int)cudaErrorInvalidValue;
        if (a.dep_radius != nullptr && a.dep_velocity != nullptr &&
            (a.dep_table_b == nullptr || a.dep_table_c == nullptr))
            return (int)cudaErrorInvalidValue;
        return dispatch_fused<KERNEL_HYDRODYNAMIC>(a, n_batches, philox, warp, aggregate, adaptive, st);
    default: return (int)cudaErrorInvalidValue;
    }
}

PYSDM_EXPORT int pysdm_b200_shuffle_local_batched(i64 *idx, const double *u01, const i64 *cell_start,
                                                  i64 n_cell, i64 n_sd, i64 max_cell_size,
                                                  i64 *status, void *stream)
{
    if (n_sd <= 0 || n_sd >= ((i64)1 << 32) || n_cell <= 0) return (int)cudaErrorInvalidValue;
    if (max_cell_size > kCtaPos) return (int)cudaErrorInvalidConfiguration;
    FusedArgs a = {};
    a.idx_in = idx;
    a.idx_out = idx;  // in place: a batch stages its whole range before it writes
    a.u01 = u01;
    a.cell_start = cell_start;
    a.status = status;
    a.n_sd = n_sd;
    a.n_cell = n_cell;
    a.substeps = 1.0;
    a.shuffle_only = 1;
    bool warp;
    a.cells_per_batch = cells_per_batch(max_cell_size, n_cell, warp);
    const i64 n_batches = (n_cell + a.cells_per_batch - 1) / a.cells_per_batch;
    return dispatch_fused<KERNEL_CONSTANT_K>(a, n_batches, false, warp, false, false,
                                             (cudaStream_t)stream);
}
