// fused_kernel.cuh -- one-launch SDM collision step for cell-sorted populations (sm_100a).
// The kernel template; instantiated per collision kernel in fused_k*.cu (one translation unit
// each, compiled in parallel), dispatched and exported by fused.cu.
//
// Executes, for every cell of a population whose `idx` is sorted by cell
// (ParticleAttributes.cell_start valid), literally the same sequence the reference runs
// as ~15 separate backend calls in Collision.step (PySDM/dynamics/collisions/collision.py:196-234):
//
//   shuffle_local            impl_numba/methods/index_methods.py:32-44
//   find_pairs               impl_numba/methods/pair_methods.py:34-55
//   sort_within_pair_by_attr pair_methods.py:126-140          (attr = multiplicity)
//   collision kernel         collision_kernels/{golovin,constantK,linear}.py  (sum_pair, *=, +=),
//                            geometric.py:15-22 (sum_pair of radius, **= 2, *=, distance_pair of
//                            the fall velocity, *=) on the per-SD radius / velocity attributes
//   max_pair, *= kernel_temp pair_methods.py:57-75, collision.py:249-254
//   normalize                impl_numba/methods/collisions_methods.py:633-662 (incl. quirk A1)
//   prob /= substeps         collision.py:279       (non-adaptive), or
//   scale_prob_for_adaptive_sdm_gamma  collisions_methods.py:330-405  (kAdaptive: one launch = one
//                            sub-step of the loop collision.py:178-190; per-cell dt reduction in
//                            shared memory, no global atomics)
//   compute_gamma            collisions_methods.py:522-585
//   collision_coalescence    collisions_methods.py:418-477
//
// and writes every API-visible result: idx, is_first_in_pair, gamma, multiplicity,
// attributes, norm_factor, the per-cell rate counters and the `healthy` flag -- bit
// for bit what the method-level kernels (and the reference) produce.
//
// Data movement (DESIGN.md "fused step"): a GROUP of threads -- one warp when every cell
// holds <= kWarpPos super-droplets, else a whole CTA -- owns a batch of consecutive cells
// whose SDs occupy one contiguous range of positions.  The range of `idx` and of the
// uniforms is streamed in with 16-byte loads and staged in shared memory as 32-bit SD
// numbers + 16-bit swap targets; one lane per cell replays the cell's swap chain in shared
// memory; then each lane handles two pair slots: gathers multiplicity / mass (the only
// random accesses, confined to the cell's own SDs), evaluates probability and gamma and
// scatters the updates.  Nothing is written to HBM between the stages, warps of the
// warp-group variant never wait for each other.
//
// Instruction diet (ncu, profiles/): the step is bound by instruction issue and by the
// conversion/transcendental pipe (XU) long before HBM unless 64-bit int<->fp conversions,
// fp64 divisions and ceil() are kept out of the per-SD path; see the helpers below --
// each is exact (bit-identical to the plain expression), with a guarded slow path.
#pragma once
#include "berry.cuh"
#include "dependents.cuh"
#include "philox.cuh"
#include "pysdm_b200.h"
#include "sdm_device.cuh"

// C-ABI argument block: include/pysdm_b200.h (mirrored by ctypes in pysdm_b200/_lib.py)
typedef pysdm_b200_fused_args FusedArgs;

enum { KERNEL_GOLOVIN = 0, KERNEL_CONSTANT_K = 1, KERNEL_LINEAR = 2, KERNEL_GEOMETRIC = 3,
       KERNEL_HYDRODYNAMIC = 4 };

#ifndef PYSDM_WARP_POS
#define PYSDM_WARP_POS 1024
#endif
#ifndef PYSDM_KWIN
#define PYSDM_KWIN 4  // phase 2: windows of idx loads in flight per thread
#endif
#ifndef PYSDM_MIN_CTAS
#define PYSDM_MIN_CTAS 3
#endif
constexpr int kFusedThreads = 256;
constexpr int kCtaPos = 8192, kCtaCells = 256;  // CTA group: positions / cells per batch
constexpr int kWarpPos = PYSDM_WARP_POS, kWarpCells = 32;  // warp group
static_assert(kWarpCells <= 32, "cell number and pair role share one byte (5 + 2 bits)");
constexpr double kTwo52 = 4503599627370496.0;    // 2^52
constexpr i64 kCidMask = ((i64)1 << 61) - 1;     // rank_info / cidv (adaptive): cell id | idle << 61 | parity << 62

// ---- exact arithmetic helpers without XU-pipe instructions ---------------------------------
// (double)n for 0 <= n < 2^52 by mantissa insertion; other values take the conversion
__device__ __forceinline__ double i64_to_f64(i64 n)
{
    if ((u64)n < ((u64)1 << 52))
        return __longlong_as_double((long long)(0x4330000000000000ull | (u64)n)) - kTwo52;
    return (double)n;
}
// (i64)x (truncation) for integer-valued or fractional 0 <= x < 2^51; else the conversion
__device__ __forceinline__ i64 f64_to_i64(double x)
{
    if (x >= 0.0 && x < 2251799813685248.0) {
        double t = __dadd_rd(x, kTwo52);  // 2^52 + floor(x), exact
        return (i64)((u64)__double_as_longlong(t) & 0x000FFFFFFFFFFFFFull);
    }
    return (i64)x;
}
// ceil(x) with the sign of zero of the IEEE result, for |x| < 2^51; else ceil()
__device__ __forceinline__ double ceil_exact(double x)
{
    if (fabs(x) < 2251799813685248.0) {
        const double c = 6755399441055744.0;  // 1.5 * 2^52: ulp = 1 around it
        double r = __dadd_ru(x, c) - c;
        if (r == 0.0) r = copysign(0.0, x);
        return r;
    }
    return ceil(x);
}
// a / b given r = RN(1/b): Markstein's correction q' = fma(fma(-q, b, a), r, q) is the
// correctly rounded quotient when q = RN(a*r) is faithful and nothing under/overflows
// (`ok` = b is finite, normal, its significand not all ones; checked once on the host).
__device__ __forceinline__ double div_by(double a, double b, double r, bool ok)
{
    const double m = fabs(a);
    if (ok && m > 1e-280 && m < 1e280) {
        const double q = a * r;
        return fma(fma(-q, b, a), r, q);
    }
    return a / b;
}

// kernel_temp of a pair from the pair's volume sum (sum_pair result; 0 for non-pairs)
template <int K>
__device__ __forceinline__ double kernel_temp_of(double vsum, double a, double b)
{
    if (K == KERNEL_GOLOVIN) return vsum * b;   // golovin.py:14-16
    if (K == KERNEL_CONSTANT_K) return a;       // constantK.py: output.fill(a)
    return vsum * b + a;                        // linear.py: sum; *= b; += a  (no FMA: -fmad=false)
}

// Geometric (collision_kernels/geometric.py:15-22): output = r_j + r_k (sum_pair); output **= 2
// (Storage.__ipow__ = sign(x) |x|^e, impl_numba/storage_impl.py power; csrc/storage_ops.cu
// signed_pow); output *= PI * collection_efficiency (= a, multiplied on the host as the
// reference does); pair_tmp = |u_j - u_k| (distance_pair); output *= pair_tmp
__device__ __forceinline__ double geometric_kernel_temp(double r_j, double r_k, double u_j,
                                                        double u_k, double a)
{
    const double x = r_j + r_k;
    const double sgn = x > 0 ? 1.0 : (x < 0 ? -1.0 : (x == 0 ? 0.0 : x));
    double out = sgn * (fabs(x) * fabs(x));
    out = out * a;
    return out * fabs(u_j - u_k);
}

__device__ __forceinline__ double signed_square(double x)  // Storage `**= 2`
{
    const double sgn = x > 0 ? 1.0 : (x < 0 ? -1.0 : (x == 0 ? 0.0 : x));
    return sgn * (fabs(x) * fabs(x));
}

// Parameterized / Hydrodynamic (collision_kernels/impl/parameterized.py:13-30):
// output = linear_collection_efficiency(radius pair); output **= 2; output *= PI (= a);
// pair_tmp = max_pair(radius); pair_tmp **= 2; output *= pair_tmp;
// pair_tmp = distance_pair(fall velocity); output *= pair_tmp
__device__ __forceinline__ double hydrodynamic_kernel_temp(const FusedArgs &a, double r_j,
                                                           double r_k, double u_j, double u_k)
{
    Berry q;
#pragma unroll
    for (int e = 0; e < 13; ++e) q.p[e] = a.berry_params[e];
    q.int_mask = a.berry_int_mask;
    double out = signed_square(berry_efficiency(q, r_j, r_k, a.berry_unit));
    out = out * a.kernel_a;
    out = out * signed_square(py_max(r_j, r_k));  // max(in[idx[i]], in[idx[i+1]]), pair_methods.py:57-75
    return out * fabs(u_j - u_k);
}

// ---- group geometry -----------------------------------------------------------------------
template <bool kWarp>
struct Geo {
    static constexpr int kPos = kWarp ? kWarpPos : kCtaPos;        // positions per batch
    static constexpr int kCells = kWarp ? kWarpCells : kCtaCells;  // cells per batch
    static constexpr int kSize = kWarp ? 32 : kFusedThreads;       // threads per group
    static constexpr int kGroups = kFusedThreads / kSize;          // groups per CTA
    // shared-memory carve-up per group (bytes, 16-byte aligned sections)
    // sd / jl are indexed through phys(): 2 pad elements per 64 positions, so that the chain
    // lanes of equally sized cells (stride 128, 64, ...) fall into different banks
    static constexpr int kPhys = kPos + 4 + 2 * ((kPos + 4) / 64 + 1);
    static constexpr int kSd = (kPhys * 4 + 15) / 16 * 16;         // u32 SD numbers
    static constexpr int kJl = (kPhys * 2 + 15) / 16 * 16;         // u16 swap targets
    static constexpr int kCo = ((kPos + 4) * 1 + 15) / 16 * 16;    // u8 position -> cell
    static constexpr int kCs = ((kCells + 1) * 4 + 15) / 16 * 16;  // i32 local cell starts
    static constexpr int kFd = kCells * 16;                        // double S, len per cell
    static constexpr int kCid = kCells * 8;                        // i64 cell id per cell
    static constexpr int kQueue = (kSize / 32) * 64 * 12;          // per warp: 64 x (f64 + u32)
    static constexpr int kBytes = kSd + kJl + kCo + kCs + kFd + kCid + kQueue;
    static constexpr int kSmem = kBytes * kGroups;
};

// skewed shared-memory index of local position lp (keeps even positions 8-byte aligned)
__device__ __forceinline__ int phys(int lp) { return lp + 2 * (lp >> 6); }

template <bool kWarp>
__device__ __forceinline__ void group_sync()
{
    if (kWarp)
        __syncwarp();
    else
        __syncthreads();
}

// per-cell int64 counter: per-lane RED when cells are many (no contention), warp-aggregated
// when few (boxes); the aggregated form needs the whole warp converged at the call
template <bool kAggregate>
__device__ __forceinline__ void counter_add(i64 *counters, i64 cid, i64 value, bool active)
{
    if (kAggregate)
        cell_counter_add(counters, cid, value, active);
    else if (active && value != 0)
        atomicAdd((unsigned long long *)&counters[cid], (unsigned long long)value);
}

// ---- colliding pairs, one per lane: gamma cap, rate counters, coalescence ---------------------
// (compute_gamma :546-558 and collision_coalescence/coalesce :44-59,:418-477 of
// impl_numba/methods/collisions_methods.py).  Called with the whole warp converged; `mine`
// tells whether this lane holds an entry: pf = local position of the pair's first member,
// g_raw = ceil(prob - rand) != 0.  The pair is re-read from shared memory and re-sorted by
// multiplicity (cheap, cached) instead of being carried through the queue.
// ---- one colliding pair: gamma cap, coalescence, derived attributes ----------------------------
// (compute_gamma :546-558 and collision_coalescence/coalesce :44-59,:418-477 of
// impl_numba/methods/collisions_methods.py).  (sj, sk) sorted by multiplicity (n_j >= n_k),
// g_raw = ceil(prob - rand) != 0; *gamma_slot is rewritten when the cap n_j // n_k applies.
// Returns the three counter increments.
struct HitResult {
    i64 add_rate, add_deficit, add_coal;
};
template <int K>
__device__ __forceinline__ HitResult hit_pair(const FusedArgs &a, u32 sj, u32 sk, i64 n_j, i64 n_k,
                                              double g_raw, double *gamma_slot)
{
    HitResult r = {0, 0, 0};
    // g = min(int(gamma), n_j // n_k); n_j >= n_k > 0 here, so int(gamma) <= 1 needs no division
    const i64 gi = f64_to_i64(g_raw);
    i64 g = gi;
    if (gi > 1 || n_k <= 0) {
        const i64 prop = floordiv(n_j, n_k);
        g = gi < prop ? gi : prop;
    }
    const double gamma = i64_to_f64(g);
    if (g != gi) *gamma_slot = gamma;
    r.add_rate = g * n_k;
    r.add_deficit = (gi - g) * n_k;
    if (gamma != 0.0 && a.skip_coalescence == 0) {
        // coalesce: float arithmetic stored to int64 (quirk A7), attribute update fused (fma)
        const double f_nj = i64_to_f64(n_j), f_nk = i64_to_f64(n_k);
        const double gnk = gamma * f_nk;
        r.add_coal = f64_to_i64(gnk);
        const double new_n = f_nj - gnk;
        i64 nj_new, nk_new = n_k;
        const bool keep = new_n > 0;
        if (keep) {
            nj_new = f64_to_i64(new_n);
            a.multiplicity[sj] = nj_new;
        } else {
            nj_new = n_k >> 1;  // n_k // 2 (floor, also for negatives)
            nk_new = n_k - nj_new;
            a.multiplicity[sj] = nj_new;
            a.multiplicity[sk] = nk_new;
        }
        double new_mass = 0.0;
        for (i64 r_ = 0; r_ < a.n_attr; ++r_) {
            double *row = a.attributes + r_ * a.n_sd;
            const double v = fma(gamma, row[sj], row[sk]);
            row[sk] = v;
            if (!keep) row[sj] = v;
            if (r_ == a.mass_row) new_mass = v;
        }
        if (a.dep_radius == nullptr && a.dep_volume != nullptr) {
            // only the "volume" attribute is kept current (Golovin / Linear read it)
            const double vol = new_mass / a.rho_w;
            a.dep_volume[sk] = vol;
            if (!keep) a.dep_volume[sj] = vol;
        }
        if ((K == KERNEL_GEOMETRIC || K == KERNEL_HYDRODYNAMIC) && a.dep_radius != nullptr) {
            // the attributes derived from the water mass, for the changed super-droplets only
            DependentParams dp;
            dp.rho_w = a.rho_w;
            dp.radius_scale = a.dep_radius_scale;
            dp.radius_exponent = a.dep_radius_exponent;
            dp.table_factor = a.dep_table_factor;
            dp.radius_max = a.dep_radius_max;
            dp.table_b = a.dep_table_b;
            dp.table_c = a.dep_table_c;
            double vol, rad, vel;
            const bool ok = dependents_of(new_mass, dp, vol, rad, vel);
            if (!ok && a.dep_status != nullptr) a.dep_status[0] = 1;
            if (a.dep_volume != nullptr) a.dep_volume[sk] = vol;
            a.dep_radius[sk] = rad;
            if (ok && a.dep_velocity != nullptr) a.dep_velocity[sk] = vel;
            if (!keep) {
                if (a.dep_volume != nullptr) a.dep_volume[sj] = vol;
                a.dep_radius[sj] = rad;
                if (ok && a.dep_velocity != nullptr) a.dep_velocity[sj] = vel;
            }
        }
        if (nj_new == 0 || nk_new == 0) {
            a.healthy[0] = 0;
            if (a.gate) a.gate[0] = a.epoch;
        }
    }
    return r;
}

template <int K, bool kAggregate, bool kAdaptive>
__device__ __forceinline__ void process_hits(const FusedArgs &a, u32 pf, double g_raw, bool mine,
                                             const u32 *sd, const u8 *cellof, u32 cell_mask,
                                             const i64 *cidv, double *gamma_b)
{
    bool active = false;
    i64 cid = 0;
    HitResult r = {0, 0, 0};
    if (mine) {
        u32 sj = sd[phys((int)pf)], sk = sd[phys((int)pf + 1)];
        i64 n_j = a.multiplicity[sj], n_k = a.multiplicity[sk];
        if (n_j < n_k) {  // sort_within_pair_by_attr (ties keep order)
            const u32 t = sj; sj = sk; sk = t;
            const i64 tn = n_j; n_j = n_k; n_k = tn;
        }
        cid = cidv[cellof[pf] & cell_mask];
        if (kAdaptive) cid &= kCidMask;
        active = true;
        r = hit_pair<K>(a, sj, sk, n_j, n_k, g_raw, gamma_b + (pf >> 1));
    }
    counter_add<kAggregate>(a.collision_rate, cid, r.add_rate, active);
    counter_add<kAggregate>(a.collision_rate_deficit, cid, r.add_deficit, active);
    counter_add<kAggregate>(a.coalescence_rate, cid, r.add_coal, active);
}

// ---- adaptive time step helpers ------------------------------------------------------------

// n_j // n_k for 0 < n_k <= n_j (pairs are sorted by multiplicity): through one fp64 division
// when both are below 2^52 (the correctly rounded quotient of two such integers truncates to
// the exact floor: it could only round up to the next integer if n_j + 1 > 2^53)
__device__ __forceinline__ i64 floordiv_sorted(i64 n_j, i64 n_k)
{
    if (n_j == n_k) return 1;
    if (n_k > 0 && n_j < ((i64)1 << 52)) return f64_to_i64(i64_to_f64(n_j) / i64_to_f64(n_k));
    return floordiv(n_j, n_k);
}

// per-cell minimum of the lanes' values (positive doubles: their bit patterns order like the
// values) into fd[2 * cell + 1]; lanes of one warp mostly share a cell, so: one pair of integer
// warp reductions per distinct cell among the lanes, one shared-memory update by its leader
template <bool kWarp>
__device__ __forceinline__ void warp_cell_min(double v, int cell, bool valid, double *fd, int lane)
{
    unsigned rem = __ballot_sync(0xffffffffu, valid);
    while (rem) {
        const int leader = __ffs(rem) - 1;
        const int c = __shfl_sync(0xffffffffu, cell, leader);
        const bool mine = valid && cell == c;
        const unsigned grp = __ballot_sync(0xffffffffu, mine);
        const u32 hi = mine ? (u32)__double2hiint(v) : 0x7ff00000u;
        const u32 mhi = __reduce_min_sync(0xffffffffu, hi);
        const u32 lo = (mine && hi == mhi) ? (u32)__double2loint(v) : 0xffffffffu;
        const u32 mlo = __reduce_min_sync(0xffffffffu, lo);
        if (lane == leader) {
            if (kWarp) {  // the warp is the only writer of its batch's cells
                const double m = __hiloint2double((int)mhi, (int)mlo);
                if (m < fd[2 * c + 1]) fd[2 * c + 1] = m;
            } else {
                atomicMin(reinterpret_cast<unsigned long long *>(fd + 2 * c + 1),
                          ((unsigned long long)mhi << 32) | mlo);
            }
        }
        rem &= ~grp;
        __syncwarp();
    }
}

// the factor normalize() multiplies prob[d] with: norm_factor[cell_idx[cell_id[d]]], the slot number
// d used as a PARTICLE number (collisions_methods.py:647-648, quirk A1).  Fixed time step: cached
// per slot across steps (k_fused_nf_slot).  Adaptive: cell_idx is re-sorted before every sub-step,
// so the chain is followed here -- consecutive slots share a cell id, the two gathers hit the same
// line for (nearly) the whole warp.
template <bool kAdaptive>
__device__ __forceinline__ double slot_norm_factor(const FusedArgs &a, i64 d)
{
    return kAdaptive ? a.norm_factor[a.cell_idx[a.cell_id[d]]] : a.nf_slot[d];
}

// adaptive: per rank r of cell_start, the cell id it holds | parity of cell_start[cell_idx[cell id]]
// << 62 (the position find_pairs counts "even" from under the re-sorted cell_idx, quirk A2).
// Prepared per rank by the pre-pass (k_fused_norm_factor), or -- prepass_done: the sub-step was
// opened by pysdm_b200_adaptive_presort, no pre-pass ran -- followed here: three dependent
// loads per cell at the head of a batch, spread over all SMs, instead of a launch on the
// critical path between two sub-steps
__device__ __forceinline__ i64 rank_info_of(const FusedArgs &a, i64 r)
{
    if (!a.prepass_done) return a.rank_info[r];
    const i64 cid = a.rank_to_cell[r];
    return cid | ((a.cell_start[a.cell_idx[cid]] & 1) << 62);
}

template <int K, bool kPhilox, bool kWarp, bool kAggregate, bool kAdaptive>
__global__ void __launch_bounds__(kFusedThreads, PYSDM_MIN_CTAS) k_fused_step(const FusedArgs a)
{
    typedef Geo<kWarp> G;
    constexpr u32 kCellMask = kWarp ? 0x1fu : 0xffu;  // cellof[] entry -> cell number
    // (warp groups) bit 7: position beyond the working length (adaptive: shuffled, nothing else)
    constexpr u32 kRoleFirst = 0x20u, kRoleSingle = 0x40u, kRoleBeyond = 0x80u;
    extern __shared__ __align__(16) unsigned char smem_all[];
    const int tid = threadIdx.x;
    const int group = kWarp ? (tid >> 5) : 0;
    const int gt = kWarp ? (tid & 31) : tid;  // thread number within the group
    unsigned char *smem = smem_all + group * G::kBytes;
    u32 *sd = reinterpret_cast<u32 *>(smem);
    u16 *jl = reinterpret_cast<u16 *>(smem + G::kSd);  // swap targets as byte offsets into sd
    u8 *cellof = smem + G::kSd + G::kJl;
    i32 *cs = reinterpret_cast<i32 *>(smem + G::kSd + G::kJl + G::kCo);
    double *fd = reinterpret_cast<double *>(smem + G::kSd + G::kJl + G::kCo + G::kCs);
    i64 *cidv = reinterpret_cast<i64 *>(smem + G::kSd + G::kJl + G::kCo + G::kCs + G::kFd);
    // per-warp queue of colliding pairs: g_raw (f64) and local position of the pair's first (u32)
    unsigned char *qbase = smem + G::kSd + G::kJl + G::kCo + G::kCs + G::kFd + G::kCid +
                           (kWarp ? 0 : (tid >> 5) * 64 * 12);
    double *q_g = reinterpret_cast<double *>(qbase);
    u32 *q_pf = reinterpret_cast<u32 *>(qbase + 64 * 8);
    int qn = 0;  // warp-uniform fill level

    // a previous launch left zero multiplicities behind and the caller has not removed them
    // yet: this step was queued speculatively and must not happen (pysdm_b200.h, `gate`)
    if (a.gate) {
        const i64 g = *(volatile const i64 *)a.gate;
        if (g != 0 && g != a.epoch) return;
    }
    const i64 length = (kAdaptive && a.length_dev) ? *a.length_dev : a.length;
    if (kAdaptive && length == 0 && !a.shuffle_only) return;  // the sub-step loop has ended
    const i64 n_pairs_active = length / 2;
    // literal chain for a slot without a pair: max_pair -> 0; *= kernel_temp; normalize
    const double prob_nopair =
        0.0 * (K == KERNEL_GEOMETRIC || K == KERNEL_HYDRODYNAMIC
                   ? 0.0  // zero-filled pair slots: 0 ** 2 * a * 0 ** 2 * 0
                   : kernel_temp_of<K>(0.0, a.kernel_a, a.kernel_b));
    const bool unit_substeps = a.substeps == 1.0;
    const double rho_rcp = 1.0 / a.rho_w;
    const bool rho_ok = a.reserved != 0;  // host-checked precondition of div_by

    // ---- tail: pair slots beyond the working length keep `0 * norm_factor` ------------
    if (!a.shuffle_only)
        for (i64 d = n_pairs_active + (i64)blockIdx.x * kFusedThreads + tid; d < a.n_sd / 2;
             d += (i64)gridDim.x * kFusedThreads)
            a.gamma[d] = prob_nopair * slot_norm_factor<kAdaptive>(a, d) / a.substeps;

    const i64 batch = (i64)blockIdx.x * G::kGroups + group;
    const i64 c0 = batch * a.cells_per_batch;
    if (c0 >= a.n_cell) return;  // (warp groups: a whole warp leaves; no CTA barrier below)
    const i64 c1 = (c0 + a.cells_per_batch < a.n_cell) ? c0 + a.cells_per_batch : a.n_cell;
    const int ncb = (int)(c1 - c0);
    const i64 base = a.cell_start[c0], end = a.cell_start[c1];
    if (end - base > G::kPos || ncb > G::kCells) {  // caller's max_cell_size hint was wrong
        if (gt == 0) a.status[0] = 1;
        return;
    }
    const i64 base_al = base & ~(i64)3;  // local position 0 <-> a global multiple of 4
    const int lo = (int)(base - base_al);
    const int P = (int)(end - base_al);

    // ---- phase 0: cell boundaries ---------------------------------------------------------
    for (int c = gt; c <= ncb; c += G::kSize) cs[c] = (i32)(a.cell_start[c0 + c] - base_al);
    // adaptive: the cell id of every rank (also of empty ones: their dt_left is updated too) and
    // the parity of cell_start[cell_idx[cell id]], the reference point find_pairs counts even
    // positions from (pair_methods.py:40; cell_idx is consumed as cell -> rank but was re-sorted
    // by dt_left since idx was sorted, quirk A2), both prepared per rank by k_fused_norm_factor
    // A cell that has used up its time step (dt_left == 0) takes dt_todo = 0: its probabilities
    // are scaled to zero whatever they are, and they would only matter for stats_dt_min -- which
    // cannot change while it is NaN (its initial value; quirk A9).  Such "idle" cells (most cells
    // from the second sub-step on) skip the kernel evaluation and its gathers altogether.
    if (kAdaptive)
        for (int c = gt; c < ncb; c += G::kSize) {
            const i64 info = rank_info_of(a, c0 + c);
            const i64 cid = info & kCidMask;
            const double s = a.stats_dt_min[cid];
            const bool idle = a.dt_left[cid] == 0.0 && s != s;
            cidv[c] = info | ((i64)idle << 61);
        }
    group_sync<kWarp>();
    // local positions below len_local are inside the working length
    const i64 len_rel = length - base_al;
    const int len_local = !kAdaptive ? P : (len_rel < 0 ? 0 : (len_rel > (i64)P ? P : (int)len_rel));
    // parity of the position find_pairs takes as "even" within local cell c
    auto ref_parity = [&](int c) -> int {
        return kAdaptive ? (int)((cidv[c] >> 62) & 1) : (cs[c] & 1);
    };
    // roles of local position lp of cell c = [s, e): first of a pair / not covered by any pair
    // (its idx entry is written on its own) / beyond the working length
    auto role_of = [&](int lp, int c, int s, int e) -> u32 {
        if (kAdaptive && lp >= len_local) return kRoleSingle | kRoleBeyond;
        const bool even = ((lp ^ ref_parity(c)) & 1) == 0, partner = lp + 1 < e;
        if (even && partner) return kRoleFirst;
        if (even || (kAdaptive && lp == s)) return kRoleSingle;  // (an odd lp > s is a pair's second)
        return 0u;
    };

    // ---- phase 1: position -> cell map, per-cell fp64 constants of the shuffle ------------
    if (kWarp) {
        for (int c = 0; c < ncb; ++c) {
            const int s = cs[c], e = cs[c + 1];
            // warp groups (<= 32 cells per batch): bits 0-4 = cell, bit 5 = first of a pair,
            // bit 6 = unpaired last SD of an odd cell (find_pairs, pair_methods.py:38-41)
            for (int lp = s + gt; lp < e; lp += 32) cellof[lp] = (u8)((u32)c | role_of(lp, c, s, e));
        }
    } else {
        for (int c = tid >> 5; c < ncb; c += kFusedThreads / 32) {
            const int s = cs[c], e = cs[c + 1];
            for (int lp = s + (tid & 31); lp < e; lp += 32) cellof[lp] = (u8)c;
        }
    }
    for (int c = gt; c < ncb; c += G::kSize) {
        fd[2 * c] = i64_to_f64(base_al + cs[c]);           // (double)cell_start[c]
        fd[2 * c + 1] = i64_to_f64((i64)(cs[c + 1] - cs[c]));  // (double)(cell_end - cell_start)
    }
    group_sync<kWarp>();

    // Work mapping of phases 2 and 4: an iteration covers 4 * kSize consecutive positions;
    // thread gt owns two windows of two positions, (lp0, lp0 + 1) with
    // lp0 = 4*kSize*it + 2*kSize*h + 2*gt, h = 0, 1  -> a warp touches 64 consecutive
    // positions per window: 512-byte coalesced idx accesses, conflict-free 8-byte smem accesses.
    const int n_iter = (P + 4 * G::kSize - 1) / (4 * G::kSize);

    // ---- phase 2: stream idx + uniforms in, stage SD numbers and swap targets -------------
    // kFast: every window of the batch (also the padding up to a whole number of window
    // groups) lies inside the arrays and is 16-byte aligned, so the loads are unconditional
    // and the loads of kWin windows are issued together, ahead of the Philox work, to have
    // kWin * 512 B per warp in flight (the exposed idx latency was 19 % of the warp time,
    // profiles/r1h); validity is applied when staging.  Otherwise: guarded scalar loads.
    {
        const i64 *idx_b = a.idx_in + base_al;  // batch-local bases: 32-bit offsets below
        const double *u01_b = kPhilox ? nullptr : a.u01 + base_al;
        const u64 blk_b = (u64)base_al >> 1;    // Philox block of local position 0
        constexpr int kWin = PYSDM_KWIN;  // windows (of 2 * kSize positions) per group
        const int n_win = 2 * n_iter;           // window w covers lp = 2*kSize*w + 2*gt, +1
        const int n_grp = (n_win + kWin - 1) / kWin;
        const bool fast = (((uintptr_t)a.idx_in) & 15) == 0 &&
                          base_al + (i64)n_grp * kWin * 2 * G::kSize <= a.n_sd;
        // (a view shifted by an odd number of elements, as `optimized_random` makes from its
        // second sub-step on, random_generator_optimizer.py:48, is only 8-byte aligned)
        const bool u01_vec = kPhilox || (((uintptr_t)a.u01) & 15) == 0;
        auto stage = [&](int lp0, i64 s0, i64 s1, double u0, double u1) {
            const i64 svq[2] = {s0, s1};
            const double uvq[2] = {u0, u1};
            u32 jrel[2] = {0, 0};
#pragma unroll
            for (int q = 0; q < 2; ++q) {
                const int lp = lp0 + q;
                if (lp < lo || lp >= P) continue;
                const int c = cellof[lp] & kCellMask;
                const int s = cs[c], len = cs[c + 1] - s;
                // index_methods.py:40: j = int(cell_start[c] + u01[i] * (cell_end - cell_start));
                // 0 <= x < 2^33: floor by a round-down add, integer in the low mantissa bits
                const double x = fd[2 * c] + uvq[q] * fd[2 * c + 1];
                const double t = __dadd_rd(x, kTwo52);
                // j < 2^33: low word minus the low word of the cell's global start is exact
                int jr = (int)((u32)__double2loint(t) - (u32)(base_al + s));
                if (jr >= len) jr = len - 1;  // fp round-up at huge cell_start (see index.cu)
                jrel[q] = 4u * (u32)phys(s + jr);  // byte offset of the swap target in sd[]
            }
            if (lp0 >= lo && lp0 + 1 < P) {
                const int ph = phys(lp0);
                *reinterpret_cast<uint2 *>(sd + ph) = make_uint2((u32)svq[0], (u32)svq[1]);
                *reinterpret_cast<u32 *>(jl + ph) = jrel[0] | (jrel[1] << 16);
            } else {
                if (lp0 >= lo && lp0 < P) {
                    sd[phys(lp0)] = (u32)svq[0];
                    jl[phys(lp0)] = (u16)jrel[0];
                }
                if (lp0 + 1 >= lo && lp0 + 1 < P) {
                    sd[phys(lp0 + 1)] = (u32)svq[1];
                    jl[phys(lp0 + 1)] = (u16)jrel[1];
                }
            }
        };
        auto philox_pair = [&](int lp0, double &u0, double &u1) {
            u32 w[4];
            const u64 b = blk_b + (u32)(lp0 >> 1);  // block = elements 2b, 2b + 1
            philox4x32_10_rk((u32)b, (u32)(b >> 32), (u32)a.call_u01, (u32)(a.call_u01 >> 32),
                             a.philox_round_keys, w);
            u0 = u01_from_words(w[0], w[1]);
            u1 = u01_from_words(w[2], w[3]);
        };
        if (fast) {
            for (int g = 0; g < n_grp; ++g) {
                longlong2 sv[kWin];
                double2 uv[kWin];
#pragma unroll
                for (int w = 0; w < kWin; ++w) {
                    const int lp0 = 2 * G::kSize * (g * kWin + w) + 2 * gt;
                    sv[w] = *reinterpret_cast<const longlong2 *>(idx_b + (u32)lp0);
                    if (!kPhilox) {
                        if (u01_vec) {
                            uv[w] = *reinterpret_cast<const double2 *>(u01_b + (u32)lp0);
                        } else {
                            uv[w].x = u01_b[(u32)lp0];
                            uv[w].y = u01_b[(u32)lp0 + 1];
                        }
                    }
                }
#pragma unroll
                for (int w = 0; w < kWin; ++w) {
                    const int lp0 = 2 * G::kSize * (g * kWin + w) + 2 * gt;
                    if (kPhilox && lp0 < P) philox_pair(lp0, uv[w].x, uv[w].y);
                }
#pragma unroll
                for (int w = 0; w < kWin; ++w) {
                    const int lp0 = 2 * G::kSize * (g * kWin + w) + 2 * gt;
                    if (lp0 < P) stage(lp0, sv[w].x, sv[w].y, uv[w].x, uv[w].y);
                }
            }
        } else {
            for (int w = 0; w < n_win; ++w) {
                const int lp0 = 2 * G::kSize * w + 2 * gt;
                const bool v0 = lp0 >= lo && lp0 < P, v1 = lp0 + 1 < P;
                i64 s0 = 0, s1 = 0;
                double u0 = 0.0, u1 = 0.0;
                if (v0) s0 = idx_b[(u32)lp0];
                if (v1) s1 = idx_b[(u32)lp0 + 1];
                if (kPhilox) {
                    if (lp0 < P) philox_pair(lp0, u0, u1);
                } else {
                    if (v0) u0 = u01_b[(u32)lp0];
                    if (v1) u1 = u01_b[(u32)lp0 + 1];
                }
                if (lp0 < P) stage(lp0, s0, s1, u0, u1);
            }
        }
    }
    group_sync<kWarp>();

    // ---- phase 3: one lane per cell replays the swap chain in shared memory ---------------
    for (int c = gt; c < ncb; c += G::kSize) {
        const int s = cs[c], e = cs[c + 1];
        i64 cid = 0;
        if (!kAdaptive && e > s && !a.shuffle_only)
            cid = a.cell_id[sd[phys(s)]];  // issued now, consumed after the chain
        // positions of a 64-block are contiguous in the skewed layout: walk block by block
        // with unit stride (the bookkeeping of the skew was half of the chain's instructions)
        int i = e - 1;
        while (i > s) {
            const int blk = i & ~63;
            const int lowest = blk > s ? blk : s + 1;  // last position handled in this block
            const int n = i - lowest + 1;
            u32 *psd = sd + phys(i);
            const u16 *pjl = jl + phys(i);
#pragma unroll 4
            for (int t = 0; t < n; ++t) {
                u32 *q = reinterpret_cast<u32 *>(reinterpret_cast<unsigned char *>(sd) + pjl[-t]);
                const u32 x = psd[-t], y = *q;
                psd[-t] = y;
                *q = x;
            }
            i = lowest - 1;
        }
        if (!kAdaptive) cidv[c] = cid;
        // adaptive: fd[] (constants of the shuffle) is free now; [2c + 1] collects the cell's
        // minimal dt_optimal, [2c] will hold dt_todo / dt
        if (kAdaptive) fd[2 * c + 1] = __longlong_as_double(0x7ff0000000000000ll);
    }
    group_sync<kWarp>();

    if (a.shuffle_only) {  // pysdm_b200_shuffle_local_batched: the permuted order, nothing else
        for (int lp = lo + gt; lp < P; lp += G::kSize) a.idx_out[base_al + lp] = (i64)sd[phys(lp)];
        return;
    }

    // ---- phase 4: two pair slots per thread and iteration ---------------------------------
    // per-SD operand of the collision kernel: water mass, or the radius attribute (Geometric)
    constexpr bool kGravitational = K == KERNEL_GEOMETRIC || K == KERNEL_HYDRODYNAMIC;
    const double *mass = kGravitational ? a.radius : a.attributes + a.mass_row * a.n_sd;
    const bool flag_al = (((uintptr_t)a.is_first_in_pair) & 1) == 0;
    const bool vec_out = (((uintptr_t)a.idx_out) & 15) == 0;
    // batch-local bases (32-bit offsets in the loop); dl = local pair slot = lp0 / 2
    u8 *flag_b = a.is_first_in_pair + base_al;
    i64 *out_b = a.idx_out + base_al;
    double *gamma_b = a.gamma + (base_al >> 1);
    const double *nf_b = a.nf_slot + (base_al >> 1);
    const double *rand_b = kPhilox ? nullptr : a.rand + (base_al >> 1);
    const i64 dl_left = n_pairs_active - (base_al >> 1);  // slots below go through compute_gamma
    const int dl_active = dl_left > (i64)G::kPos ? G::kPos : (dl_left < 0 ? 0 : (int)dl_left);
    if constexpr (!kAdaptive) {
    for (int it = 0; it < n_iter; ++it) {  // uniform trip count: warp-collective code inside
        double rnd[2] = {0.0, 0.0};

        // stage A: find_pairs (pair_methods.py:38-41, :55) for my positions, pair members
        bool has[2], dflt[2];
        int pf[2];
        u32 j[2], k[2];
        i64 nj[2], nk[2];
        double mj[2], mk[2], nf[2];
        double uj[2] = {0.0, 0.0}, uk[2] = {0.0, 0.0};  // fall velocities (Geometric)
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const int lp0 = 4 * G::kSize * it + 2 * G::kSize * h + 2 * gt, lp1 = lp0 + 1;
            const bool v0 = lp0 >= lo && lp0 < P, v1 = lp1 >= lo && lp1 < P;
            bool first0 = false, first1 = false, single0 = false, single1 = false;
            if (kWarp) {  // roles were packed next to the cell number in phase 1
                const u32 both = *reinterpret_cast<const u16 *>(cellof + lp0);  // lp0 is even
                const u32 b0 = v0 ? (both & 0xffu) : 0u, b1 = v1 ? (both >> 8) : 0u;
                first0 = (b0 & kRoleFirst) != 0;
                single0 = (b0 & kRoleSingle) != 0;
                first1 = (b1 & kRoleFirst) != 0;
                single1 = (b1 & kRoleSingle) != 0;
            } else {
                if (v0) {
                    const int c0l = cellof[lp0];
                    const u32 r0 = role_of(lp0, c0l, cs[c0l], cs[c0l + 1]);
                    first0 = (r0 & kRoleFirst) != 0;
                    single0 = (r0 & kRoleSingle) != 0;
                }
                if (v1) {
                    const int c1l = cellof[lp1];
                    const u32 r1 = role_of(lp1, c1l, cs[c1l], cs[c1l + 1]);
                    first1 = (r1 & kRoleFirst) != 0;
                    single1 = (r1 & kRoleSingle) != 0;
                }
            }
            if (v0 && v1 && flag_al) {
                *reinterpret_cast<u16 *>(flag_b + (u32)lp0) =
                    (u16)((first0 ? 1u : 0u) | (first1 ? 0x100u : 0u));
            } else {
                if (v0) flag_b[(u32)lp0] = (u8)first0;
                if (v1) flag_b[(u32)lp1] = (u8)first1;
            }
            if (single0) out_b[(u32)lp0] = (i64)sd[phys(lp0)];
            if (single1) out_b[(u32)lp1] = (i64)sd[phys(lp1)];

            const u32 dl = (u32)lp0 >> 1;
            has[h] = first0 || first1;
            pf[h] = first0 ? lp0 : lp1;
            dflt[h] = false;
            j[h] = k[h] = 0;
            nj[h] = nk[h] = 1;
            mj[h] = mk[h] = nf[h] = 0.0;
            if (has[h]) {
                if (first0) {
                    const uint2 t = *reinterpret_cast<const uint2 *>(sd + phys(lp0));
                    j[h] = t.x;
                    k[h] = t.y;
                } else {
                    j[h] = sd[phys(lp1)];
                    k[h] = sd[phys(lp1 + 1)];
                }
            } else if (v0 && (int)dl < dl_active) {
                // no pair starts inside this window; the slot still goes through compute_gamma
                // unless a pair starting at its second position, in the NEXT batch, owns it
                bool next_first = false;
                if (!v1) {  // lp0 is this batch's last position and p0 + 1 < length
                    i64 rr = c1;
                    while (rr < a.n_cell && a.cell_start[rr + 1] == end) ++rr;
                    next_first = rr < a.n_cell && a.cell_start[rr + 1] - end >= 2;
                }
                dflt[h] = !next_first;
            }
            if (has[h] || dflt[h]) {
                nf[h] = nf_b[dl];  // norm_factor[cell_idx[cell_id[slot]]] (quirk A1), cached
                if (!kPhilox) rnd[h] = rand_b[dl];
            }
        }
#pragma unroll
        for (int h = 0; h < 2; ++h)
            if (has[h]) {
                nj[h] = a.multiplicity[j[h]];
                nk[h] = a.multiplicity[k[h]];
                if (K != KERNEL_CONSTANT_K) {
                    mj[h] = mass[j[h]];
                    mk[h] = mass[k[h]];
                }
                if (kGravitational) {
                    uj[h] = a.fall_velocity[j[h]];
                    uk[h] = a.fall_velocity[k[h]];
                }
            }

        // uniforms of this warp's 2 x 32 slots: 32 Philox blocks, one per lane (block b holds
        // elements 2b, 2b + 1); lane x gets slot (h, x) from lane 16h + x/2
        if (kPhilox) {
            const int lane = tid & 31, wig = (gt >> 5) << 5;  // first thread of my warp in group
            // local slot (even) whose block this lane computes; global slot = base_al/2 + it
            const int dl_mine = 2 * G::kSize * it + G::kSize * (lane >> 4) + wig + 2 * (lane & 15);
            u32 w[4];
            const u64 b = ((u64)base_al >> 2) + (u32)(dl_mine >> 1);
            philox4x32_10_rk((u32)b, (u32)(b >> 32), (u32)a.call_rand, (u32)(a.call_rand >> 32),
                             a.philox_round_keys, w);
            const double r0 = u01_from_words(w[0], w[1]), r1 = u01_from_words(w[2], w[3]);
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const int src = 16 * h + (lane >> 1);
                const double e0 = __shfl_sync(0xffffffffu, r0, src);
                const double e1 = __shfl_sync(0xffffffffu, r1, src);
                rnd[h] = (lane & 1) ? e1 : e0;
            }
        }
        // (issued after the gathers above so that the Philox arithmetic overlaps their latency)

        // stage B: sort within pair, probability, gamma.  Pairs that do collide (gamma != 0, a
        // few per cent) are queued per warp and handled with full lanes by process_hits()
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const u32 dl = (u32)(4 * G::kSize * it + 2 * G::kSize * h + 2 * gt) >> 1;
            double g_raw = 0.0;
            if (has[h]) {
                u32 sj = j[h], sk = k[h];
                i64 n_j = nj[h];
                double m_j = mj[h], m_k = mk[h];
                if (nj[h] < nk[h]) {  // sort_within_pair_by_attr (ties keep order)
                    sj = k[h]; sk = j[h];
                    n_j = nk[h];
                    m_j = mk[h]; m_k = mj[h];
                }
                if (vec_out && (pf[h] & 1) == 0) {  // base_al is even: alignment of pf = of base_al + pf
                    *reinterpret_cast<longlong2 *>(out_b + (u32)pf[h]) =
                        make_longlong2((i64)sj, (i64)sk);
                } else {
                    out_b[(u32)pf[h]] = (i64)sj;
                    out_b[(u32)pf[h] + 1] = (i64)sk;
                }
                // volume = mass / rho_w (physics_methods.py:107-119); sum_pair; kernel; max_pair
                double kernel_temp;
                if (K == KERNEL_GEOMETRIC) {  // |u_j - u_k| is symmetric: no need to swap uj / uk
                    kernel_temp = geometric_kernel_temp(m_j, m_k, uj[h], uk[h], a.kernel_a);
                } else if (K == KERNEL_HYDRODYNAMIC) {
                    kernel_temp = hydrodynamic_kernel_temp(a, m_j, m_k, uj[h], uk[h]);
                } else {
                    double vsum = 0.0;
                    if (K != KERNEL_CONSTANT_K)
                        vsum = div_by(m_j, a.rho_w, rho_rcp, rho_ok) + div_by(m_k, a.rho_w, rho_rcp, rho_ok);
                    kernel_temp = kernel_temp_of<K>(vsum, a.kernel_a, a.kernel_b);
                }
                double prob = i64_to_f64(n_j) * kernel_temp;
                prob = prob * nf[h];
                if (!unit_substeps) prob = prob / a.substeps;
                g_raw = ceil_exact(prob - rnd[h]);
                gamma_b[dl] = g_raw;  // final unless capped by n_j // n_k (process_hits)
            } else if (dflt[h]) {
                double prob = prob_nopair * nf[h];
                if (!unit_substeps) prob = prob / a.substeps;
                gamma_b[dl] = ceil_exact(prob - rnd[h]);
            }
            // queue the hits of this warp (g_raw == 0, also -0, means no collision)
            const bool hit = has[h] && g_raw != 0.0;
            const unsigned hits = __ballot_sync(0xffffffffu, hit);
            if (hit) {
                const int slot = qn + __popc(hits & ((1u << (tid & 31)) - 1u));
                q_pf[slot] = (u32)pf[h];
                q_g[slot] = g_raw;
            }
            qn += __popc(hits);
            __syncwarp();
            if (qn >= 32) {
                qn -= 32;
                process_hits<K, kAggregate, false>(a, q_pf[qn + (tid & 31)], q_g[qn + (tid & 31)], true, sd,
                                            cellof, kCellMask, cidv, gamma_b);
                __syncwarp();
            }
        }
    }
    if (qn > 0) {
        const bool mine = (tid & 31) < qn;
        process_hits<K, kAggregate, false>(a, mine ? q_pf[tid & 31] : 0u, mine ? q_g[tid & 31] : 0.0, mine, sd,
                                    cellof, kCellMask, cidv, gamma_b);
    }
    } else {
    // ================= adaptive sub-step: two passes over the batch ==========================
    // pass A: pairs, probabilities (left in `gamma`, as the reference leaves them in `prob`) and the
    //         per-cell minimum of dt_optimal = max(dt * (n_j // n_k) / prob, dt_min)
    //         (scale_prob_for_adaptive_sdm_gamma, collisions_methods.py:357-367);
    // per cell: dt_todo, dt_left, stats (:355-356, :373-376), adaptive_sdm_end (:313-328);
    // pass B: prob *= dt_todo / dt (:368-372), compute_gamma, coalescence.
    const int lane = tid & 31;
    // roles of my two positions of window (it, h); returns the pair slot's local number
    auto roles = [&](int it, int h, bool &v0, bool &first0, bool &first1, bool &single0,
                     bool &single1, bool &a0, bool &a1) -> int {
        const int lp0 = 4 * G::kSize * it + 2 * G::kSize * h + 2 * gt, lp1 = lp0 + 1;
        v0 = lp0 >= lo && lp0 < P;
        const bool v1 = lp1 >= lo && lp1 < P;
        u32 r0 = 0, r1 = 0;
        if (kWarp) {
            const u32 both = *reinterpret_cast<const u16 *>(cellof + lp0);  // lp0 is even
            r0 = v0 ? (both & 0xffu) : 0u;
            r1 = v1 ? (both >> 8) : 0u;
        } else {
            if (v0) {
                const int c = cellof[lp0];
                r0 = role_of(lp0, c, cs[c], cs[c + 1]);
            }
            if (v1) {
                const int c = cellof[lp1];
                r1 = role_of(lp1, c, cs[c], cs[c + 1]);
            }
        }
        first0 = (r0 & kRoleFirst) != 0;
        first1 = (r1 & kRoleFirst) != 0;
        single0 = (r0 & kRoleSingle) != 0;
        single1 = (r1 & kRoleSingle) != 0;
        a0 = v0 && (r0 & kRoleBeyond) == 0;  // inside the working length
        a1 = v1 && (r1 & kRoleBeyond) == 0;
        return lp0;
    };
    // a slot in which no pair starts still goes through compute_gamma (with prob = 0 * nf) unless
    // a pair starting at its second position, in the NEXT batch, owns it
    auto default_slot = [&](int lp0, bool v0, bool a1valid) -> bool {
        if (!v0 || (int)((u32)lp0 >> 1) >= dl_active) return false;
        if (a1valid) return true;
        if (lp0 + 1 < P) return true;  // second position is mine and starts no pair
        i64 rr = c1;
        while (rr < a.n_cell && a.cell_start[rr + 1] == end) ++rr;
        if (rr >= a.n_cell || a.cell_start[rr + 1] - end < 2) return true;
        return (((i64)end ^ (rank_info_of(a, rr) >> 62)) & 1) != 0;  // its first SD is not "even"
    };

    // ---- pass A ----------------------------------------------------------------------------
    for (int it = 0; it < n_iter; ++it) {
        bool has[2], dflt[2], idle[2];
        int pf[2];
        u32 j[2], k[2];
        i64 nj[2], nk[2];
        double mj[2], mk[2], nf[2];
        double uj[2] = {0.0, 0.0}, uk[2] = {0.0, 0.0};
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            bool v0, first0, first1, single0, single1, a0, a1;
            const int lp0 = roles(it, h, v0, first0, first1, single0, single1, a0, a1);
            const int lp1 = lp0 + 1;
            if (a0 && a1 && flag_al) {
                *reinterpret_cast<u16 *>(flag_b + (u32)lp0) =
                    (u16)((first0 ? 1u : 0u) | (first1 ? 0x100u : 0u));
            } else {
                if (a0) flag_b[(u32)lp0] = (u8)first0;
                if (a1) flag_b[(u32)lp1] = (u8)first1;
            }
            if (single0) out_b[(u32)lp0] = (i64)sd[phys(lp0)];
            if (single1) out_b[(u32)lp1] = (i64)sd[phys(lp1)];
            const u32 dl = (u32)lp0 >> 1;
            has[h] = first0 || first1;
            pf[h] = first0 ? lp0 : lp1;
            dflt[h] = false;
            j[h] = k[h] = 0;
            nj[h] = nk[h] = 1;
            mj[h] = mk[h] = nf[h] = 0.0;
            idle[h] = false;
            if (has[h]) {
                j[h] = sd[phys(pf[h])];
                k[h] = sd[phys(pf[h] + 1)];
                idle[h] = ((cidv[cellof[pf[h]] & kCellMask] >> 61) & 1) != 0;
            } else {
                dflt[h] = default_slot(lp0, v0, lp1 < P);
            }
            if ((has[h] && !idle[h]) || dflt[h])
                nf[h] = slot_norm_factor<true>(a, (base_al >> 1) + (i64)dl);
        }
#pragma unroll
        for (int h = 0; h < 2; ++h)
            if (has[h]) {
                nj[h] = a.multiplicity[j[h]];
                nk[h] = a.multiplicity[k[h]];
                if (idle[h]) continue;
                if (K != KERNEL_CONSTANT_K) {
                    mj[h] = mass[j[h]];
                    mk[h] = mass[k[h]];
                }
                if (kGravitational) {
                    uj[h] = a.fall_velocity[j[h]];
                    uk[h] = a.fall_velocity[k[h]];
                }
            }
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const u32 dl = (u32)(4 * G::kSize * it + 2 * G::kSize * h + 2 * gt) >> 1;
            double dt_opt = 0.0;
            bool candidate = false;
            if (has[h]) {
                u32 sj = j[h], sk = k[h];
                i64 n_j = nj[h], n_k = nk[h];
                double m_j = mj[h], m_k = mk[h];
                if (nj[h] < nk[h]) {  // sort_within_pair_by_attr (ties keep order)
                    sj = k[h]; sk = j[h];
                    n_j = nk[h]; n_k = nj[h];
                    m_j = mk[h]; m_k = mj[h];
                }
                if (vec_out && (pf[h] & 1) == 0) {
                    *reinterpret_cast<longlong2 *>(out_b + (u32)pf[h]) =
                        make_longlong2((i64)sj, (i64)sk);
                } else {
                    out_b[(u32)pf[h]] = (i64)sj;
                    out_b[(u32)pf[h] + 1] = (i64)sk;
                }
                double prob = 0.0;  // idle cell: the scaled probability is 0 in any case
                if (!idle[h]) {
                    double kernel_temp;
                    if (K == KERNEL_GEOMETRIC) {
                        kernel_temp = geometric_kernel_temp(m_j, m_k, uj[h], uk[h], a.kernel_a);
                    } else if (K == KERNEL_HYDRODYNAMIC) {
                        kernel_temp = hydrodynamic_kernel_temp(a, m_j, m_k, uj[h], uk[h]);
                    } else {
                        double vsum = 0.0;
                        if (K != KERNEL_CONSTANT_K)
                            vsum = div_by(m_j, a.rho_w, rho_rcp, rho_ok) +
                                   div_by(m_k, a.rho_w, rho_rcp, rho_ok);
                        kernel_temp = kernel_temp_of<K>(vsum, a.kernel_a, a.kernel_b);
                    }
                    prob = i64_to_f64(n_j) * kernel_temp;
                    prob = prob * nf[h];
                }
                gamma_b[dl] = prob;
                if (prob != 0.0) {  // pair_indices: prob == 0 skips the pair (:360-362)
                    const i64 prop = floordiv_sorted(n_j, n_k);
                    dt_opt = a.timestep * i64_to_f64(prop) / prob;
                    dt_opt = py_max(dt_opt, a.dt_min);
                    candidate = dt_opt == dt_opt;  // (a NaN never wins a `<` comparison)
                }
            } else if (dflt[h]) {
                gamma_b[dl] = prob_nopair * nf[h];
            }
            warp_cell_min<kWarp>(dt_opt, (int)(cellof[has[h] ? pf[h] : 0] & kCellMask), candidate,
                                 fd, lane);
        }
    }
    group_sync<kWarp>();

    // ---- per cell: time step to take, bookkeeping by cell id -------------------------------------
    {
        u32 my_end = 0;  // adaptive_sdm_end: cell_start[i + 1] of cells i with dt_left[i] != 0
        for (int c = gt; c < ncb; c += G::kSize) {
            const i64 cid = cidv[c] & kCidMask;
            const double pm = fd[2 * c + 1];  // +inf when no pair of the cell was a candidate
            const double left = a.dt_left[cid];
            double todo = py_min(left, a.dt_max);  // :355-356
            todo = pm < todo ? pm : todo;          // :366
            const double cur = a.stats_dt_min[cid];
            if (pm < cur) a.stats_dt_min[cid] = pm;  // :367, a NaN stays (quirk A9)
            fd[2 * c] = todo / a.timestep;           // factor of pass B
            const double new_left = left - todo;     // :373-376
            a.dt_left[cid] = new_left;
            if (todo > 0) a.stats_n_substep[cid] += 1;
            if (a.adaptive_end != nullptr && new_left != 0) {
                const u32 e = (u32)a.cell_start[cid + 1];  // (cell id used as a rank: quirk A2)
                my_end = e > my_end ? e : my_end;
            }
        }
        if (a.adaptive_end != nullptr) {
            const u32 w_end = __reduce_max_sync(0xffffffffu, my_end);
            if (lane == 0 && w_end != 0)
                atomicMax(reinterpret_cast<unsigned long long *>(a.adaptive_end),
                          (unsigned long long)w_end);
        }
    }
    group_sync<kWarp>();

    // ---- pass B ----------------------------------------------------------------------------
    for (int it = 0; it < n_iter; ++it) {
        double rnd[2] = {0.0, 0.0};
        bool has[2], dflt[2];
        int pf[2];
        double prob[2] = {0.0, 0.0};
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            bool v0, first0, first1, single0, single1, a0, a1;
            const int lp0 = roles(it, h, v0, first0, first1, single0, single1, a0, a1);
            const u32 dl = (u32)lp0 >> 1;
            has[h] = first0 || first1;
            pf[h] = first0 ? lp0 : lp0 + 1;
            dflt[h] = !has[h] && default_slot(lp0, v0, lp0 + 1 < P);
            if (has[h] || dflt[h]) {
                prob[h] = gamma_b[dl];  // written by this thread in pass A
                if (!kPhilox) rnd[h] = rand_b[dl];
            }
        }
        if (kPhilox) {  // as in the non-adaptive loop: 32 blocks per warp, one per lane
            const int wig = (gt >> 5) << 5;
            const int dl_mine = 2 * G::kSize * it + G::kSize * (lane >> 4) + wig + 2 * (lane & 15);
            u32 w[4];
            const u64 b = ((u64)base_al >> 2) + (u32)(dl_mine >> 1);
            philox4x32_10_rk((u32)b, (u32)(b >> 32), (u32)a.call_rand, (u32)(a.call_rand >> 32),
                             a.philox_round_keys, w);
            const double r0 = u01_from_words(w[0], w[1]), r1 = u01_from_words(w[2], w[3]);
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const int src = 16 * h + (lane >> 1);
                const double e0 = __shfl_sync(0xffffffffu, r0, src);
                const double e1 = __shfl_sync(0xffffffffu, r1, src);
                rnd[h] = (lane & 1) ? e1 : e0;
            }
        }
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const u32 dl = (u32)(4 * G::kSize * it + 2 * G::kSize * h + 2 * gt) >> 1;
            double g_raw = 0.0;
            if (has[h] || dflt[h]) {
                double pr = prob[h];
                if (has[h] && pr != 0.0) pr = pr * fd[2 * (cellof[pf[h]] & kCellMask)];
                g_raw = ceil_exact(pr - rnd[h]);
                gamma_b[dl] = g_raw;
            }
            const bool hit = has[h] && g_raw != 0.0;
            const unsigned hits = __ballot_sync(0xffffffffu, hit);
            if (hit) {
                const int slot = qn + __popc(hits & ((1u << lane) - 1u));
                q_pf[slot] = (u32)pf[h];
                q_g[slot] = g_raw;
            }
            qn += __popc(hits);
            __syncwarp();
            if (qn >= 32) {
                qn -= 32;
                process_hits<K, kAggregate, kAdaptive>(a, q_pf[qn + lane], q_g[qn + lane], true, sd,
                                                       cellof, kCellMask, cidv, gamma_b);
                __syncwarp();
            }
        }
    }
    if (qn > 0) {
        const bool mine = lane < qn;
        process_hits<K, kAggregate, kAdaptive>(a, mine ? q_pf[lane] : 0u, mine ? q_g[lane] : 0.0, mine,
                                               sd, cellof, kCellMask, cidv, gamma_b);
    }
    }
}

// ---- launcher -----------------------------------------------------------------------------
template <int K, bool kPhilox, bool kWarp, bool kAggregate, bool kAdaptive>
static int launch_fused(const FusedArgs &a, i64 n_batches, cudaStream_t st)
{
    // the attribute is per device and cheap to set: no caching (a process may drive several GPUs)
    const int smem = Geo<kWarp>::kSmem;
    int err = (int)cudaFuncSetAttribute(k_fused_step<K, kPhilox, kWarp, kAggregate, kAdaptive>,
                                        cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (err) return err;
    const i64 grid = (n_batches + Geo<kWarp>::kGroups - 1) / Geo<kWarp>::kGroups;
    k_fused_step<K, kPhilox, kWarp, kAggregate, kAdaptive>
        <<<(unsigned)(grid < 1 ? 1 : grid), kFusedThreads, smem, st>>>(a);
    return LAUNCH_CHECK();
}

// one definition per collision kernel K, each in its own translation unit (fused_k*.cu)
template <int K>
int dispatch_fused(const FusedArgs &a, i64 n_batches, bool philox, bool warp, bool aggregate,
                   bool adaptive, cudaStream_t st);

#define PYSDM_FUSED_DEFINE_DISPATCH(K)                                                         \
    template <>                                                                                \
    int dispatch_fused<K>(const FusedArgs &a, i64 n_batches, bool philox, bool warp,           \
                          bool aggregate, bool adaptive, cudaStream_t st)                      \
    {                                                                                          \
        const int code = (philox ? 8 : 0) | (warp ? 4 : 0) | (aggregate ? 2 : 0) | (adaptive ? 1 : 0); \
        switch (code) {                                                                        \
        case 0: return launch_fused<K, false, false, false, false>(a, n_batches, st);          \
        case 1: return launch_fused<K, false, false, false, true>(a, n_batches, st);           \
        case 2: return launch_fused<K, false, false, true, false>(a, n_batches, st);           \
        case 3: return launch_fused<K, false, false, true, true>(a, n_batches, st);            \
        case 4: return launch_fused<K, false, true, false, false>(a, n_batches, st);           \
        case 5: return launch_fused<K, false, true, false, true>(a, n_batches, st);            \
        case 6: return launch_fused<K, false, true, true, false>(a, n_batches, st);            \
        case 7: return launch_fused<K, false, true, true, true>(a, n_batches, st);             \
        case 8: return launch_fused<K, true, false, false, false>(a, n_batches, st);           \
        case 9: return launch_fused<K, true, false, false, true>(a, n_batches, st);            \
        case 10: return launch_fused<K, true, false, true, false>(a, n_batches, st);           \
        case 11: return launch_fused<K, true, false, true, true>(a, n_batches, st);            \
        case 12: return launch_fused<K, true, true, false, false>(a, n_batches, st);           \
        case 13: return launch_fused<K, true, true, false, true>(a, n_batches, st);            \
        case 14: return launch_fused<K, true, true, true, false>(a, n_batches, st);            \
        default: return launch_fused<K, true, true, true, true>(a, n_batches, st);             \
        }                                                                                      \
    }
