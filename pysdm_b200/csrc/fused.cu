// fused.cu -- one-launch SDM collision step for cell-sorted populations (sm_100a).
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
//   prob /= substeps         collision.py:279
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
#include "berry.cuh"
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

// ---- pre-pass: norm_factor per cell rank (normalize, collisions_methods.py:639-646) ---------
__global__ void k_fused_norm_factor(const i64 *__restrict__ cell_start, i64 n_cell,
                                    double *__restrict__ norm_factor, double timestep, double dv)
{
    i64 r = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (r < n_cell)
        norm_factor[r] = sdm_norm_factor(cell_start[r + 1] - cell_start[r], timestep, dv);
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
template <int K, bool kAggregate>
__device__ __forceinline__ void process_hits(const FusedArgs &a, u32 pf, double g_raw, bool mine,
                                             const u32 *sd, const u8 *cellof, u32 cell_mask,
                                             const i64 *cidv, double *gamma_b)
{
    bool active = false;
    i64 cid = 0, add_rate = 0, add_deficit = 0, add_coal = 0;
    if (mine) {
        u32 sj = sd[phys((int)pf)], sk = sd[phys((int)pf + 1)];
        i64 n_j = a.multiplicity[sj], n_k = a.multiplicity[sk];
        if (n_j < n_k) {  // sort_within_pair_by_attr (ties keep order)
            const u32 t = sj; sj = sk; sk = t;
            const i64 tn = n_j; n_j = n_k; n_k = tn;
        }
        // g = min(int(gamma), n_j // n_k); n_j >= n_k > 0 here, so int(gamma) <= 1 needs no division
        const i64 gi = f64_to_i64(g_raw);
        i64 g = gi;
        if (gi > 1 || n_k <= 0) {
            const i64 prop = floordiv(n_j, n_k);
            g = gi < prop ? gi : prop;
        }
        const double gamma = i64_to_f64(g);
        if (g != gi) gamma_b[pf >> 1] = gamma;
        cid = cidv[cellof[pf] & cell_mask];
        add_rate = g * n_k;
        add_deficit = (gi - g) * n_k;
        active = true;
        if (gamma != 0.0) {
            // coalesce: float arithmetic stored to int64 (quirk A7), attribute update fused (fma)
            const double f_nj = i64_to_f64(n_j), f_nk = i64_to_f64(n_k);
            const double gnk = gamma * f_nk;
            add_coal = f64_to_i64(gnk);
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
            for (i64 r_ = 0; r_ < a.n_attr; ++r_) {
                double *row = a.attributes + r_ * a.n_sd;
                const double v = fma(gamma, row[sj], row[sk]);
                row[sk] = v;
                if (!keep) row[sj] = v;
            }
            if (nj_new == 0 || nk_new == 0) {
                a.healthy[0] = 0;
                if (a.gate) a.gate[0] = a.epoch;
            }
        }
    }
    counter_add<kAggregate>(a.collision_rate, cid, add_rate, active);
    counter_add<kAggregate>(a.collision_rate_deficit, cid, add_deficit, active);
    counter_add<kAggregate>(a.coalescence_rate, cid, add_coal, active);
}

template <int K, bool kPhilox, bool kWarp, bool kAggregate>
__global__ void __launch_bounds__(kFusedThreads, PYSDM_MIN_CTAS) k_fused_step(const FusedArgs a)
{
    typedef Geo<kWarp> G;
    constexpr u32 kCellMask = kWarp ? 0x1fu : 0xffu;  // cellof[] entry -> cell number
    constexpr u32 kRoleFirst = 0x20u, kRoleSingle = 0x40u;
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
    const i64 n_pairs_active = a.length / 2;
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
            a.gamma[d] = prob_nopair * a.nf_slot[d] / a.substeps;

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
    group_sync<kWarp>();

    // ---- phase 1: position -> cell map, per-cell fp64 constants of the shuffle ------------
    if (kWarp) {
        for (int c = 0; c < ncb; ++c) {
            const int s = cs[c], e = cs[c + 1];
            // warp groups (<= 32 cells per batch): bits 0-4 = cell, bit 5 = first of a pair,
            // bit 6 = unpaired last SD of an odd cell (find_pairs, pair_methods.py:38-41)
            for (int lp = s + gt; lp < e; lp += 32) {
                const bool even = ((lp - s) & 1) == 0, partner = lp + 1 < e;
                cellof[lp] = (u8)(c | (even && partner ? kRoleFirst : 0) |
                                  (even && !partner ? kRoleSingle : 0));
            }
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
                          (kPhilox || (((uintptr_t)a.u01) & 15) == 0) &&
                          base_al + (i64)n_grp * kWin * 2 * G::kSize <= a.n_sd;
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
                    if (!kPhilox) uv[w] = *reinterpret_cast<const double2 *>(u01_b + (u32)lp0);
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
        if (e > s && !a.shuffle_only)
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
        cidv[c] = cid;
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
                    const bool even = ((lp0 - cs[c0l]) & 1) == 0, partner = lp0 + 1 < cs[c0l + 1];
                    first0 = even && partner;
                    single0 = even && !partner;
                }
                if (v1) {
                    const int c1l = cellof[lp1];
                    const bool even = ((lp1 - cs[c1l]) & 1) == 0, partner = lp1 + 1 < cs[c1l + 1];
                    first1 = even && partner;
                    single1 = even && !partner;
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
                process_hits<K, kAggregate>(a, q_pf[qn + (tid & 31)], q_g[qn + (tid & 31)], true, sd,
                                            cellof, kCellMask, cidv, gamma_b);
                __syncwarp();
            }
        }
    }
    if (qn > 0) {
        const bool mine = (tid & 31) < qn;
        process_hits<K, kAggregate>(a, mine ? q_pf[tid & 31] : 0u, mine ? q_g[tid & 31] : 0.0, mine, sd,
                                    cellof, kCellMask, cidv, gamma_b);
    }
}

// ---- launcher -----------------------------------------------------------------------------
PYSDM_EXPORT int64_t pysdm_b200_fused_max_cell_size(void) { return kCtaPos; }

template <int K, bool kPhilox, bool kWarp, bool kAggregate>
static int launch_fused(const FusedArgs &a, i64 n_batches, cudaStream_t st)
{
    static bool configured = false;  // one caller thread per process (SURVEY 8b "Threading")
    const int smem = Geo<kWarp>::kSmem;
    if (!configured) {
        int err = (int)cudaFuncSetAttribute(k_fused_step<K, kPhilox, kWarp, kAggregate>,
                                            cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        if (err) return err;
        configured = true;
    }
    const i64 grid = (n_batches + Geo<kWarp>::kGroups - 1) / Geo<kWarp>::kGroups;
    k_fused_step<K, kPhilox, kWarp, kAggregate>
        <<<(unsigned)(grid < 1 ? 1 : grid), kFusedThreads, smem, st>>>(a);
    return LAUNCH_CHECK();
}

template <int K>
static int dispatch_fused(const FusedArgs &a, i64 n_batches, bool philox, bool warp, bool aggregate,
                          cudaStream_t st)
{
#define PYSDM_FUSED_CASE(P, W, A)                                                     \
    if (philox == P && warp == W && aggregate == A) return launch_fused<K, P, W, A>(a, n_batches, st);
    PYSDM_FUSED_CASE(false, false, false)
    PYSDM_FUSED_CASE(false, false, true)
    PYSDM_FUSED_CASE(false, true, false)
    PYSDM_FUSED_CASE(false, true, true)
    PYSDM_FUSED_CASE(true, false, false)
    PYSDM_FUSED_CASE(true, false, true)
    PYSDM_FUSED_CASE(true, true, false)
    PYSDM_FUSED_CASE(true, true, true)
#undef PYSDM_FUSED_CASE
    return (int)cudaErrorInvalidValue;
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

PYSDM_EXPORT int pysdm_b200_fused_collision_step(const pysdm_b200_fused_args *host_args,
                                                 void *stream)
{
    FusedArgs a = *host_args;
    if (a.n_sd <= 0 || a.n_sd >= ((i64)1 << 32) || a.n_cell <= 0) return (int)cudaErrorInvalidValue;
    if (a.max_cell_size > kCtaPos) return (int)cudaErrorInvalidConfiguration;
    if ((a.u01 == nullptr) != (a.rand == nullptr)) return (int)cudaErrorInvalidValue;
    cudaStream_t st = (cudaStream_t)stream;
    if (a.nf_slot == nullptr) return (int)cudaErrorInvalidValue;
    if (a.refresh_nf) {
        k_fused_norm_factor<<<grid_for(a.n_cell), kThreads, 0, st>>>(
            a.cell_start, a.n_cell, a.norm_factor, a.timestep, a.dv);
        if (a.n_sd / 2 > 0)
            k_fused_nf_slot<<<grid_for(a.n_sd / 2), kThreads, 0, st>>>(
                a.cell_id, a.cell_idx, a.norm_factor, a.nf_slot, a.n_sd / 2);
    }
    const i64 mcs = a.max_cell_size < 1 ? 1 : a.max_cell_size;
    const bool warp = mcs <= kWarpPos;
    i64 cpb = (warp ? kWarpPos : kCtaPos) / mcs;
    const i64 cap = warp ? kWarpCells : kCtaCells;
    if (cpb > cap) cpb = cap;
    a.cells_per_batch = cpb;
    a.reserved = divisor_ok(a.rho_w) ? 1 : 0;
    for (int r = 0; r < 10; ++r) {  // Philox4x32 key schedule (philox.cuh)
        a.philox_round_keys[2 * r] = (u32)a.seed + (u32)r * 0x9E3779B9u;
        a.philox_round_keys[2 * r + 1] = (u32)(a.seed >> 32) + (u32)r * 0xBB67AE85u;
    }
    const i64 n_batches = (a.n_cell + cpb - 1) / cpb;
    const bool philox = a.u01 == nullptr;
    const bool aggregate = a.n_cell < 64;  // few counters: aggregate within the warp first
    switch (a.kernel) {
    case KERNEL_GOLOVIN: return dispatch_fused<KERNEL_GOLOVIN>(a, n_batches, philox, warp, aggregate, st);
    case KERNEL_CONSTANT_K:
        return dispatch_fused<KERNEL_CONSTANT_K>(a, n_batches, philox, warp, aggregate, st);
    case KERNEL_LINEAR: return dispatch_fused<KERNEL_LINEAR>(a, n_batches, philox, warp, aggregate, st);
    case KERNEL_GEOMETRIC:
        if (a.radius == nullptr || a.fall_velocity == nullptr) return (int)cudaErrorInvalidValue;
        return dispatch_fused<KERNEL_GEOMETRIC>(a, n_batches, philox, warp, aggregate, st);
    case KERNEL_HYDRODYNAMIC:
        if (a.radius == nullptr || a.fall_velocity == nullptr) return (int)cudaErrorInvalidValue;
        return dispatch_fused<KERNEL_HYDRODYNAMIC>(a, n_batches, philox, warp, aggregate, st);
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
    const i64 mcs = max_cell_size < 1 ? 1 : max_cell_size;
    const bool warp = mcs <= kWarpPos;
    i64 cpb = (warp ? kWarpPos : kCtaPos) / mcs;
    const i64 cap = warp ? kWarpCells : kCtaCells;
    if (cpb > cap) cpb = cap;
    a.cells_per_batch = cpb;
    const i64 n_batches = (n_cell + cpb - 1) / cpb;
    cudaStream_t st = (cudaStream_t)stream;
    return warp ? launch_fused<KERNEL_CONSTANT_K, false, true, false>(a, n_batches, st)
                : launch_fused<KERNEL_CONSTANT_K, false, false, false>(a, n_batches, st);
}
