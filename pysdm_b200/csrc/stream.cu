// stream.cu -- the collision (sub-)step for populations with cells too large for the shared-memory
// batches of fused_kernel.cuh: single-cell boxes (BASELINE configs[0..2]: 2^17 .. 2^22
// super-droplets in ONE cell) and anything in between.  Same contract and argument block as
// the fused step (pysdm_b200_fused_collision_step dispatches here), same results bit for bit:
//
//   1. shuffle_local: exact parallel replay of every cell's serial swap chain (index.cu;
//      index_methods.py:32-44), out of place into scratch;
//   2. ONE streaming kernel over the pair slots: find_pairs (pair_methods.py:34-55, literally,
//      with its cell_idx / cell_start look-ups), sort_within_pair_by_attr (:126-140), collision
//      kernel, max_pair, normalize (quirk A1), /= substeps, compute_gamma, collision_coalescence
//      (collisions_methods.py:418-585) -- every thread owns four consecutive positions = two
//      pair slots, 16-byte idx loads and stores, one Philox block for its two `rand` values;
//   2'. adaptive time step: the kernel runs twice -- pass 1 leaves the probabilities in `gamma`
//      and reduces dt_optimal per cell (block-level reduction first: a box is ONE cell), a
//      per-cell kernel takes dt_todo / dt_left / stats / adaptive_sdm_end, pass 2 scales,
//      draws gamma and coalesces (scale_prob_for_adaptive_sdm_gamma, :330-405).
//
// The random accesses (multiplicity, mass / radius / velocity of the pair members, over the
// whole cell) are what bounds it: 32-byte sectors for 8-byte values.
#include "fused_kernel.cuh"

int pysdm_shuffle_local_to(const i64 *idx_in, i64 *idx_out, const double *u01,
                           const i64 *cell_start, i64 n_cell, i64 length, void *ws,
                           size_t ws_bytes, cudaStream_t st);
size_t pysdm_shuffle_ws_bytes(i64 length);
extern "C" int pysdm_b200_philox_uniform(double *out, i64 n, u64 seed, u64 call, void *stream);

namespace {

constexpr int kStreamThreads = 256;

struct StreamScratch {
    const i64 *shuffled;  // idx after shuffle_local, [sorted_length]
    double *dt_todo;      // [n_cell] adaptive: min(dt_left, dt_max) then min with dt_optimal
    double *ratio;        // [n_cell] adaptive: dt_todo / dt
};

static inline size_t al(size_t x) { return (x + 255) / 256 * 256; }

// positive doubles order like their bit patterns (dt_optimal >= dt_min >= 0)
__device__ __forceinline__ void atomic_min_pos(double *addr, double v)
{
    atomicMin(reinterpret_cast<unsigned long long *>(addr),
              (unsigned long long)__double_as_longlong(v));
}

// kPass: 0 = fixed time step (everything in one pass), 1 = adaptive pass 1 (pairs, prob,
// dt_optimal), 2 = adaptive pass 2 (scale, gamma, coalescence; pairs re-read as written by pass 1)
template <int K, bool kPhilox, int kPass, bool kOneCell>
__global__ void __launch_bounds__(kStreamThreads) k_stream_step(const FusedArgs a, const StreamScratch sc)
{
    constexpr bool kGravitational = K == KERNEL_GEOMETRIC || K == KERNEL_HYDRODYNAMIC;
    const i64 t = (i64)blockIdx.x * kStreamThreads + threadIdx.x;
    const i64 length = a.length, sorted_length = a.sorted_length > 0 ? a.sorted_length : a.length;
    const i64 n_slots = a.n_sd / 2, n_active = length / 2;
    const i64 p0 = 4 * t;  // my positions p0 .. p0 + 3, slots 2t, 2t + 1
    const double prob_nopair =
        0.0 * (kGravitational ? 0.0 : kernel_temp_of<K>(0.0, a.kernel_a, a.kernel_b));
    const bool unit_substeps = a.substeps == 1.0;
    const double rho_rcp = 1.0 / a.rho_w;
    const bool rho_ok = a.reserved != 0;
    const double *mass = kGravitational ? a.radius : a.attributes + a.mass_row * a.n_sd;
    const i64 *src = kPass == 2 ? a.idx_out : sc.shuffled;  // pass 2: the pairs pass 1 sorted

    // ---- positions, cells, find_pairs ---------------------------------------------------
    // s[q], c[q], f[q] for positions p0 - 1 + q, q = 0 .. 5 (my four + both neighbours)
    i64 s[6], c[6];
    bool f[6];
#pragma unroll
    for (int q = 0; q < 6; ++q) {
        const i64 p = p0 - 1 + q;
        s[q] = (p >= 0 && p < sorted_length) ? src[p] : 0;
        c[q] = 0;
    }
    if (!kOneCell) {
#pragma unroll
        for (int q = 0; q < 6; ++q) {
            const i64 p = p0 - 1 + q;
            if (p >= 0 && p < length) c[q] = a.cell_id[s[q]];
        }
    }
    if (kPass == 2) {  // the flags pass 1 wrote
#pragma unroll
        for (int q = 0; q < 5; ++q) {
            const i64 p = p0 - 1 + q;
            f[q] = p >= 0 && p < length && a.is_first_in_pair[p] != 0;
        }
    } else {
#pragma unroll
        for (int q = 0; q < 5; ++q) {
            const i64 p = p0 - 1 + q;
            bool first = false;
            if (p >= 0 && p + 1 < length && c[q] == c[q + 1]) {  // (flag[length - 1] = False)
                const i64 ref = kOneCell ? a.cell_start[a.cell_idx[0]]
                                         : a.cell_start[a.cell_idx[c[q]]];
                first = ((p - ref) & 1) == 0;  // Python's % 2 == 0, also for negative offsets
            }
            f[q] = first;
        }
    }
    f[5] = false;

    // ---- flags and unpaired positions ------------------------------------------------------
    if (kPass != 2) {
#pragma unroll
        for (int q = 1; q <= 4; ++q) {
            const i64 p = p0 - 1 + q;
            if (p < length) a.is_first_in_pair[p] = (u8)f[q];
            // not in any pair (neither first nor the second of the pair before): idx goes through
            if (p < sorted_length && !f[q] && !f[q - 1]) a.idx_out[p] = s[q];
        }
    }

    // ---- my two slots -------------------------------------------------------------------------
    double rnd[2] = {0.0, 0.0};
    if (kPass != 1 && p0 < 2 * n_active + 4) {
        if (kPhilox) {  // block t holds the `rand` elements 2t, 2t + 1
            u32 w[4];
            philox4x32_10_rk((u32)t, (u32)((u64)t >> 32), (u32)a.call_rand,
                             (u32)(a.call_rand >> 32), a.philox_round_keys, w);
            rnd[0] = u01_from_words(w[0], w[1]);
            rnd[1] = u01_from_words(w[2], w[3]);
        } else {
            if (2 * t < n_active) rnd[0] = a.rand[2 * t];
            if (2 * t + 1 < n_active) rnd[1] = a.rand[2 * t + 1];
        }
    }
    i64 cid[2] = {0, 0}, add_rate[2] = {0, 0}, add_deficit[2] = {0, 0}, add_coal[2] = {0, 0};
    bool counted[2] = {false, false};
    double dt_opt[2] = {0.0, 0.0};
    bool candidate[2] = {false, false};
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        const i64 d = 2 * t + h;
        if (d >= n_slots) continue;
        if (d >= n_active) {  // beyond the working length: `0 * norm_factor`, no compute_gamma
            if (kPass != 2) {
                double pr = prob_nopair * a.nf_slot[d];
                if (!unit_substeps) pr = pr / a.substeps;
                a.gamma[d] = pr;
            }
            continue;
        }
        const int q0 = 1 + 2 * h;  // s[] index of the slot's first position
        const bool has = f[q0] || f[q0 + 1];
        const bool at0 = f[q0];  // the pair starts at the slot's first position, else its second
        const double nf = (kPass == 2) ? 0.0 : a.nf_slot[d];
        if (!has) {
            if (kPass == 1) {
                a.gamma[d] = prob_nopair * nf;
            } else {
                double pr = kPass == 2 ? a.gamma[d] : prob_nopair * nf;
                if (kPass == 0 && !unit_substeps) pr = pr / a.substeps;
                a.gamma[d] = ceil_exact(pr - rnd[h]);
            }
            continue;
        }
        u32 sj = (u32)(at0 ? s[q0] : s[q0 + 1]), sk = (u32)(at0 ? s[q0 + 1] : s[q0 + 2]);
        i64 n_j = 0, n_k = 0;
        double prob;
        if (kPass != 2) {
            n_j = a.multiplicity[sj];
            n_k = a.multiplicity[sk];
            double m_j = 0.0, m_k = 0.0, u_j = 0.0, u_k = 0.0;
            if (K != KERNEL_CONSTANT_K) {
                m_j = mass[sj];
                m_k = mass[sk];
            }
            if (kGravitational) {
                u_j = a.fall_velocity[sj];
                u_k = a.fall_velocity[sk];
            }
            if (n_j < n_k) {  // sort_within_pair_by_attr (ties keep order)
                const u32 ts = sj; sj = sk; sk = ts;
                const i64 tn = n_j; n_j = n_k; n_k = tn;
                const double tm = m_j; m_j = m_k; m_k = tm;
            }
            const i64 pf = p0 - 1 + q0 + (at0 ? 0 : 1);
            a.idx_out[pf] = (i64)sj;
            a.idx_out[pf + 1] = (i64)sk;
            double kernel_temp;
            if (K == KERNEL_GEOMETRIC) {
                kernel_temp = geometric_kernel_temp(m_j, m_k, u_j, u_k, a.kernel_a);
            } else if (K == KERNEL_HYDRODYNAMIC) {
                kernel_temp = hydrodynamic_kernel_temp(a, m_j, m_k, u_j, u_k);
            } else {
                double vsum = 0.0;
                if (K != KERNEL_CONSTANT_K)
                    vsum = div_by(m_j, a.rho_w, rho_rcp, rho_ok) + div_by(m_k, a.rho_w, rho_rcp, rho_ok);
                kernel_temp = kernel_temp_of<K>(vsum, a.kernel_a, a.kernel_b);
            }
            prob = i64_to_f64(n_j) * kernel_temp;
            prob = prob * nf;
        } else {
            prob = a.gamma[d];
        }
        cid[h] = kOneCell ? 0 : (at0 ? c[q0] : c[q0 + 1]);
        if (kPass == 1) {
            a.gamma[d] = prob;
            if (prob != 0.0) {  // pair_indices: prob == 0 skips the pair (:360-362)
                const i64 prop = floordiv_sorted(n_j, n_k);
                double o = a.timestep * i64_to_f64(prop) / prob;
                o = py_max(o, a.dt_min);
                dt_opt[h] = o;
                candidate[h] = o == o;
            }
            continue;
        }
        if (kPass == 0) {
            if (!unit_substeps) prob = prob / a.substeps;
        } else if (prob != 0.0) {
            prob = prob * sc.ratio[cid[h]];
        }
        const double g_raw = ceil_exact(prob - rnd[h]);
        a.gamma[d] = g_raw;
        if (g_raw != 0.0) {
            if (kPass == 2) {
                n_j = a.multiplicity[sj];
                n_k = a.multiplicity[sk];
            }
            const HitResult r = hit_pair<K>(a, sj, sk, n_j, n_k, g_raw, a.gamma + d);
            add_rate[h] = r.add_rate;
            add_deficit[h] = r.add_deficit;
            add_coal[h] = r.add_coal;
            counted[h] = true;
        }
    }

    // ---- per-cell reductions (whole warps arrive here: no early returns above) --------------------
    if constexpr (kPass == 1) {
        if (kOneCell) {  // one block-level minimum, one atomic per block
            __shared__ unsigned long long block_min;
            if (threadIdx.x == 0) block_min = 0x7ff0000000000000ull;
            __syncthreads();
            double m = __longlong_as_double(0x7ff0000000000000ll);
            if (candidate[0]) m = dt_opt[0];
            if (candidate[1] && dt_opt[1] < m) m = dt_opt[1];
            u32 hi = (u32)__double2hiint(m);
            const u32 mhi = __reduce_min_sync(0xffffffffu, hi);
            const u32 lo = hi == mhi ? (u32)__double2loint(m) : 0xffffffffu;
            const u32 mlo = __reduce_min_sync(0xffffffffu, lo);
            if ((threadIdx.x & 31) == 0)
                atomicMin(&block_min, ((unsigned long long)mhi << 32) | mlo);
            __syncthreads();
            if (threadIdx.x == 0 && block_min != 0x7ff0000000000000ull)
                atomicMin(reinterpret_cast<unsigned long long *>(sc.dt_todo), block_min);
        } else {
#pragma unroll
            for (int h = 0; h < 2; ++h)
                if (candidate[h]) atomic_min_pos(sc.dt_todo + a.n_cell + cid[h], dt_opt[h]);
        }
    } else {
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            cell_counter_add(a.collision_rate, cid[h], add_rate[h], counted[h]);
            cell_counter_add(a.collision_rate_deficit, cid[h], add_deficit[h], counted[h]);
            cell_counter_add(a.coalescence_rate, cid[h], add_coal[h], counted[h]);
        }
    }
}

// adaptive, before pass 1: the pairs' minimum starts at +inf (kept apart from dt_left so that
// stats_dt_min sees the pairs only): pair_min[c] = sc.dt_todo[n_cell + c]; one-cell boxes use
// sc.dt_todo[0] for it
__global__ void k_stream_adaptive_init(double *pair_min, i64 n)
{
    i64 c = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (c < n) pair_min[c] = __longlong_as_double(0x7ff0000000000000ll);
}

// adaptive, between the passes, per cell id: dt_todo, dt_left, stats, adaptive_sdm_end
__global__ void k_stream_adaptive_cells(const FusedArgs a, const double *pair_min, double *ratio)
{
    i64 cid = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    u32 my_end = 0;
    if (cid < a.n_cell) {
        const double pm = pair_min[cid];
        const double left = a.dt_left[cid];
        double todo = py_min(left, a.dt_max);
        todo = pm < todo ? pm : todo;
        const double cur = a.stats_dt_min[cid];
        if (pm < cur) a.stats_dt_min[cid] = pm;
        ratio[cid] = todo / a.timestep;
        const double new_left = left - todo;
        a.dt_left[cid] = new_left;
        if (todo > 0) a.stats_n_substep[cid] += 1;
        if (a.adaptive_end != nullptr && new_left != 0) my_end = (u32)a.cell_start[cid + 1];
    }
    if (a.adaptive_end != nullptr) {
        const u32 w_end = __reduce_max_sync(0xffffffffu, my_end);
        if ((threadIdx.x & 31) == 0 && w_end != 0)
            atomicMax(reinterpret_cast<unsigned long long *>(a.adaptive_end),
                      (unsigned long long)w_end);
    }
}

template <int K, bool kPhilox, bool kOneCell>
int launch_stream(const FusedArgs &a, const StreamScratch &sc, double *pair_min, cudaStream_t st)
{
    const i64 sorted_length = a.sorted_length > 0 ? a.sorted_length : a.length;
    const i64 n_pos = sorted_length > a.n_sd / 2 * 2 ? sorted_length : a.n_sd / 2 * 2;
    const unsigned grid = (unsigned)(((n_pos + 3) / 4 + kStreamThreads - 1) / kStreamThreads);
    if (!a.adaptive) {
        k_stream_step<K, kPhilox, 0, kOneCell><<<grid ? grid : 1, kStreamThreads, 0, st>>>(a, sc);
        return LAUNCH_CHECK();
    }
    k_stream_adaptive_init<<<grid_for(a.n_cell), kThreads, 0, st>>>(pair_min, a.n_cell);
    k_stream_step<K, kPhilox, 1, kOneCell><<<grid ? grid : 1, kStreamThreads, 0, st>>>(a, sc);
    k_stream_adaptive_cells<<<grid_for(a.n_cell), kThreads, 0, st>>>(a, pair_min, sc.ratio);
    k_stream_step<K, kPhilox, 2, kOneCell><<<grid ? grid : 1, kStreamThreads, 0, st>>>(a, sc);
    return LAUNCH_CHECK();
}

template <int K>
int dispatch_stream(const FusedArgs &a, const StreamScratch &sc, double *pair_min, bool philox,
                    bool one_cell, cudaStream_t st)
{
    if (philox)
        return one_cell ? launch_stream<K, true, true>(a, sc, pair_min, st)
                        : launch_stream<K, true, false>(a, sc, pair_min, st);
    return one_cell ? launch_stream<K, false, true>(a, sc, pair_min, st)
                    : launch_stream<K, false, false>(a, sc, pair_min, st);
}

}  // namespace

size_t pysdm_stream_ws_bytes(i64 n_sd, i64 n_cell)
{
    return al(sizeof(double) * (size_t)n_sd)        // u01 when generated by Philox
           + al(sizeof(i64) * (size_t)n_sd)         // shuffled idx
           + al(pysdm_shuffle_ws_bytes(n_sd))       // shuffle replay
           + 3 * al(sizeof(double) * (size_t)(n_cell + 1));  // dt_todo / pair_min, ratio
}

// called by pysdm_b200_fused_collision_step (fused.cu) after the norm_factor / nf_slot pre-pass,
// with the launcher-filled fields of `a` (round keys, divisor check) in place
int pysdm_stream_collision_step(const FusedArgs &a, cudaStream_t st)
{
    if (a.ws == nullptr || a.ws_bytes < pysdm_stream_ws_bytes(a.n_sd, a.n_cell))
        return (int)cudaErrorInvalidConfiguration;
    const i64 sorted_length = a.sorted_length > 0 ? a.sorted_length : a.length;
    if (sorted_length > a.n_sd || a.length > sorted_length) return (int)cudaErrorInvalidValue;
    char *w = (char *)a.ws;
    double *u01_ws = (double *)w;
    w += al(sizeof(double) * (size_t)a.n_sd);
    i64 *shuffled = (i64 *)w;
    w += al(sizeof(i64) * (size_t)a.n_sd);
    void *shuffle_ws = w;
    const size_t shuffle_bytes = al(pysdm_shuffle_ws_bytes(a.n_sd));
    w += shuffle_bytes;
    double *pair_min = (double *)w;
    w += 2 * al(sizeof(double) * (size_t)(a.n_cell + 1));
    double *ratio = (double *)w;
    const bool philox = a.u01 == nullptr;
    const double *u01 = a.u01;
    if (philox) {  // the shuffle's uniforms, as the Random kernel would have written them
        int err = pysdm_b200_philox_uniform(u01_ws, sorted_length, a.seed, a.call_u01, st);
        if (err) return err;
        u01 = u01_ws;
    }
    int err = pysdm_shuffle_local_to(a.idx_in, shuffled, u01, a.cell_start, a.n_cell, sorted_length,
                                     shuffle_ws, shuffle_bytes, st);
    if (err) return err;
    StreamScratch sc;
    sc.shuffled = shuffled;
    sc.dt_todo = pair_min - a.n_cell;  // k_stream_step addresses pair_min as dt_todo[n_cell + c]
    sc.ratio = ratio;
    const bool one_cell = a.n_cell == 1;
    if (one_cell) sc.dt_todo = pair_min;
    switch (a.kernel) {
    case KERNEL_GOLOVIN: return dispatch_stream<KERNEL_GOLOVIN>(a, sc, pair_min, philox, one_cell, st);
    case KERNEL_CONSTANT_K: return dispatch_stream<KERNEL_CONSTANT_K>(a, sc, pair_min, philox, one_cell, st);
    case KERNEL_LINEAR: return dispatch_stream<KERNEL_LINEAR>(a, sc, pair_min, philox, one_cell, st);
    case KERNEL_GEOMETRIC: return dispatch_stream<KERNEL_GEOMETRIC>(a, sc, pair_min, philox, one_cell, st);
    case KERNEL_HYDRODYNAMIC:
        return dispatch_stream<KERNEL_HYDRODYNAMIC>(a, sc, pair_min, philox, one_cell, st);
    default: return (int)cudaErrorInvalidValue;
    }
}
