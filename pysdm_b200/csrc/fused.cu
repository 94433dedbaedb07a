// fused.cu -- one-launch SDM collision step for cell-sorted populations (sm_100a).
//
// Executes, for every cell of a population whose `idx` is sorted by cell
// (ParticleAttributes.cell_start valid), literally the same sequence the reference runs
// as ~15 separate backend calls in Collision.step (PySDM/dynamics/collisions/collision.py:196-234):
//
//   shuffle_local            impl_numba/methods/index_methods.py:32-44
//   find_pairs               impl_numba/methods/pair_methods.py:34-55
//   sort_within_pair_by_attr pair_methods.py:126-140          (attr = multiplicity)
//   collision kernel         collision_kernels/{golovin,constantK,linear}.py  (sum_pair, *=, +=)
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
// Data movement (DESIGN.md "fused step"): a CTA owns a batch of consecutive cells whose
// SDs occupy one contiguous range of positions.  The range of `idx` and of the uniforms
// is streamed in with 16-byte loads and staged in shared memory as 32-bit SD numbers +
// 16-bit swap targets; one thread per cell replays the cell's swap chain in shared
// memory; then one thread per pair slot gathers multiplicity / mass (the only random
// accesses, confined to the cell's own SDs), evaluates probability and gamma and
// scatters the updates.  Nothing is written to HBM between those stages.
#include "philox.cuh"
#include "pysdm_b200.h"
#include "sdm_device.cuh"

constexpr int kFusedThreads = 256;
constexpr int kPosMax = 8192;   // positions (SDs) per batch
constexpr int kCellMax = 256;   // cells per batch (one chain thread per cell)

// C-ABI argument block: include/pysdm_b200.h (mirrored by ctypes in pysdm_b200/_lib.py)
typedef pysdm_b200_fused_args FusedArgs;

enum { KERNEL_GOLOVIN = 0, KERNEL_CONSTANT_K = 1, KERNEL_LINEAR = 2 };

// kernel_temp of a pair from the pair's volume sum (sum_pair result; 0 for non-pairs)
template <int K>
__device__ __forceinline__ double kernel_temp_of(double vsum, double a, double b)
{
    if (K == KERNEL_GOLOVIN) return vsum * b;   // golovin.py:14-16
    if (K == KERNEL_CONSTANT_K) return a;       // constantK.py: output.fill(a)
    return vsum * b + a;                        // linear.py: sum; *= b; += a  (no FMA: -fmad=false)
}

__device__ __forceinline__ double philox_element(u64 seed, u64 call, i64 e)
{
    u32 w[4];
    u64 b = (u64)e >> 1;
    philox4x32_10((u32)b, (u32)(b >> 32), (u32)call, (u32)(call >> 32), (u32)seed,
                  (u32)(seed >> 32), w);
    return (e & 1) ? u01_from_words(w[2], w[3]) : u01_from_words(w[0], w[1]);
}

// norm_factor of the cell with rank r (normalize, collisions_methods.py:639-646)
__device__ __forceinline__ double norm_factor_of_rank(const i64 *__restrict__ cell_start, i64 r,
                                                      double timestep, double dv)
{
    return sdm_norm_factor(cell_start[r + 1] - cell_start[r], timestep, dv);
}

// shared-memory carve-up (bytes, 16-byte aligned sections)
constexpr int kSdBytes = ((kPosMax + 2) * 4 + 15) / 16 * 16;
constexpr int kJlBytes = ((kPosMax + 2) * 2 + 15) / 16 * 16;
constexpr int kCoBytes = ((kPosMax + 2) * 1 + 15) / 16 * 16;
constexpr int kCsBytes = ((kCellMax + 1) * 4 + 15) / 16 * 16;
constexpr int kCidBytes = kCellMax * 8;
constexpr int kFusedSmem = kSdBytes + kJlBytes + kCoBytes + kCsBytes + kCidBytes;

template <int K, bool kPhilox>
__global__ void __launch_bounds__(kFusedThreads, 3) k_fused_step(const FusedArgs a)
{
    extern __shared__ __align__(16) unsigned char smem[];
    u32 *sd = reinterpret_cast<u32 *>(smem);
    u16 *jl = reinterpret_cast<u16 *>(smem + kSdBytes);
    u8 *cellof = smem + kSdBytes + kJlBytes;
    i32 *cs = reinterpret_cast<i32 *>(smem + kSdBytes + kJlBytes + kCoBytes);
    i64 *cidv = reinterpret_cast<i64 *>(smem + kSdBytes + kJlBytes + kCoBytes + kCsBytes);

    const int tid = threadIdx.x;
    const i64 n_pairs_active = a.length / 2;
    // literal chain for a slot without a pair: max_pair -> 0; *= kernel_temp; normalize
    const double prob_nopair = 0.0 * kernel_temp_of<K>(0.0, a.kernel_a, a.kernel_b);

    // ---- tail: pair slots beyond the working length keep `0 * norm_factor` ------------
    for (i64 d = n_pairs_active + (i64)blockIdx.x * kFusedThreads + tid; d < a.n_sd / 2;
         d += (i64)gridDim.x * kFusedThreads) {
        double nf = norm_factor_of_rank(a.cell_start, a.cell_idx[a.cell_id[d]], a.timestep, a.dv);
        a.gamma[d] = prob_nopair * nf / a.substeps;
    }

    const i64 c0 = (i64)blockIdx.x * a.cells_per_batch;
    if (c0 >= a.n_cell) return;
    const i64 c1 = (c0 + a.cells_per_batch < a.n_cell) ? c0 + a.cells_per_batch : a.n_cell;
    const int ncb = (int)(c1 - c0);
    const i64 base = a.cell_start[c0], end = a.cell_start[c1];
    if (end - base > kPosMax || ncb > kCellMax) {  // caller's max_cell_size hint was wrong
        if (tid == 0) a.status[0] = 1;
        return;
    }
    const i64 base_al = base & ~(i64)1;  // local position 0 <-> an even global position
    const int lo = (int)(base - base_al);
    const int P = (int)(end - base_al);

    // ---- phase 0: cell boundaries ---------------------------------------------------------
    for (int c = tid; c <= ncb; c += kFusedThreads) cs[c] = (i32)(a.cell_start[c0 + c] - base_al);
    __syncthreads();

    // ---- phase 1: position -> cell map; per-cell norm_factor (API-visible) ----------------
    {
        const int warp = tid >> 5, lane = tid & 31;
        for (int c = warp; c < ncb; c += kFusedThreads / 32) {
            const int s = cs[c], e = cs[c + 1];
            for (int lp = s + lane; lp < e; lp += 32) cellof[lp] = (u8)c;
        }
        for (int c = tid; c < ncb; c += kFusedThreads)
            a.norm_factor[c0 + c] = sdm_norm_factor((i64)(cs[c + 1] - cs[c]), a.timestep, a.dv);
    }
    __syncthreads();

    // ---- phase 2: stream idx + uniforms in, stage SD numbers and swap targets -------------
    {
        const bool vec_idx = (((uintptr_t)a.idx_in) & 15) == 0;
        const bool vec_u = !kPhilox && (((uintptr_t)a.u01) & 15) == 0;
#pragma unroll 4
        for (int lp0 = 2 * tid; lp0 < P; lp0 += 2 * kFusedThreads) {
            const i64 p0 = base_al + lp0;
            const bool v0 = lp0 >= lo, v1 = lp0 + 1 < P;
            i64 s0 = 0, s1 = 0;
            double u0 = 0.0, u1 = 0.0;
            if (v0 && v1 && vec_idx) {
                longlong2 t = *reinterpret_cast<const longlong2 *>(a.idx_in + p0);
                s0 = t.x;
                s1 = t.y;
            } else {
                if (v0) s0 = a.idx_in[p0];
                if (v1) s1 = a.idx_in[p0 + 1];
            }
            if (kPhilox) {
                u32 w[4];
                u64 b = (u64)p0 >> 1;
                philox4x32_10((u32)b, (u32)(b >> 32), (u32)a.call_u01, (u32)(a.call_u01 >> 32),
                              (u32)a.seed, (u32)(a.seed >> 32), w);
                u0 = u01_from_words(w[0], w[1]);
                u1 = u01_from_words(w[2], w[3]);
            } else if (v0 && v1 && vec_u) {
                double2 t = *reinterpret_cast<const double2 *>(a.u01 + p0);
                u0 = t.x;
                u1 = t.y;
            } else {
                if (v0) u0 = a.u01[p0];
                if (v1) u1 = a.u01[p0 + 1];
            }
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const int lp = lp0 + h;
                if (h == 0 ? !v0 : !v1) continue;
                const int c = cellof[lp];
                const int s = cs[c], len = cs[c + 1] - s;
                const i64 S = base_al + s;
                // index_methods.py:40: j = int(cell_start[c] + u01[i] * (cell_end - cell_start))
                i64 j = (i64)((double)S + (h == 0 ? u0 : u1) * (double)len);
                if (j >= S + len) j = S + len - 1;  // fp round-up at huge S (see index.cu)
                jl[lp] = (u16)(j - S);
                sd[lp] = (u32)(h == 0 ? s0 : s1);
            }
        }
    }
    __syncthreads();

    // ---- phase 3: one thread per cell replays the swap chain in shared memory -------------
    for (int c = tid; c < ncb; c += kFusedThreads) {
        const int s = cs[c], e = cs[c + 1];
        i64 cid = 0;
        if (e > s) cid = a.cell_id[sd[s]];  // issued now, consumed after the chain
        for (int i = e - 1; i > s; --i) {
            const int j = s + jl[i];
            const u32 x = sd[i], y = sd[j];
            sd[i] = y;
            sd[j] = x;
        }
        cidv[c] = cid;
    }
    __syncthreads();

    // ---- phase 4: one thread per pair slot (local window = positions lp0, lp0 + 1) --------
    const double *mass = a.attributes + a.mass_row * a.n_sd;
    const bool flag_al = (((uintptr_t)a.is_first_in_pair) & 1) == 0;
    const int n_iter = (P + 2 * kFusedThreads - 1) / (2 * kFusedThreads);
    for (int it = 0; it < n_iter; ++it) {  // uniform trip count: warp-collective counter adds
        const int lp0 = 2 * (it * kFusedThreads + tid), lp1 = lp0 + 1;
        const i64 p0 = base_al + lp0;
        const bool v0 = lp0 >= lo && lp0 < P, v1 = lp1 < P;
        int cA = 0, cB = 0;
        bool first0 = false, first1 = false, single0 = false, single1 = false;
        if (v0) {
            cA = cellof[lp0];
            const bool even = ((lp0 - cs[cA]) & 1) == 0, partner = lp0 + 1 < cs[cA + 1];
            first0 = even && partner;
            single0 = even && !partner;
        }
        if (v1) {
            cB = cellof[lp1];
            const bool even = ((lp1 - cs[cB]) & 1) == 0, partner = lp1 + 1 < cs[cB + 1];
            first1 = even && partner;
            single1 = even && !partner;
        }
        // find_pairs output (pair_methods.py:38-41, :55)
        if (v0 && v1 && flag_al) {
            *reinterpret_cast<u16 *>(a.is_first_in_pair + p0) =
                (u16)((first0 ? 1u : 0u) | (first1 ? 0x100u : 0u));
        } else {
            if (v0) a.is_first_in_pair[p0] = (u8)first0;
            if (v1) a.is_first_in_pair[p0 + 1] = (u8)first1;
        }
        if (single0) a.idx_out[p0] = (i64)sd[lp0];
        if (single1) a.idx_out[p0 + 1] = (i64)sd[lp1];

        const bool has = first0 || first1;
        const i64 d = p0 >> 1;  // global pair slot of this window
        bool active = false;
        i64 cid = 0, add_rate = 0, add_deficit = 0, add_coal = 0;
        if (has) {
            const int pf = first0 ? lp0 : lp1;
            const int c = first0 ? cA : cB;
            i64 j = (i64)sd[pf], k = (i64)sd[pf + 1];
            i64 nj = a.multiplicity[j], nk = a.multiplicity[k];
            double mj = 0.0, mk = 0.0;
            if (K != KERNEL_CONSTANT_K) {
                mj = mass[j];
                mk = mass[k];
            }
            const double r = kPhilox ? philox_element(a.seed, a.call_rand, d) : a.rand[d];
            const double nf =
                norm_factor_of_rank(a.cell_start, a.cell_idx[a.cell_id[d]], a.timestep, a.dv);
            if (nj < nk) {  // sort_within_pair_by_attr (ties keep order)
                i64 t = j; j = k; k = t;
                t = nj; nj = nk; nk = t;
                double tm = mj; mj = mk; mk = tm;
            }
            const i64 pg = base_al + pf;
            a.idx_out[pg] = j;
            a.idx_out[pg + 1] = k;
            // volume = mass / rho_w (physics_methods.py:107-119); sum_pair; kernel; max_pair
            double vsum = 0.0;
            if (K != KERNEL_CONSTANT_K) vsum = mj / a.rho_w + mk / a.rho_w;
            double prob = (double)nj * kernel_temp_of<K>(vsum, a.kernel_a, a.kernel_b);
            prob = prob * nf;
            prob = prob / a.substeps;
            const double g_raw = ceil(prob - r);
            if (g_raw == 0.0) {
                a.gamma[d] = g_raw;  // keeps the sign of zero the reference stores
            } else {
                const GammaResult g = sdm_gamma(g_raw, nj, nk);
                a.gamma[d] = g.gamma;
                cid = cidv[c];
                add_rate = g.rate_add;
                add_deficit = g.deficit_add;
                active = true;
                if (g.gamma != 0.0) {
                    // coalesce, collisions_methods.py:44-59 (float arithmetic stored to int64)
                    add_coal = (i64)(g.gamma * (double)nk);
                    const double new_n = (double)nj - g.gamma * (double)nk;
                    i64 nj_new, nk_new = nk;
                    if (new_n > 0) {
                        nj_new = (i64)new_n;
                        a.multiplicity[j] = nj_new;
                        for (i64 r_ = 0; r_ < a.n_attr; ++r_) {
                            double *row = a.attributes + r_ * a.n_sd;
                            const double aj = (r_ == a.mass_row && K != KERNEL_CONSTANT_K) ? mj : row[j];
                            const double ak = (r_ == a.mass_row && K != KERNEL_CONSTANT_K) ? mk : row[k];
                            row[k] = fma(g.gamma, aj, ak);
                        }
                    } else {
                        nj_new = floordiv(nk, 2);
                        nk_new = nk - nj_new;
                        a.multiplicity[j] = nj_new;
                        a.multiplicity[k] = nk_new;
                        for (i64 r_ = 0; r_ < a.n_attr; ++r_) {
                            double *row = a.attributes + r_ * a.n_sd;
                            const double aj = (r_ == a.mass_row && K != KERNEL_CONSTANT_K) ? mj : row[j];
                            const double ak = (r_ == a.mass_row && K != KERNEL_CONSTANT_K) ? mk : row[k];
                            const double v = fma(g.gamma, aj, ak);
                            row[j] = v;
                            row[k] = v;
                        }
                    }
                    if (nj_new == 0 || nk_new == 0) a.healthy[0] = 0;
                }
            }
        } else if (v0 && d < n_pairs_active) {
            // no pair starts inside this window; the slot still goes through compute_gamma.
            // A pair starting at p0 + 1 in the NEXT batch owns the slot instead.
            bool next_first = false;
            if (!v1) {  // p0 is this batch's last position and p0 + 1 < length
                i64 rr = c1;
                while (rr < a.n_cell && a.cell_start[rr + 1] == end) ++rr;
                next_first = rr < a.n_cell && a.cell_start[rr + 1] - end >= 2;
            }
            if (!next_first) {
                const double r = kPhilox ? philox_element(a.seed, a.call_rand, d) : a.rand[d];
                const double nf =
                    norm_factor_of_rank(a.cell_start, a.cell_idx[a.cell_id[d]], a.timestep, a.dv);
                a.gamma[d] = ceil(prob_nopair * nf / a.substeps - r);
            }
        }
        cell_counter_add(a.collision_rate, cid, add_rate, active);
        cell_counter_add(a.collision_rate_deficit, cid, add_deficit, active);
        cell_counter_add(a.coalescence_rate, cid, add_coal, active);
    }
}

// ---- launcher -----------------------------------------------------------------------------
PYSDM_EXPORT int64_t pysdm_b200_fused_max_cell_size(void) { return kPosMax; }

template <int K, bool kPhilox>
static int launch_fused(const FusedArgs &a, i64 n_batches, cudaStream_t st)
{
    static bool configured = false;  // one caller thread per process (SURVEY 8b "Threading")
    if (!configured) {
        int err = (int)cudaFuncSetAttribute(k_fused_step<K, kPhilox>,
                                            cudaFuncAttributeMaxDynamicSharedMemorySize,
                                            kFusedSmem);
        if (err) return err;
        configured = true;
    }
    k_fused_step<K, kPhilox><<<(unsigned)n_batches, kFusedThreads, kFusedSmem, st>>>(a);
    return LAUNCH_CHECK();
}

PYSDM_EXPORT int pysdm_b200_fused_collision_step(const pysdm_b200_fused_args *host_args,
                                                 void *stream)
{
    FusedArgs a = *host_args;
    if (a.n_sd <= 0 || a.n_sd >= ((i64)1 << 32) || a.n_cell <= 0) return (int)cudaErrorInvalidValue;
    if (a.max_cell_size > kPosMax) return (int)cudaErrorInvalidConfiguration;
    if ((a.u01 == nullptr) != (a.rand == nullptr)) return (int)cudaErrorInvalidValue;
    i64 mcs = a.max_cell_size < 1 ? 1 : a.max_cell_size;
    i64 cpb = kPosMax / mcs;
    if (cpb > kCellMax) cpb = kCellMax;
    a.cells_per_batch = cpb;
    i64 n_batches = (a.n_cell + cpb - 1) / cpb;
    cudaStream_t st = (cudaStream_t)stream;
    const bool philox = a.u01 == nullptr;
    switch (a.kernel) {
    case KERNEL_GOLOVIN:
        return philox ? launch_fused<KERNEL_GOLOVIN, true>(a, n_batches, st)
                      : launch_fused<KERNEL_GOLOVIN, false>(a, n_batches, st);
    case KERNEL_CONSTANT_K:
        return philox ? launch_fused<KERNEL_CONSTANT_K, true>(a, n_batches, st)
                      : launch_fused<KERNEL_CONSTANT_K, false>(a, n_batches, st);
    case KERNEL_LINEAR:
        return philox ? launch_fused<KERNEL_LINEAR, true>(a, n_batches, st)
                      : launch_fused<KERNEL_LINEAR, false>(a, n_batches, st);
    default: return (int)cudaErrorInvalidValue;
    }
}
