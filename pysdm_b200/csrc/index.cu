// index.cu -- permutation bookkeeping of the SDM collision step on B200:
// identity_index, shuffle_global / shuffle_local (exact parallel replay of the
// reference's sequential swap chains), stable sort by cell, sort_by_key,
// remove_zero_n_or_flagged, cell_id.  C-ABI declared in include/pysdm_b200.h.
#include "primitives.cuh"
#include "pysdm_b200.h"
#include "sdm_device.cuh"

#include <stdio.h>

// ---------------------------------------------------------------------------------
// identity_index  (reference: impl_numba/methods/index_methods.py:13-20)
// ---------------------------------------------------------------------------------
__global__ void k_identity_index(i64 *__restrict__ idx, i64 n)
{
    i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) idx[i] = i;
}

PYSDM_EXPORT int pysdm_b200_identity_index(i64 *idx, i64 n, void *stream)
{
    if (n <= 0) return 0;
    k_identity_index<<<grid_for(n), kThreads, 0, (cudaStream_t)stream>>>(idx, n);
    return LAUNCH_CHECK();
}

// ---------------------------------------------------------------------------------
// shuffle_global / shuffle_local
// (reference: index_methods.py:22-30 serial chain, :32-44 one serial chain per cell)
//
// The reference applies transpositions t_i = (i, j_i) for i = e-1 ... s+1 inside each
// segment [s, e) (global: one segment [0, length), j_i = int(u01[i]*(i+1)); local:
// j_i = int(s + u01[i]*(e-s)), drawn over the WHOLE cell -- quirk A4).  We replay
// the chain exactly but in parallel: the content of slot q after all swaps is found
// by walking the chain backwards in time, i.e. through increasing i: at slot p, the
// next transposition touching p is either its own (i = p) or one that targets it
// (j_i = p); the walk hops to the partner slot and continues.  Every transposition
// is visited by exactly two walks, so the total work is O(n); targets are grouped
// per slot with a histogram + exclusive scan + fill (lists are unordered, the walk
// takes the minimum above its current time).
// ---------------------------------------------------------------------------------
__device__ __forceinline__ void segment_of(i64 i, const i64 *__restrict__ cell_start, i64 n_cell,
                                           i64 &s, i64 &e)
{
    // last c with cell_start[c] <= i  (empty cells share a start: take the last one)
    i64 lo = 0, hi = n_cell;  // invariant: cell_start[lo] <= i < cell_start[hi]
    while (hi - lo > 1) {
        i64 mid = (lo + hi) >> 1;
        if (cell_start[mid] <= i)
            lo = mid;
        else
            hi = mid;
    }
    s = cell_start[lo];
    e = cell_start[lo + 1];
}

template <bool Local>
__global__ void k_shuffle_targets(const double *__restrict__ u01, const i64 *__restrict__ cell_start,
                                  i64 n_cell, i64 length, i32 *__restrict__ tgt,
                                  u32 *__restrict__ cnt)
{
    i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= length) return;
    i64 s = 0, e = length, j;
    if (Local) {
        segment_of(i, cell_start, n_cell, s, e);
        j = (i64)((double)s + u01[i] * (double)(e - s));
        // (double)s + x can round up to e for huge s (reference would index outside the
        // cell, undefined there); keep the draw inside the cell -- see DESIGN.md
        if (j >= e) j = e - 1;
    } else {
        j = (i64)(u01[i] * (double)(i + 1));
        if (j > i) j = i;
    }
    if (i == s || j == i) {
        tgt[i] = -1;  // no transposition at the first slot of a segment; (i,i) is a no-op
    } else {
        tgt[i] = (i32)j;
        atomicAdd(&cnt[j], 1u);
    }
}

__global__ void k_shuffle_fill(const i32 *__restrict__ tgt, const u32 *__restrict__ off,
                               u32 *__restrict__ fill, i32 *__restrict__ list, i64 length)
{
    i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= length) return;
    i32 j = tgt[i];
    if (j >= 0) list[off[j] + atomicAdd(&fill[j], 1u)] = (i32)i;
}

__global__ void k_shuffle_walk(const i64 *__restrict__ idx_in, i64 *__restrict__ idx_out,
                               const i32 *__restrict__ tgt, const u32 *__restrict__ off,
                               const i32 *__restrict__ list, i64 length)
{
    i64 q = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= length) return;
    i32 p = (i32)q;
    i32 c = -1;  // index of the last transposition already undone
    for (;;) {
        i32 best = 0x7fffffff;
        if (p > c && tgt[p] >= 0) best = p;  // own transposition t_p
        u32 b = off[p], en = off[p + 1];
        for (u32 t = b; t < en; ++t) {
            i32 i = list[t];
            if (i > c && i < best) best = i;
        }
        if (best == 0x7fffffff) break;
        p = (best == p) ? tgt[p] : best;
        c = best;
    }
    idx_out[q] = idx_in[p];
}

__global__ void k_copy_i64(i64 *__restrict__ dst, const i64 *__restrict__ src, i64 n)
{
    i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) dst[i] = src[i];
}

static inline size_t align_up(size_t x, size_t a = 256) { return (x + a - 1) / a * a; }

// workspace layout for the shuffles (bytes)
static size_t shuffle_ws_bytes(i64 n)
{
    size_t b = 0;
    b += align_up(sizeof(i32) * (size_t)n);        // tgt
    b += align_up(sizeof(u32) * (size_t)(n + 1));  // cnt -> off
    b += align_up(sizeof(u32) * (size_t)n);        // fill
    b += align_up(sizeof(i32) * (size_t)n);        // list
    b += align_up(sizeof(i64) * (size_t)n);        // idx_out
    b += align_up(sizeof(u32) * scan_scratch_elems<u32>(n + 1));
    return b;
}

template <bool Local>
static int shuffle_impl(i64 *idx, const double *u01, const i64 *cell_start, i64 n_cell, i64 length,
                        void *ws, size_t ws_bytes, cudaStream_t st)
{
    if (length < 2) return 0;
    if (length >= ((i64)1 << 31)) return (int)cudaErrorInvalidValue;
    if (ws_bytes < shuffle_ws_bytes(length)) return (int)cudaErrorMemoryAllocation;
    char *w = (char *)ws;
    i32 *tgt = (i32 *)w;
    w += align_up(sizeof(i32) * (size_t)length);
    u32 *cnt = (u32 *)w;
    w += align_up(sizeof(u32) * (size_t)(length + 1));
    u32 *fill = (u32 *)w;
    w += align_up(sizeof(u32) * (size_t)length);
    i32 *list = (i32 *)w;
    w += align_up(sizeof(i32) * (size_t)length);
    i64 *out = (i64 *)w;
    w += align_up(sizeof(i64) * (size_t)length);
    u32 *scan_ws = (u32 *)w;
    // cnt and fill are adjacent: clear both with one memset
    cudaMemsetAsync(cnt, 0, (size_t)((char *)list - (char *)cnt), st);
    const int g = grid_for(length);
    k_shuffle_targets<Local><<<g, kThreads, 0, st>>>(u01, cell_start, n_cell, length, tgt, cnt);
    device_exclusive_scan<u32>(cnt, cnt, length, 1, scan_ws, st);
    k_shuffle_fill<<<g, kThreads, 0, st>>>(tgt, cnt, fill, list, length);
    k_shuffle_walk<<<g, kThreads, 0, st>>>(idx, out, tgt, cnt, list, length);
    k_copy_i64<<<g, kThreads, 0, st>>>(idx, out, length);
    return LAUNCH_CHECK();
}

PYSDM_EXPORT size_t pysdm_b200_shuffle_workspace_bytes(i64 length) { return shuffle_ws_bytes(length); }

// shuffle_local out of place (idx_out != idx_in): what the streaming collision step (stream.cu)
// starts with; same kernels as above without the copy back.  Workspace: shuffle_ws_bytes(length).
int pysdm_shuffle_local_to(const i64 *idx_in, i64 *idx_out, const double *u01,
                           const i64 *cell_start, i64 n_cell, i64 length, void *ws,
                           size_t ws_bytes, cudaStream_t st)
{
    if (length <= 0) return 0;
    const int g = grid_for(length);
    if (length < 2) {
        k_copy_i64<<<g, kThreads, 0, st>>>(idx_out, idx_in, length);
        return LAUNCH_CHECK();
    }
    if (length >= ((i64)1 << 31)) return (int)cudaErrorInvalidValue;
    if (ws_bytes < shuffle_ws_bytes(length)) return (int)cudaErrorMemoryAllocation;
    char *w = (char *)ws;
    i32 *tgt = (i32 *)w;
    w += align_up(sizeof(i32) * (size_t)length);
    u32 *cnt = (u32 *)w;
    w += align_up(sizeof(u32) * (size_t)(length + 1));
    u32 *fill = (u32 *)w;
    w += align_up(sizeof(u32) * (size_t)length);
    i32 *list = (i32 *)w;
    w += align_up(sizeof(i32) * (size_t)length);
    w += align_up(sizeof(i64) * (size_t)length);  // (the in-place variant's output buffer)
    u32 *scan_ws = (u32 *)w;
    cudaMemsetAsync(cnt, 0, (size_t)((char *)list - (char *)cnt), st);
    k_shuffle_targets<true><<<g, kThreads, 0, st>>>(u01, cell_start, n_cell, length, tgt, cnt);
    device_exclusive_scan<u32>(cnt, cnt, length, 1, scan_ws, st);
    k_shuffle_fill<<<g, kThreads, 0, st>>>(tgt, cnt, fill, list, length);
    k_shuffle_walk<<<g, kThreads, 0, st>>>(idx_in, idx_out, tgt, cnt, list, length);
    return LAUNCH_CHECK();
}
size_t pysdm_shuffle_ws_bytes(i64 length) { return shuffle_ws_bytes(length); }

PYSDM_EXPORT int pysdm_b200_shuffle_global(i64 *idx, i64 length, const double *u01, void *ws,
                                           size_t ws_bytes, void *stream)
{
    return shuffle_impl<false>(idx, u01, nullptr, 0, length, ws, ws_bytes, (cudaStream_t)stream);
}

// `length` = cell_start[n_cell] (the reference walks every cell up to its last start)
PYSDM_EXPORT int pysdm_b200_shuffle_local(i64 *idx, const double *u01, const i64 *cell_start,
                                          i64 n_cell, i64 length, void *ws, size_t ws_bytes,
                                          void *stream)
{
    return shuffle_impl<true>(idx, u01, cell_start, n_cell, length, ws, ws_bytes,
                              (cudaStream_t)stream);
}

// ---------------------------------------------------------------------------------
// shuffle_local for callers that may not allocate (CUDA graph capture, serving loops that
// forbid cudaMalloc): same kernels, same result bit for bit as pysdm_b200_shuffle_local, with
// the workspace in a file-scope device array instead of a caller-provided buffer.  No
// counterpart in the reference.
//
// Every GPU that loads the library holds the array (6 MiB) until the process exits.  All calls
// share it: calls that overlap on the GPU (two streams, two threads) race; `cnt` / `fill` are
// cleared by shuffle_impl at the start of every call, the rest is fully rewritten before it is
// read.  In host code the array's NAME is only a host-side stand-in: the device address comes
// from cudaGetSymbolAddress, looked up once per device with that device current.
// ---------------------------------------------------------------------------------
constexpr size_t kShuffleStaticBytes = 24 * (size_t)PYSDM_B200_SHUFFLE_STATIC_MAX_LENGTH + 4096;
static __device__ __align__(256) unsigned char g_shuffle_static_scratch[kShuffleStaticBytes];

constexpr int kShuffleStaticMaxDevices = 64;
static void *g_shuffle_static_addr[kShuffleStaticMaxDevices];  // device address, per device

PYSDM_EXPORT int pysdm_b200_shuffle_local_static(i64 *idx, const double *u01, const i64 *cell_start,
                                                 i64 n_cell, i64 length, void *stream)
{
    if (length < 0 || length > (i64)PYSDM_B200_SHUFFLE_STATIC_MAX_LENGTH
        || shuffle_ws_bytes(length) > kShuffleStaticBytes) {
        fprintf(stderr,
                "pysdm_b200_shuffle_local_static: length %lld is outside [0, %lld], the size the "
                "static scratch array was built for; use pysdm_b200_shuffle_local with a "
                "caller-provided workspace\n",
                (long long)length, (long long)PYSDM_B200_SHUFFLE_STATIC_MAX_LENGTH);
        return (int)cudaErrorInvalidValue;
    }
    int device = -1;
    cudaError_t err = cudaGetDevice(&device);
    if (err != cudaSuccess) return (int)err;
    if (device < 0 || device >= kShuffleStaticMaxDevices) return (int)cudaErrorInvalidDevice;
    void *scratch = g_shuffle_static_addr[device];
    if (scratch == nullptr) {  // set-up, once per device (a call with length 0 does only this)
        err = cudaGetSymbolAddress(&scratch, g_shuffle_static_scratch);
        if (err != cudaSuccess) return (int)err;
        g_shuffle_static_addr[device] = scratch;
    }
    return shuffle_impl<true>(idx, u01, cell_start, n_cell, length, scratch, kShuffleStaticBytes,
                              (cudaStream_t)stream);
}

// ---------------------------------------------------------------------------------
// stable sort of idx[:length] by cell rank + cell_start
// (reference: CellCaretaker, collisions_methods.py:587-631; counting sorts :682-741)
//   key(i) = cell_idx[cell_id[idx[i]]];  new_idx = stable sort; cell_start[c] = first
//   position whose key >= c (exclusive prefix sums of the per-cell counts).
// LSD radix passes of 8 bits over ceil(log2(n_cell)) bits; one cell => plain copy.
// ---------------------------------------------------------------------------------
__global__ void k_cell_keys(const i64 *__restrict__ idx, const i64 *__restrict__ cell_id,
                            const i64 *__restrict__ cell_idx, u32 *__restrict__ keys, i64 length)
{
    i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < length) keys[i] = (u32)cell_idx[cell_id[idx[i]]];
}

// sorted keys -> cell_start[0..n_cell]; thread i fills the starts of the cells in
// (key[i-1], key[i]]; the last thread also fills the tail up to n_cell
__global__ void k_cell_start_from_sorted(const u32 *__restrict__ keys, i64 length,
                                         i64 *__restrict__ cell_start, i64 n_cell)
{
    i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= length) return;
    i64 prev = i == 0 ? -1 : (i64)keys[i - 1];
    i64 cur = (i64)keys[i];
    for (i64 c = prev + 1; c <= cur; ++c) cell_start[c] = i;
    if (i == length - 1)
        for (i64 c = cur + 1; c <= n_cell; ++c) cell_start[c] = length;
}

__global__ void k_fill_i64(i64 *__restrict__ dst, i64 value, i64 n)
{
    i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) dst[i] = value;
}

static int bits_for(i64 n_values)
{
    int b = 0;
    while (((i64)1 << b) < n_values) ++b;
    return b;
}

static size_t sort_by_cell_ws_bytes(i64 length)
{
    size_t b = 0;
    b += 2 * align_up(sizeof(u32) * (size_t)length);  // keys ping-pong
    b += align_up(sizeof(i64) * (size_t)length);      // values ping-pong partner
    b += align_up(sizeof(u32) * radix_scratch_u32(length));
    return b;
}

PYSDM_EXPORT size_t pysdm_b200_sort_by_cell_workspace_bytes(i64 length)
{
    return sort_by_cell_ws_bytes(length);
}

PYSDM_EXPORT int pysdm_b200_sort_by_cell_id(i64 *new_idx, const i64 *idx, const i64 *cell_id,
                                            const i64 *cell_idx, i64 length, i64 *cell_start,
                                            i64 n_cell, void *ws, size_t ws_bytes, void *stream)
{
    cudaStream_t st = (cudaStream_t)stream;
    if (length <= 0) {
        k_fill_i64<<<grid_for(n_cell + 1), kThreads, 0, st>>>(cell_start, 0, n_cell + 1);
        return LAUNCH_CHECK();
    }
    if (length >= ((i64)1 << 31)) return (int)cudaErrorInvalidValue;
    const int g = grid_for(length);
    if (n_cell == 1) {
        k_copy_i64<<<g, kThreads, 0, st>>>(new_idx, idx, length);
        i64 h[2] = {0, length};
        // two scalars: written by tiny kernels to stay stream-ordered without host buffers
        k_fill_i64<<<1, 32, 0, st>>>(cell_start, h[0], 1);
        k_fill_i64<<<1, 32, 0, st>>>(cell_start + 1, h[1], 1);
        return LAUNCH_CHECK();
    }
    if (ws_bytes < sort_by_cell_ws_bytes(length)) return (int)cudaErrorMemoryAllocation;
    char *w = (char *)ws;
    u32 *keys_a = (u32 *)w;
    w += align_up(sizeof(u32) * (size_t)length);
    u32 *keys_b = (u32 *)w;
    w += align_up(sizeof(u32) * (size_t)length);
    i64 *vals_ws = (i64 *)w;
    w += align_up(sizeof(i64) * (size_t)length);
    u32 *scratch = (u32 *)w;

    k_cell_keys<<<g, kThreads, 0, st>>>(idx, cell_id, cell_idx, keys_a, length);
    const int passes = (bits_for(n_cell) + 7) / 8;
    // values: idx -> (ws | new_idx) alternating so that the last pass lands in new_idx
    const i64 *vin = idx;
    u32 *kin = keys_a, *kout = keys_b;
    for (int p = 0; p < passes; ++p) {
        bool to_new = ((passes - 1 - p) % 2) == 0;
        i64 *vout = to_new ? new_idx : vals_ws;
        radix_pass<u32, i64>(kin, vin, kout, vout, length, 8 * p, scratch, 0, st);
        vin = vout;
        u32 *t = kin;
        kin = kout;
        kout = t;
    }
    k_cell_start_from_sorted<<<g, kThreads, 0, st>>>(kin, length, cell_start, n_cell);
    return LAUNCH_CHECK();
}

// ---------------------------------------------------------------------------------
// sort_by_key  (reference: index_methods.py:46-48)
//   idx[:] = argsort(attr, kind="stable")[::-1]      (NaN sorts last in numpy)
// fp64 keys are mapped to order-preserving u64, 8 stable passes, the last one writes
// the reversed order directly.
// ---------------------------------------------------------------------------------
__global__ void k_f64_sort_keys(const double *__restrict__ attr, u64 *__restrict__ keys,
                                i64 *__restrict__ vals, i64 n)
{
    i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    double v = attr[i];
    if (v == 0.0) v = 0.0;  // -0.0 and +0.0 are ties for numpy
    u64 b = (u64)__double_as_longlong(v);
    if (v != v)
        b = ~(u64)0;  // every NaN last, like numpy
    else
        b = (b >> 63) ? ~b : (b | ((u64)1 << 63));
    keys[i] = b;
    vals[i] = i;
}

// Small arrays (the per-cell keys of one GPU's slab: a few 10^4 cells): the whole sort in ONE
// launch of one 1024-thread CTA.  The keys of the adaptive loop (`dt_left` per cell,
// collision.py:181) are far from arbitrary: all equal before the first sub-step, and from the
// second one on most cells are finished (dt_left == 0).  So:
//   1. the elements holding the MINIMAL key need no sorting -- they come last, in descending
//      index order (stable ascending argsort, reversed); the others are compacted (stable);
//   2. the compacted keys go through stable 8-bit LSD passes, but only over the digits in which
//      they actually differ (OR / AND of the keys), the last pass writing the reversed order.
// All keys equal: no pass at all.  General keys: the 8 passes of before on n - (count of min).
// Per pass: warp-private digit histograms in shared memory (warp w owns a contiguous chunk),
// exclusive scan over (digit-major, warp-minor), in-order replay of every chunk -- the same
// stable construction as radix_histogram / radix_scatter.
constexpr int kSmallSortThreads = 1024, kSmallSortWarps = kSmallSortThreads / 32;
constexpr i64 kSmallSortMax = 65536;

__device__ __forceinline__ u64 f64_sort_key(double v)
{
    if (v == 0.0) v = 0.0;  // -0.0 and +0.0 are ties for numpy
    const u64 b = (u64)__double_as_longlong(v);
    if (v != v) return ~(u64)0;  // every NaN last, like numpy
    return (b >> 63) ? ~b : (b | ((u64)1 << 63));
}

__device__ __forceinline__ u64 warp_reduce_u64(u64 v, int op)  // 0 min, 1 or, 2 and
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const u64 t = __shfl_xor_sync(0xffffffffu, v, o);
        v = op == 0 ? (t < v ? t : v) : (op == 1 ? (v | t) : (v & t));
    }
    return v;
}

__global__ void __launch_bounds__(kSmallSortThreads)
    k_sort_by_key_small(const double *__restrict__ attr, u64 *ka, i64 *va, u64 *kb, i64 *vb,
                        i64 *__restrict__ idx, int n, const i64 *__restrict__ skip_if_zero,
                        i64 *__restrict__ adaptive_end)
{
    // speculative adaptive sub-steps (pysdm_b200_sort_by_key_if): the loop ended before this one.
    // The sub-step this sort opens then does nothing either, and leaves the same verdict behind
    // (adaptive_end: the word the fused sub-step this sort opens accumulates adaptive_sdm_end in,
    // pysdm_b200_adaptive_presort)
    if (threadIdx.x == 0 && adaptive_end != nullptr) adaptive_end[0] = 0;
    if (skip_if_zero != nullptr && *skip_if_zero == 0) return;
    __shared__ u32 hist[kRadixBins][kSmallSortWarps + 1];  // [digit][warp], padded
    __shared__ u32 warp_sums[kSmallSortWarps];
    __shared__ unsigned long long s_min, s_or, s_and;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const unsigned lt = (1u << lane) - 1u;
    if (tid == 0) {
        s_min = ~0ull;
        s_or = 0ull;
        s_and = ~0ull;
    }
    __syncthreads();
    {  // minimal key
        u64 m = ~(u64)0;
        for (int i = tid; i < n; i += kSmallSortThreads) {
            const u64 k = f64_sort_key(attr[i]);
            m = k < m ? k : m;
        }
        m = warp_reduce_u64(m, 0);
        if (lane == 0) atomicMin(&s_min, (unsigned long long)m);
    }
    __syncthreads();
    const u64 kmin = s_min;
    // stable compaction of the elements above the minimum: warp w owns a contiguous chunk
    const int chunk0 = ((n + kSmallSortWarps - 1) / kSmallSortWarps + 31) / 32 * 32;
    const int begin0 = warp * chunk0, end0 = begin0 + chunk0 < n ? begin0 + chunk0 : n;
    {
        u32 cnt = 0;
        u64 o = 0, a = ~(u64)0;
        for (int i = begin0 + lane; i < end0; i += 32) {
            const u64 k = f64_sort_key(attr[i]);
            if (k != kmin) {
                ++cnt;
                o |= k;
                a &= k;
            }
        }
        cnt = __reduce_add_sync(0xffffffffu, cnt);
        o = warp_reduce_u64(o, 1);
        a = warp_reduce_u64(a, 2);
        if (lane == 0) {
            warp_sums[warp] = cnt;
            atomicOr(&s_or, (unsigned long long)o);
            atomicAnd(&s_and, (unsigned long long)a);
        }
    }
    __syncthreads();
    if (warp == 0) {
        const u32 w = warp_sums[lane];
        u32 winc = w;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const u32 t = __shfl_up_sync(0xffffffffu, winc, o);
            if (lane >= o) winc += t;
        }
        warp_sums[lane] = winc - w;
        if (lane == 31) hist[0][0] = winc;  // total
    }
    __syncthreads();
    const int n_act = (int)hist[0][0];
    const u64 differ = n_act > 1 ? (s_or & ~s_and) : 0;  // bits in which the compacted keys differ
    int passes[8], n_pass = 0;
    for (int p = 0; p < 8; ++p)
        if ((differ >> (8 * p)) & 0xffu) passes[n_pass++] = p;
    {
        u32 off = warp_sums[warp];  // compacted elements before my chunk
        for (int i0 = begin0; i0 < end0; i0 += 32) {
            const int i = i0 + lane;
            const bool in = i < end0;
            const u64 k = in ? f64_sort_key(attr[i]) : kmin;
            const bool act = in && k != kmin;
            const unsigned m = __ballot_sync(0xffffffffu, act);
            const u32 pos = off + __popc(m & lt);
            if (act) {
                if (n_pass == 0) {
                    idx[n_act - 1 - (int)pos] = i;  // already in order: reversed
                } else {
                    ka[pos] = k;
                    va[pos] = i;
                }
            } else if (in) {
                idx[n - 1 - (i - (int)pos)] = i;  // i - pos minimal-key elements precede i
            }
            off += __popc(m);
        }
    }
    __syncthreads();  // orders the global writes above before the reads of the passes
    const int chunk = ((n_act + kSmallSortWarps - 1) / kSmallSortWarps + 31) / 32 * 32;
    const int begin = warp * chunk < n_act ? warp * chunk : n_act;
    const int end = begin + chunk < n_act ? begin + chunk : n_act;
    for (int q = 0; q < n_pass; ++q) {
        const int shift = 8 * passes[q];
        for (int e = tid; e < kRadixBins * (kSmallSortWarps + 1); e += kSmallSortThreads)
            (&hist[0][0])[e] = 0;
        __syncthreads();  // also orders the previous pass's global writes before these reads
        for (int i = begin + lane; i < end; i += 32)
            atomicAdd(&hist[(u32)(ka[i] >> shift) & 0xffu][warp], 1u);
        __syncthreads();
        // exclusive scan of the 256 x 32 counts in (digit, warp) order: thread t owns the 8
        // consecutive entries 8t .. 8t + 7 = digit t / 4, warps 8 (t % 4) ..
        u32 c[8], sum = 0;
        {
            const int d = tid >> 2, w0 = (tid & 3) * 8;
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                c[u] = hist[d][w0 + u];
                sum += c[u];
            }
        }
        u32 inc = sum;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const u32 t = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += t;
        }
        if (lane == 31) warp_sums[warp] = inc;
        __syncthreads();
        if (warp == 0) {
            const u32 w = warp_sums[lane];
            u32 winc = w;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const u32 t = __shfl_up_sync(0xffffffffu, winc, o);
                if (lane >= o) winc += t;
            }
            warp_sums[lane] = winc - w;
        }
        __syncthreads();
        {
            u32 ex = warp_sums[warp] + inc - sum;
            const int d = tid >> 2, w0 = (tid & 3) * 8;
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                hist[d][w0 + u] = ex;
                ex += c[u];
            }
        }
        __syncthreads();
        // in-order replay of my chunk; hist[d][warp] is the running output offset
        const bool last = q == n_pass - 1;
        for (int i0 = begin; i0 < end; i0 += 32) {
            const int i = i0 + lane;
            const bool active = i < end;
            const u64 key = active ? ka[i] : 0;
            const i64 val = active ? va[i] : 0;
            const int d = active ? (int)((u32)(key >> shift) & 0xffu) : kRadixBins + lane;
            const unsigned peers = __match_any_sync(0xffffffffu, d);
            u32 pos = 0;
            if (active) pos = hist[d][warp] + __popc(peers & lt);
            __syncwarp();
            if (active && (peers & lt) == 0) hist[d][warp] += __popc(peers);
            __syncwarp();
            if (active) {
                if (last) {
                    idx[n_act - 1 - (int)pos] = val;  // [::-1]
                } else {
                    kb[pos] = key;
                    vb[pos] = val;
                }
            }
        }
        __syncthreads();  // every warp is done with hist[] and with this pass's input
        u64 *t = ka;
        ka = kb;
        kb = t;
        i64 *tv = va;
        va = vb;
        vb = tv;
    }
}

static size_t sort_by_key_ws_bytes(i64 n)
{
    return 2 * align_up(sizeof(u64) * (size_t)n) + 2 * align_up(sizeof(i64) * (size_t)n) +
           align_up(sizeof(u32) * radix_scratch_u32(n));
}

PYSDM_EXPORT size_t pysdm_b200_sort_by_key_workspace_bytes(i64 n) { return sort_by_key_ws_bytes(n); }

// fused.cu: the pre-pass on its own
int pysdm_fused_prepass(const i64 *cell_start, i64 n_cell, double *norm_factor, double timestep,
                        double dv, const i64 *rank_to_cell, const i64 *cell_idx, i64 *rank_info,
                        i64 *adaptive_end, cudaStream_t st);

static int sort_by_key_impl(i64 *idx, const double *attr, i64 n, void *ws, size_t ws_bytes,
                            const i64 *skip_if_zero, i64 *adaptive_end, cudaStream_t st)
{
    if (n <= 0) return 0;
    if (n == 1) {
        k_fill_i64<<<1, 32, 0, st>>>(idx, 0, 1);
        if (adaptive_end != nullptr) cudaMemsetAsync(adaptive_end, 0, sizeof(i64), st);
        return LAUNCH_CHECK();
    }
    if (ws_bytes < sort_by_key_ws_bytes(n)) return (int)cudaErrorMemoryAllocation;
    char *w = (char *)ws;
    u64 *ka = (u64 *)w;
    w += align_up(sizeof(u64) * (size_t)n);
    u64 *kb = (u64 *)w;
    w += align_up(sizeof(u64) * (size_t)n);
    i64 *va = (i64 *)w;
    w += align_up(sizeof(i64) * (size_t)n);
    i64 *vb = (i64 *)w;
    w += align_up(sizeof(i64) * (size_t)n);
    u32 *scratch = (u32 *)w;
    if (n <= kSmallSortMax) {
        k_sort_by_key_small<<<1, kSmallSortThreads, 0, st>>>(attr, ka, va, kb, vb, idx, (int)n,
                                                             skip_if_zero, adaptive_end);
        return LAUNCH_CHECK();
    }
    // (more keys than one CTA sorts: the skip word is not looked at, the sort always runs)
    k_f64_sort_keys<<<grid_for(n), kThreads, 0, st>>>(attr, ka, va, n);
    for (int p = 0; p < 8; ++p) {
        const bool last = p == 7;
        radix_pass<u64, i64>(ka, va, kb, last ? idx : vb, n, 8 * p, scratch, last ? 1 : 0, st);
        u64 *t = ka;
        ka = kb;
        kb = t;
        i64 *tv = va;
        va = vb;
        vb = tv;
    }
    if (adaptive_end != nullptr) cudaMemsetAsync(adaptive_end, 0, sizeof(i64), st);
    return LAUNCH_CHECK();
}

PYSDM_EXPORT int pysdm_b200_sort_by_key(i64 *idx, const double *attr, i64 n, void *ws,
                                        size_t ws_bytes, void *stream)
{
    return sort_by_key_impl(idx, attr, n, ws, ws_bytes, nullptr, nullptr, (cudaStream_t)stream);
}

// sort_by_key queued ahead of the read-back that decides whether it is needed (the adaptive
// sub-step loop, collision.py:178-190: `cell_idx.sort_by_key(dt_left)` opens every sub-step):
// does nothing when the device word *skip_if_zero -- the working length its predecessor's
// adaptive_sdm_end left -- is 0, i.e. the loop has ended; idx then keeps its content.
PYSDM_EXPORT int pysdm_b200_sort_by_key_if(i64 *idx, const double *attr, i64 n, void *ws,
                                           size_t ws_bytes, const i64 *skip_if_zero, void *stream)
{
    return sort_by_key_impl(idx, attr, n, ws, ws_bytes, skip_if_zero, nullptr, (cudaStream_t)stream);
}

// Opens one sub-step of the adaptive loop (collision.py:181): the sort above, which also zeroes
// the word the fused sub-step accumulates adaptive_sdm_end in; the per-rank part of that step's
// pre-pass is followed inside the fused kernel (prepass_done), so ONE launch sits between two
// launches of the fused step.  norm_factor != NULL: the per-cell factors of normalize() are
// (re)computed first (they only depend on cell_start, timestep and dv).
PYSDM_EXPORT int pysdm_b200_adaptive_presort(i64 *cell_idx, const double *dt_left, i64 n_cell,
                                             void *ws, size_t ws_bytes, const i64 *skip_if_zero,
                                             const i64 *cell_start, double *norm_factor,
                                             double timestep, double dv, i64 *adaptive_end,
                                             void *stream)
{
    if (n_cell <= 0 || !cell_idx || !dt_left || !adaptive_end) return (int)cudaErrorInvalidValue;
    cudaStream_t st = (cudaStream_t)stream;
    if (norm_factor != nullptr) {
        if (!cell_start) return (int)cudaErrorInvalidValue;
        const int err = pysdm_fused_prepass(cell_start, n_cell, norm_factor, timestep, dv, nullptr,
                                            nullptr, nullptr, nullptr, st);
        if (err) return err;
    }
    return sort_by_key_impl(cell_idx, dt_left, n_cell, ws, ws_bytes, skip_if_zero, adaptive_end, st);
}

// ---------------------------------------------------------------------------------
// remove_zero_n_or_flagged  (reference: collisions_methods.py:664-680)
// The reference compacts serially: a dead slot i is overwritten by the current last
// element, which is re-examined in place.  Net effect (derived in DESIGN.md): with D
// dead entries in [0, L) and L' = L - D, the k-th dead slot below L' (ascending)
// receives the k-th live entry at or above L' counted from the END; every slot in
// [L', L) becomes the flag value n_sd.  Reproduced with a scan + two scatters.
// ---------------------------------------------------------------------------------
__global__ void k_dead_flags(const i64 *__restrict__ mult, const i64 *__restrict__ idx, i64 length,
                             i64 n_sd, u32 *__restrict__ dead)
{
    i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= length) return;
    i64 v = idx[i];
    dead[i] = (v == n_sd || mult[v] == 0) ? 1u : 0u;
}

// dead_ps = exclusive prefix sums of the dead flags (dead_ps[length] = D)
__global__ void k_dead_positions(const u32 *__restrict__ dead_ps, i64 length,
                                 u32 *__restrict__ dead_pos)
{
    i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= length) return;
    u32 D = dead_ps[length];
    i64 new_len = length - (i64)D;
    bool is_dead = dead_ps[i + 1] != dead_ps[i];
    if (is_dead && i < new_len) dead_pos[dead_ps[i]] = (u32)i;
}

__global__ void k_move_live_tail(const u32 *__restrict__ dead_ps, i64 length,
                                 const u32 *__restrict__ dead_pos, const i64 *__restrict__ idx_in,
                                 i64 *__restrict__ idx, i64 *__restrict__ new_length_out)
{
    // reads idx_in (a snapshot of the tail) so that flag-filling cannot race with it
    i64 q = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= length) return;
    i64 D = (i64)dead_ps[length];
    i64 new_len = length - D;
    if (q == 0) *new_length_out = new_len;
    if (q < new_len) return;
    bool is_dead = dead_ps[q + 1] != dead_ps[q];
    if (!is_dead) {
        i64 dead_after = D - (i64)dead_ps[q + 1];
        i64 r = (length - 1 - q) - dead_after;  // rank among live entries, from the end
        idx[dead_pos[r]] = idx_in[q - new_len];
    }
}

__global__ void k_snapshot_tail(const u32 *__restrict__ dead_ps, i64 length,
                                const i64 *__restrict__ idx, i64 *__restrict__ tail,
                                i64 *__restrict__ idx_w, i64 n_sd)
{
    // tail[t] = idx[new_len + t]; then the slot is overwritten with the flag
    i64 q = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= length) return;
    i64 new_len = length - (i64)dead_ps[length];
    if (q >= new_len) {
        tail[q - new_len] = idx[q];
        idx_w[q] = n_sd;
    }
}

static size_t remove_zero_ws_bytes(i64 length)
{
    return align_up(sizeof(u32) * (size_t)(length + 1)) + align_up(sizeof(u32) * (size_t)length) +
           align_up(sizeof(i64) * (size_t)length) +
           align_up(sizeof(u32) * scan_scratch_elems<u32>(length + 1)) + 256;
}

PYSDM_EXPORT size_t pysdm_b200_remove_zero_workspace_bytes(i64 length)
{
    return remove_zero_ws_bytes(length);
}

// host_new_length: host pointer (pinned or pageable); the stream is synchronised.
PYSDM_EXPORT int pysdm_b200_remove_zero_n_or_flagged(const i64 *multiplicity, i64 *idx, i64 length,
                                                     i64 n_sd, i64 *host_new_length, void *ws,
                                                     size_t ws_bytes, void *stream)
{
    cudaStream_t st = (cudaStream_t)stream;
    if (length <= 0) {
        *host_new_length = 0;
        return 0;
    }
    if (ws_bytes < remove_zero_ws_bytes(length)) return (int)cudaErrorMemoryAllocation;
    char *w = (char *)ws;
    u32 *dead = (u32 *)w;
    w += align_up(sizeof(u32) * (size_t)(length + 1));
    u32 *dead_pos = (u32 *)w;
    w += align_up(sizeof(u32) * (size_t)length);
    i64 *tail = (i64 *)w;
    w += align_up(sizeof(i64) * (size_t)length);
    u32 *scan_ws = (u32 *)w;
    w += align_up(sizeof(u32) * scan_scratch_elems<u32>(length + 1));
    i64 *dev_len = (i64 *)w;
    const int g = grid_for(length);
    k_dead_flags<<<g, kThreads, 0, st>>>(multiplicity, idx, length, n_sd, dead);
    device_exclusive_scan<u32>(dead, dead, length, 1, scan_ws, st);
    k_dead_positions<<<g, kThreads, 0, st>>>(dead, length, dead_pos);
    k_snapshot_tail<<<g, kThreads, 0, st>>>(dead, length, idx, tail, idx, n_sd);
    k_move_live_tail<<<g, kThreads, 0, st>>>(dead, length, dead_pos, tail, idx, dev_len);
    int err = LAUNCH_CHECK();
    if (err) return err;
    err = (int)cudaMemcpyAsync(host_new_length, dev_len, sizeof(i64), cudaMemcpyDeviceToHost, st);
    if (err) return err;
    return (int)cudaStreamSynchronize(st);
}

// ---------------------------------------------------------------------------------
// cell_id  (reference: collisions_methods.py:407-416): cell_id = strides . cell_origin
// ---------------------------------------------------------------------------------
__global__ void k_cell_id(i64 *__restrict__ cell_id, const i64 *__restrict__ cell_origin,
                          const i64 *__restrict__ strides, i64 n_dims, i64 n_sd)
{
    i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_sd) return;
    i64 acc = 0;
    for (i64 d = 0; d < n_dims; ++d) acc += strides[d] * cell_origin[d * n_sd + i];
    cell_id[i] = acc;
}

PYSDM_EXPORT int pysdm_b200_cell_id(i64 *cell_id, const i64 *cell_origin, const i64 *strides,
                                    i64 n_dims, i64 n_sd, void *stream)
{
    if (n_sd <= 0) return 0;
    k_cell_id<<<grid_for(n_sd), kThreads, 0, (cudaStream_t)stream>>>(cell_id, cell_origin, strides,
                                                                    n_dims, n_sd);
    return LAUNCH_CHECK();
}

// ---------------------------------------------------------------------------------
// largest cell population: max_c (cell_start[c+1] - cell_start[c]) -> *host_result
// (host-side dispatch between the fused step and the method-level kernels)
// ---------------------------------------------------------------------------------
__global__ void k_max_cell_size(const i64 *__restrict__ cell_start, i64 n_cell,
                                unsigned long long *__restrict__ result)
{
    i64 m = 0;
    for (i64 c = (i64)blockIdx.x * blockDim.x + threadIdx.x; c < n_cell;
         c += (i64)gridDim.x * blockDim.x) {
        i64 n = cell_start[c + 1] - cell_start[c];
        m = n > m ? n : m;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        i64 t = __shfl_xor_sync(0xffffffffu, m, o);
        m = t > m ? t : m;
    }
    if ((threadIdx.x & 31) == 0 && m > 0) atomicMax(result, (unsigned long long)m);
}

PYSDM_EXPORT int pysdm_b200_max_cell_size(const i64 *cell_start, i64 n_cell, i64 *host_result,
                                          i64 *dev_scratch, void *stream)
{
    cudaStream_t st = (cudaStream_t)stream;
    int err = (int)cudaMemsetAsync(dev_scratch, 0, sizeof(i64), st);
    if (err) return err;
    if (n_cell > 0) {
        int g = grid_for(n_cell);
        if (g > 1184) g = 1184;
        k_max_cell_size<<<g, kThreads, 0, st>>>(cell_start, n_cell,
                                                (unsigned long long *)dev_scratch);
        err = LAUNCH_CHECK();
        if (err) return err;
    }
    err = (int)cudaMemcpyAsync(host_result, dev_scratch, sizeof(i64), cudaMemcpyDeviceToHost, st);
    if (err) return err;
    return (int)cudaStreamSynchronize(st);
}
