// primitives.cuh -- device-wide building blocks written for this path: exclusive scan,
// stable LSD radix sort of (key, value) pairs, min/max reductions.
// No thrust / cub: everything the hot path launches is in this directory.
#pragma once
#include "common.cuh"

// ---------------------------------------------------------------------------------
// block-level exclusive scan over kThreads threads (warp shuffles + one smem hop)
// ---------------------------------------------------------------------------------
template <typename T>
__device__ __forceinline__ T block_exclusive_scan(T v, T &block_total)
{
    __shared__ T warp_sums[kThreads / 32];
    __shared__ T total;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    T inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        T t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
    }
    if (lane == 31) warp_sums[warp] = inc;
    __syncthreads();
    if (warp == 0) {
        T w = lane < kThreads / 32 ? warp_sums[lane] : T(0);
        T winc = w;
#pragma unroll
        for (int o = 1; o < kThreads / 32; o <<= 1) {
            T t = __shfl_up_sync(0xffffffffu, winc, o);
            if (lane >= o) winc += t;
        }
        if (lane < kThreads / 32) warp_sums[lane] = winc - w;
        if (lane == kThreads / 32 - 1) total = winc;
    }
    __syncthreads();
    T res = warp_sums[warp] + inc - v;
    block_total = total;
    __syncthreads();  // smem reusable by the next call
    return res;
}

// ---------------------------------------------------------------------------------
// device-wide exclusive scan (three passes: tile sums, scan of sums, rescan)
//   out[i] = sum(in[0..i)), and out[n] = total when `with_total`
// `in` may alias `out`.  Scratch: (n / kScanTile + 2) elements of T.
// ---------------------------------------------------------------------------------
constexpr int kScanItems = 8;
constexpr int kScanTile = kThreads * kScanItems;

template <typename T>
__global__ void __launch_bounds__(kThreads) scan_tile_sums(const T *__restrict__ in, i64 n,
                                                           T *__restrict__ sums)
{
    const i64 base = (i64)blockIdx.x * kScanTile;
    T s = 0;
#pragma unroll
    for (int k = 0; k < kScanItems; ++k) {
        i64 i = base + (i64)k * kThreads + threadIdx.x;
        if (i < n) s += in[i];
    }
    T total;
    block_exclusive_scan(s, total);
    if (threadIdx.x == 0) sums[blockIdx.x] = total;
}

template <typename T>
__global__ void __launch_bounds__(kThreads) scan_sums_inplace(T *sums, i64 n_tiles)
{
    // single block: chunks of kThreads with a running carry; sums[n_tiles] <- grand total
    T carry = 0;
    for (i64 base = 0; base < n_tiles; base += kThreads) {
        i64 i = base + threadIdx.x;
        T v = i < n_tiles ? sums[i] : T(0);
        T total;
        T ex = block_exclusive_scan(v, total);
        if (i < n_tiles) sums[i] = carry + ex;
        carry += total;
    }
    if (threadIdx.x == 0) sums[n_tiles] = carry;
}

template <typename T>
__global__ void __launch_bounds__(kThreads) scan_rescan(const T *in, T *out, i64 n,
                                                        const T *__restrict__ sums, int with_total)
{
    // blocked arrangement: thread t owns items [t*kScanItems, (t+1)*kScanItems) of the tile
    const i64 base = (i64)blockIdx.x * kScanTile + (i64)threadIdx.x * kScanItems;
    T v[kScanItems];
    T s = 0;
#pragma unroll
    for (int k = 0; k < kScanItems; ++k) {
        i64 i = base + k;
        v[k] = i < n ? in[i] : T(0);
        s += v[k];
    }
    T total;
    T ex = block_exclusive_scan(s, total) + sums[blockIdx.x];
#pragma unroll
    for (int k = 0; k < kScanItems; ++k) {
        i64 i = base + k;
        if (i < n) out[i] = ex;
        ex += v[k];
    }
    if (with_total && blockIdx.x == gridDim.x - 1 && threadIdx.x == 0) out[n] = sums[gridDim.x];
}

template <typename T>
static inline size_t scan_scratch_elems(i64 n)
{
    return (size_t)((n + kScanTile - 1) / kScanTile + 2);
}

template <typename T>
static inline void device_exclusive_scan(const T *in, T *out, i64 n, int with_total, T *scratch,
                                         cudaStream_t st)
{
    if (n <= 0) {
        if (with_total) cudaMemsetAsync(out, 0, sizeof(T), st);
        return;
    }
    const i64 n_tiles = (n + kScanTile - 1) / kScanTile;
    scan_tile_sums<T><<<(int)n_tiles, kThreads, 0, st>>>(in, n, scratch);
    scan_sums_inplace<T><<<1, kThreads, 0, st>>>(scratch, n_tiles);
    scan_rescan<T><<<(int)n_tiles, kThreads, 0, st>>>(in, out, n, scratch, with_total);
}

// ---------------------------------------------------------------------------------
// stable LSD radix sort of (key, value) pairs, 8-bit digits.
// One warp owns a contiguous chunk of radix_chunk(n) elements: per-(digit, chunk)
// histogram -> device scan (digit-major) -> each warp replays its chunk in order and
// scatters with warp-private running offsets (stable by construction, no inter-warp
// coordination, no atomics on global memory).
// ---------------------------------------------------------------------------------
constexpr int kRadixWarps = kThreads / 32;
constexpr int kRadixBins = 256;
constexpr int kRadixGroup = 8;  // elements per lane loaded before any is consumed

// elements per warp: small arrays (per-cell keys of sort_by_key, a few 10^4 elements) are cut
// finer so that more than a handful of warps work on them
static inline int radix_chunk(i64 n) { return n <= ((i64)1 << 18) ? 256 : 1024; }

template <typename K>
__global__ void __launch_bounds__(kThreads) radix_histogram(const K *__restrict__ keys, i64 n,
                                                            int shift, u32 *__restrict__ hist,
                                                            i64 n_chunks, int chunk_size)
{
    __shared__ u32 cnt[kRadixWarps][kRadixBins];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const i64 chunk = (i64)blockIdx.x * kRadixWarps + warp;
    for (int d = lane; d < kRadixBins; d += 32) cnt[warp][d] = 0;
    __syncwarp();
    if (chunk < n_chunks) {
        const i64 base = chunk * chunk_size;
        for (int g0 = 0; g0 < chunk_size / 32; g0 += kRadixGroup) {
            K key[kRadixGroup];
#pragma unroll
            for (int u = 0; u < kRadixGroup; ++u) {
                const i64 i = base + (g0 + u) * 32 + lane;
                key[u] = i < n ? keys[i] : K(0);
            }
#pragma unroll
            for (int u = 0; u < kRadixGroup; ++u) {
                const i64 i = base + (g0 + u) * 32 + lane;
                if (i < n) atomicAdd(&cnt[warp][(u32)(key[u] >> shift) & 0xffu], 1u);
            }
        }
        __syncwarp();
        for (int d = lane; d < kRadixBins; d += 32) hist[(i64)d * n_chunks + chunk] = cnt[warp][d];
    }
}

template <typename K, typename V>
__global__ void __launch_bounds__(kThreads)
    radix_scatter(const K *__restrict__ keys_in, const V *__restrict__ vals_in,
                  K *__restrict__ keys_out, V *__restrict__ vals_out, i64 n, int shift,
                  const u32 *__restrict__ offsets, i64 n_chunks, int reverse_output,
                  int chunk_size)
{
    __shared__ u32 running[kRadixWarps][kRadixBins];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const i64 chunk = (i64)blockIdx.x * kRadixWarps + warp;
    if (chunk >= n_chunks) return;  // whole warp leaves together
    for (int d = lane; d < kRadixBins; d += 32) running[warp][d] = offsets[(i64)d * n_chunks + chunk];
    __syncwarp();
    const i64 base = chunk * chunk_size;
    const unsigned lt = (1u << lane) - 1u;
    for (int g0 = 0; g0 < chunk_size / 32; g0 += kRadixGroup) {
        if (base + (i64)g0 * 32 >= n) break;  // uniform across the warp
        // the loads of a whole group are in flight together; the replay below is in order
        K key[kRadixGroup];
        V val[kRadixGroup];
#pragma unroll
        for (int u = 0; u < kRadixGroup; ++u) {
            const i64 i = base + (g0 + u) * 32 + lane;
            const bool active = i < n;
            key[u] = active ? keys_in[i] : K(0);
            val[u] = active ? vals_in[i] : V(0);
        }
#pragma unroll
        for (int u = 0; u < kRadixGroup; ++u) {
            const i64 i = base + (g0 + u) * 32 + lane;
            const bool active = i < n;
            const int d = active ? (int)((u32)(key[u] >> shift) & 0xffu) : kRadixBins + lane;
            const unsigned peers = __match_any_sync(0xffffffffu, d);
            u32 pos = 0;
            if (active) pos = running[warp][d] + __popc(peers & lt);
            __syncwarp();
            if (active && (peers & lt) == 0) running[warp][d] += __popc(peers);
            __syncwarp();
            if (active) {
                const i64 p = reverse_output ? (n - 1 - (i64)pos) : (i64)pos;
                keys_out[p] = key[u];
                vals_out[p] = val[u];
            }
        }
    }
}

static inline i64 radix_n_chunks(i64 n) { return (n + radix_chunk(n) - 1) / radix_chunk(n); }

// scratch (in u32 elements) needed by one radix pass over n elements
static inline size_t radix_scratch_u32(i64 n)
{
    size_t h = (size_t)kRadixBins * (size_t)radix_n_chunks(n);
    return h + 1 + scan_scratch_elems<u32>((i64)h) + 8;
}

// One stable pass on bits [shift, shift+8).  `reverse_output` writes position n-1-p
// (used by the last pass of sort_by_key to produce the reversed order).
template <typename K, typename V>
static inline void radix_pass(const K *keys_in, const V *vals_in, K *keys_out, V *vals_out, i64 n,
                              int shift, u32 *scratch, int reverse_output, cudaStream_t st)
{
    const i64 n_chunks = radix_n_chunks(n);
    const i64 h = (i64)kRadixBins * n_chunks;
    u32 *hist = scratch;
    u32 *scan_ws = scratch + h + 1;
    const int blocks = (int)((n_chunks + kRadixWarps - 1) / kRadixWarps);
    const int chunk = radix_chunk(n);
    radix_histogram<K><<<blocks, kThreads, 0, st>>>(keys_in, n, shift, hist, n_chunks, chunk);
    device_exclusive_scan<u32>(hist, hist, h, 0, scan_ws, st);
    radix_scatter<K, V><<<blocks, kThreads, 0, st>>>(keys_in, vals_in, keys_out, vals_out, n, shift,
                                                     hist, n_chunks, reverse_output, chunk);
}

// ---------------------------------------------------------------------------------
// min / max reductions (fp64 with numpy NaN propagation, int64)
// ---------------------------------------------------------------------------------
template <typename T, bool IsMax>
__device__ __forceinline__ T red_op(T a, T b)
{
    return IsMax ? (b > a ? b : a) : (b < a ? b : a);
}

template <typename T, bool IsMax>
__global__ void __launch_bounds__(kThreads) reduce_minmax(const T *__restrict__ in, i64 n,
                                                          T *__restrict__ partial, int *nan_flag)
{
    __shared__ T sm[kThreads / 32];
    T acc = in[0];
    bool has_nan = false;
    for (i64 i = (i64)blockIdx.x * kThreads + threadIdx.x; i < n; i += (i64)gridDim.x * kThreads) {
        T v = in[i];
        if (v != v) has_nan = true;
        acc = red_op<T, IsMax>(acc, v);
    }
    if (has_nan) *nan_flag = 1;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc = red_op<T, IsMax>(acc, __shfl_xor_sync(0xffffffffu, acc, o));
    if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < kThreads / 32; ++w) acc = red_op<T, IsMax>(acc, sm[w]);
        partial[blockIdx.x] = acc;
    }
}
