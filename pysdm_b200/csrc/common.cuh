// common.cuh -- shared device helpers for the B200 SDM collision kernels (sm_100a).
//
// Conventions (mirroring the reference's Storage, impl_numba/storage.py:16-19):
// all floats are fp64, all integers int64, flags are 1-byte bools.  The file is
// compiled with -fmad=false: a*b+c is never contracted unless fma() is written
// explicitly (the reference's Numba code is built with fastmath=False for the
// method bodies; see DESIGN.md "floating-point contract").
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

typedef int64_t i64;
typedef int32_t i32;
typedef uint64_t u64;
typedef uint32_t u32;
typedef uint8_t u8;
typedef uint16_t u16;

#define PYSDM_EXPORT extern "C" __attribute__((visibility("default")))

constexpr int kThreads = 256;

static inline int grid_for(i64 n, int threads = kThreads)
{
    i64 g = (n + threads - 1) / threads;
    return (int)(g < 1 ? 1 : g);
}

#define LAUNCH_CHECK() ((int)cudaGetLastError())

// ---- pair_indices (reference: impl_numba/methods/collisions_methods.py:16-35) ----
// pair slot i covers positions (2i, 2i+1) when is_first_in_pair[2i], else (2i+1, 2i+2)
struct PairJK {
    i64 j, k;
    bool skip;
};
__device__ __forceinline__ PairJK pair_indices(i64 i, const i64 *__restrict__ idx,
                                               const u8 *__restrict__ flag, double prob_like)
{
    PairJK r;
    if (prob_like == 0.0) {
        r.j = -1;
        r.k = -1;
        r.skip = true;
    } else {
        i64 offset = 1 - (i64)flag[2 * i];
        r.j = idx[2 * i + offset];
        r.k = idx[2 * i + 1 + offset];
        r.skip = false;
    }
    return r;
}

// Python floor division on int64 (multiplicities are positive in practice)
__device__ __forceinline__ i64 floordiv(i64 a, i64 b)
{
    i64 q = a / b;
    if ((a % b != 0) && ((a < 0) != (b < 0))) q -= 1;
    return q;
}

// Python/numba min(a, b), max(a, b) on floats: NaN in `a` persists (quirk A9)
__device__ __forceinline__ double py_min(double a, double b) { return b < a ? b : a; }
__device__ __forceinline__ double py_max(double a, double b) { return b > a ? b : a; }

// ---- per-cell int64 counters: one atomic per group of lanes hitting the same cell ----
// Fast path: all active lanes address the same counter (boxes, or a warp inside one
// cell) -> warp reduction + a single atomic; otherwise one atomic per lane.
__device__ __forceinline__ void cell_counter_add(i64 *counters, i64 cid, i64 value, bool active)
{
    unsigned mask = __ballot_sync(0xffffffffu, active);
    if (mask == 0) return;
    int leader = __ffs(mask) - 1;
    i64 lead_cid = __shfl_sync(0xffffffffu, cid, leader);
    bool uniform = __all_sync(0xffffffffu, !active || cid == lead_cid);
    if (uniform) {
        i64 v = active ? value : 0;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if ((int)(threadIdx.x & 31) == leader && v != 0)
            atomicAdd((unsigned long long *)&counters[lead_cid], (unsigned long long)v);
    } else if (active && value != 0) {
        atomicAdd((unsigned long long *)&counters[cid], (unsigned long long)value);
    }
}

// atomic min on fp64 through the integer ordering of non-negative doubles; a CAS
// loop covers negative values.  `cur` NaN is left untouched by the callers.
__device__ __forceinline__ void atomic_min_f64(double *addr, double v)
{
    if (v >= 0.0 && !(*addr < 0.0)) {
        atomicMin((long long *)addr, __double_as_longlong(v));
        // a concurrent negative store is impossible here: all writers of one array use
        // values of one sign in this code base (time steps are >= 0)
    } else {
        unsigned long long *a = (unsigned long long *)addr;
        unsigned long long old = *a, assumed;
        do {
            assumed = old;
            if (!(v < __longlong_as_double((long long)assumed))) break;
            old = atomicCAS(a, assumed, (unsigned long long)__double_as_longlong(v));
        } while (assumed != old);
    }
}
