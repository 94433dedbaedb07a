// philox.cuh -- counter-based Philox4x32-10 shared by the Random kernel (storage_ops.cu)
// and the fused collision step (fused.cu), so both produce the same element streams.
#pragma once
#include "common.cuh"

// ---- Philox4x32-10 -----------------------------------------------------------------
// Element e of call number `call` under `seed` is a pure function of (seed, call, e):
// counter = (e/2 lo, e/2 hi, call lo, call hi), key = (seed lo, seed hi); one Philox
// block yields two doubles (52 random bits each, see u01_from_words).
// Independent of launch geometry, so any kernel can regenerate element e in place.
__device__ __forceinline__ void philox4x32_10(u32 c0, u32 c1, u32 c2, u32 c3, u32 k0, u32 k1,
                                              u32 out[4])
{
    const u32 M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        u32 hi0 = __umulhi(M0, c0), lo0 = M0 * c0;
        u32 hi1 = __umulhi(M1, c2), lo1 = M1 * c2;
        u32 n0 = hi1 ^ c1 ^ k0, n1 = lo1, n2 = hi0 ^ c3 ^ k1, n3 = lo0;
        c0 = n0;
        c1 = n1;
        c2 = n2;
        c3 = n3;
        k0 += W0;
        k1 += W1;
    }
    out[0] = c0;
    out[1] = c1;
    out[2] = c2;
    out[3] = c3;
}

// the same block function with the 10 round keys (k0 + r W0, k1 + r W1) precomputed by the
// host: the keys are read straight from the kernel-argument constant bank by the LOP3s instead
// of being re-derived with 20 integer adds per block
__device__ __forceinline__ void philox4x32_10_rk(u32 c0, u32 c1, u32 c2, u32 c3,
                                                 const u32 (&rk)[20], u32 out[4])
{
    const u32 M0 = 0xD2511F53u, M1 = 0xCD9E8D57u;
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        u32 hi0 = __umulhi(M0, c0), lo0 = M0 * c0;
        u32 hi1 = __umulhi(M1, c2), lo1 = M1 * c2;
        u32 n0 = hi1 ^ c1 ^ rk[2 * r], n1 = lo1, n2 = hi0 ^ c3 ^ rk[2 * r + 1], n3 = lo0;
        c0 = n0;
        c1 = n1;
        c2 = n2;
        c3 = n3;
    }
    out[0] = c0;
    out[1] = c1;
    out[2] = c2;
    out[3] = c3;
}

// uniform in [0, 1) from 64 random bits: the top 52 bits fill the mantissa of a double in
// [1, 2), minus 1 (resolution 2^-52; no integer->fp conversion instruction needed)
__device__ __forceinline__ double u01_from_words(u32 hi, u32 lo)
{
    u64 x = ((u64)hi << 32) | (u64)lo;
    return __longlong_as_double((long long)(0x3FF0000000000000ull | (x >> 12))) - 1.0;
}
