// storage_ops.cu -- what `Storage` needs beyond memory plumbing: fp64 elementwise
// in-place operators, min/max reductions, and the device random number generator
// (Philox4x32-10, counter based).  Reference semantics:
//   impl_numba/storage.py:54-213 + storage_impl.py:11-83 (operators),
//   impl_numba/random.py:9-19 (Random fills a Storage with uniforms in [0,1)).
#include "pysdm_b200.h"
#include <math.h>

#include "primitives.cuh"
#include "dependents.cuh"

// ---- elementwise ---------------------------------------------------------------
enum EwOp {
    EW_ADD = 0,
    EW_SUB = 1,
    EW_MUL = 2,
    EW_DIV = 3,
    EW_DIV_IF_NOT_ZERO = 4,  // storage_impl.py divide_if_not_zero
    EW_POW = 5,              // storage_impl.py power: sign(x) * |x|**e (quirk A12)
    EW_AXPY = 6,             // storage_impl.py add_with_multiplier: out += s * b
    EW_FLOOR = 7,
    EW_EXP = 8,
    EW_ABS = 9,
    EW_MOD = 10  // Python float modulo (result takes the sign of the divisor)
};

__device__ __forceinline__ double py_mod(double a, double b)
{
    double r = fmod(a, b);
    if (r != 0 && ((r < 0) != (b < 0))) r += b;
    return r;
}

__device__ __forceinline__ double ew_apply(int op, double a, double b, double s)
{
    switch (op) {
    case EW_ADD: return a + b;
    case EW_SUB: return a - b;
    case EW_MUL: return a * b;
    case EW_DIV: return a / b;
    case EW_DIV_IF_NOT_ZERO: return b != 0.0 ? a / b : a;
    case EW_POW: return signed_pow(a, s);
    case EW_AXPY: return fma(s, b, a);
    case EW_FLOOR: return floor(b);
    case EW_EXP: return exp(b);
    case EW_ABS: return fabs(b);
    default: return py_mod(a, b);
    }
}

// out[i] = op(a[i], b ? b[i] : s, s); pointers may alias each other
__global__ void k_elementwise(int op, double *out, const double *a, const double *b, double s, i64 n)
{
    i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    double av = a ? a[i] : 0.0;
    double bv = b ? b[i] : s;
    out[i] = ew_apply(op, av, bv, s);
}

PYSDM_EXPORT int pysdm_b200_elementwise(int op, double *out, const double *a, const double *b,
                                        double scalar, i64 n, void *stream)
{
    if (n <= 0) return 0;
    k_elementwise<<<grid_for(n), kThreads, 0, (cudaStream_t)stream>>>(op, out, a, b, scalar, n);
    return LAUNCH_CHECK();
}

// row_modulo (storage_impl.py:31-35): output[d, i] %= divisor[d]
__global__ void k_row_modulo(double *__restrict__ out, const double *__restrict__ divisor,
                             i64 n_rows, i64 n_cols)
{
    i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_rows * n_cols) return;
    out[i] = py_mod(out[i], divisor[i / n_cols]);
}

PYSDM_EXPORT int pysdm_b200_row_modulo(double *out, const double *divisor, i64 n_rows, i64 n_cols,
                                       void *stream)
{
    if (n_rows * n_cols <= 0) return 0;
    k_row_modulo<<<grid_for(n_rows * n_cols), kThreads, 0, (cudaStream_t)stream>>>(out, divisor,
                                                                                  n_rows, n_cols);
    return LAUNCH_CHECK();
}

// ---- mixed int64 / fp64 operators of Displacement (dynamics/displacement.py:141-153) ------
__global__ void k_floor_to_i64(i64 *__restrict__ out, const double *__restrict__ in, i64 n)
{
    i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = (i64)floor(in[i]);
}

__global__ void k_sub_i64(double *__restrict__ inout, const i64 *__restrict__ b, i64 n)
{
    i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) inout[i] -= (double)b[i];
}

__global__ void k_row_modulo_i64(i64 *__restrict__ out, const i64 *__restrict__ divisor,
                                 i64 n_rows, i64 n_cols)
{
    i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_rows * n_cols) return;
    const i64 d = divisor[i / n_cols];
    i64 r = out[i] % d;  // numpy's %: the result takes the sign of the divisor
    if (r != 0 && ((r < 0) != (d < 0))) r += d;
    out[i] = r;
}

PYSDM_EXPORT int pysdm_b200_floor_to_i64(i64 *out, const double *in, i64 n, void *stream)
{
    if (n <= 0) return 0;
    k_floor_to_i64<<<grid_for(n), kThreads, 0, (cudaStream_t)stream>>>(out, in, n);
    return LAUNCH_CHECK();
}

PYSDM_EXPORT int pysdm_b200_sub_i64(double *inout, const i64 *b, i64 n, void *stream)
{
    if (n <= 0) return 0;
    k_sub_i64<<<grid_for(n), kThreads, 0, (cudaStream_t)stream>>>(inout, b, n);
    return LAUNCH_CHECK();
}

PYSDM_EXPORT int pysdm_b200_row_modulo_i64(i64 *out, const i64 *divisor, i64 n_rows, i64 n_cols,
                                           void *stream)
{
    if (n_rows * n_cols <= 0) return 0;
    k_row_modulo_i64<<<grid_for(n_rows * n_cols), kThreads, 0, (cudaStream_t)stream>>>(
        out, divisor, n_rows, n_cols);
    return LAUNCH_CHECK();
}

// ---- reductions: amin / amax with numpy NaN propagation ---------------------------
constexpr int kReduceBlocks = 592;  // 4 per SM on 148 SMs

template <typename T, bool IsMax>
__global__ void k_reduce_final(const T *__restrict__ partial, int n_partial,
                               const int *__restrict__ nan_flag, T *__restrict__ result)
{
    // one warp: strided pass over the partials, then a shuffle tree (min / max are
    // order-independent)
    T acc = partial[0];
    for (int i = threadIdx.x; i < n_partial; i += 32) acc = red_op<T, IsMax>(acc, partial[i]);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc = red_op<T, IsMax>(acc, __shfl_xor_sync(0xffffffffu, acc, o));
    if (threadIdx.x == 0) {
        result[0] = acc;
        ((i64 *)result)[1] = (i64)(*nan_flag);
    }
}

// scratch: >= (kReduceBlocks + 4) * 8 bytes of device memory.  host_result[0] receives the
// extremum, host_result[1] (as int64) a "NaN seen" flag; the stream is synchronised.
template <typename T, bool IsMax>
static int reduce_impl(const T *in, i64 n, void *host_result, void *scratch, cudaStream_t st)
{
    if (n <= 0) return (int)cudaErrorInvalidValue;
    T *partial = (T *)scratch;
    int *nan_flag = (int *)(partial + kReduceBlocks);
    T *result = (T *)(partial + kReduceBlocks + 1);
    int blocks = grid_for(n);
    if (blocks > kReduceBlocks) blocks = kReduceBlocks;
    cudaMemsetAsync(nan_flag, 0, sizeof(T), st);
    reduce_minmax<T, IsMax><<<blocks, kThreads, 0, st>>>(in, n, partial, nan_flag);
    k_reduce_final<T, IsMax><<<1, 32, 0, st>>>(partial, blocks, nan_flag, result);
    int err = LAUNCH_CHECK();
    if (err) return err;
    err = (int)cudaMemcpyAsync(host_result, result, 2 * sizeof(T), cudaMemcpyDeviceToHost, st);
    if (err) return err;
    return (int)cudaStreamSynchronize(st);
}

PYSDM_EXPORT size_t pysdm_b200_reduce_scratch_bytes(void) { return (kReduceBlocks + 4) * 8; }

PYSDM_EXPORT int pysdm_b200_minmax_f64(const double *in, i64 n, int is_max, double *host_result,
                                       void *scratch, void *stream)
{
    return is_max ? reduce_impl<double, true>(in, n, host_result, scratch, (cudaStream_t)stream)
                  : reduce_impl<double, false>(in, n, host_result, scratch, (cudaStream_t)stream);
}

PYSDM_EXPORT int pysdm_b200_minmax_i64(const i64 *in, i64 n, int is_max, i64 *host_result,
                                       void *scratch, void *stream)
{
    return is_max ? reduce_impl<i64, true>(in, n, host_result, scratch, (cudaStream_t)stream)
                  : reduce_impl<i64, false>(in, n, host_result, scratch, (cudaStream_t)stream);
}

#include "philox.cuh"

__global__ void k_philox_uniform(double *__restrict__ out, i64 n, u64 seed, u64 call)
{
    i64 b = (i64)blockIdx.x * blockDim.x + threadIdx.x;  // Philox block = elements 2b, 2b+1
    if (2 * b >= n) return;
    u32 w[4];
    philox4x32_10((u32)b, (u32)((u64)b >> 32), (u32)call, (u32)(call >> 32), (u32)seed,
                  (u32)(seed >> 32), w);
    double u0 = u01_from_words(w[0], w[1]), u1 = u01_from_words(w[2], w[3]);
    if (2 * b + 1 < n && (((uintptr_t)out) & 15) == 0) {
        reinterpret_cast<double2 *>(out)[b] = make_double2(u0, u1);
    } else {
        out[2 * b] = u0;
        if (2 * b + 1 < n) out[2 * b + 1] = u1;
    }
}

PYSDM_EXPORT int pysdm_b200_philox_uniform(double *out, i64 n, u64 seed, u64 call, void *stream)
{
    if (n <= 0) return 0;
    k_philox_uniform<<<grid_for((n + 1) / 2), kThreads, 0, (cudaStream_t)stream>>>(out, n, seed,
                                                                                  call);
    return LAUNCH_CHECK();
}

// raw Philox block for known-answer tests (host-visible through a device buffer)
__global__ void k_philox_raw(u32 *out, u32 c0, u32 c1, u32 c2, u32 c3, u32 k0, u32 k1)
{
    u32 w[4];
    philox4x32_10(c0, c1, c2, c3, k0, k1, w);
    for (int i = 0; i < 4; ++i) out[i] = w[i];
}

PYSDM_EXPORT int pysdm_b200_philox_raw(u32 *dev_out4, u32 c0, u32 c1, u32 c2, u32 c3, u32 k0,
                                       u32 k1, void *stream)
{
    k_philox_raw<<<1, 1, 0, (cudaStream_t)stream>>>(dev_out4, c0, c1, c2, c3, k0, k1);
    return LAUNCH_CHECK();
}

// ---- physics one-liners (impl_numba/methods/physics_methods.py:107-133 with
// physics/particle_shape_and_density/liquid_spheres.py:17-23) --------------------------
__global__ void k_volume_of_mass(double *__restrict__ volume, const double *__restrict__ mass,
                                 i64 n, double rho_w)
{
    i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) volume[i] = mass[i] / rho_w;
}
__global__ void k_mass_of_volume(double *__restrict__ mass, const double *__restrict__ volume, i64 n,
                                 double rho_w)
{
    i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) mass[i] = rho_w * volume[i];
}

PYSDM_EXPORT int pysdm_b200_volume_of_water_mass(double *volume, const double *mass, i64 n,
                                                 double rho_w, void *stream)
{
    if (n <= 0) return 0;
    k_volume_of_mass<<<grid_for(n), kThreads, 0, (cudaStream_t)stream>>>(volume, mass, n, rho_w);
    return LAUNCH_CHECK();
}

PYSDM_EXPORT int pysdm_b200_mass_of_water_volume(double *mass, const double *volume, i64 n,
                                                 double rho_w, void *stream)
{
    if (n <= 0) return 0;
    k_mass_of_volume<<<grid_for(n), kThreads, 0, (cudaStream_t)stream>>>(mass, volume, n, rho_w);
    return LAUNCH_CHECK();
}
